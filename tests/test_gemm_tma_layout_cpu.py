"""CPU: the index arithmetic of the TMA-fed FP64 GEMM (pyscf_isdf_b200/csrc/gemm_tma.cu), restated in Python.

A TMA box lands in shared memory dense and SWIZZLE_128B-swizzled (the 16-byte chunk index of every 128-byte row is
XORed with the row index mod 8), so the kernel cannot pad rows; it makes the DMMA fragment loads bank-conflict free
by permuting which tile row holds which x index.  Checked here, without a GPU:
  * the fragment addresses of load_frag<> read exactly the (x, k) elements the permutation says, from where the
    TMA unit put them;
  * every LDS.128 quarter-warp / LDS.64 half-warp touches each 8-byte bank word of a 128-byte line once;
  * the four tiles of a 32-wide warp tile cover its x indices once, and the epilogue's row_of / col_of are the
    C coordinates of what the DMMAs accumulated.
(What the GPU tests add: the same bits as the oracle's FMA chain, tests/test_g1_stages_gpu.py.)"""
import itertools
import numpy as np


def perm(kmajor, g):
    return (g >> 1) + 4 * (g & 1) if kmajor else g


def smem_word(kmajor, x, k):
    """8-byte word offset of element (x, k) of an operand tile as the TMA boxes leave it (tile base 1024-aligned)."""
    if kmajor:                      # boxes of {16 x, 16 k}: [x >> 4][k][16 x], row = k
        row, chunk, half, base = k, (x & 15) >> 1, x & 1, (x >> 4) * 256 + k * 16
    else:                           # one box of {16 k, BX x}: [x][16 k], row = x
        row, chunk, half, base = x, k >> 1, k & 1, x * 16
    return base + ((chunk ^ (row & 7)) << 1) + half


def frag_words(kmajor, wx, kk, g, t):
    """(word offsets, (x, k) read) of load_frag<KMAJOR> for lane (g, t): v[2 grp + ii]."""
    out = []
    k = kk + t
    for grp in range(2):
        if kmajor:
            sw = (perm(True, g) ^ (k & 7)) << 1
            w0 = ((wx >> 4) + grp) * 256 + k * 16 + sw           # one LDS.128: v[2 grp], v[2 grp + 1]
            out += [(w0, grp, 0), (w0 + 1, grp, 1)]
        else:
            for ii in range(2):
                x = wx + 16 * grp + 2 * g + ii
                out.append((x * 16 + ((((k >> 1) ^ (x & 7)) << 1) | (k & 1)), grp, ii))
    return out


def test_fragment_addresses_read_the_permuted_elements():
    for kmajor, wx, kk in itertools.product((True, False), (0, 32), (0, 4, 8, 12)):
        for g, t in itertools.product(range(8), range(4)):
            for word, grp, ii in frag_words(kmajor, wx, kk, g, t):
                x = wx + 16 * grp + 2 * perm(kmajor, g) + ii        # tile i = 2 grp + ii, row g
                assert word == smem_word(kmajor, x, kk + t), (kmajor, wx, kk, g, t, grp, ii)


def test_tiles_cover_the_warp_tile_once():
    for kmajor in (True, False):
        xs = sorted(16 * grp + 2 * perm(kmajor, g) + ii for grp in range(2) for ii in range(2) for g in range(8))
        assert xs == list(range(32))


def test_fragment_loads_are_bank_conflict_free():
    for wx, kk in itertools.product((0, 32), (0, 4, 8, 12)):
        # k-major: LDS.128, served a quarter-warp (8 lanes x 16 bytes = one 128-byte wavefront) at a time
        for grp, q in itertools.product(range(2), range(4)):
            lanes = [(g, t) for g in (2 * q, 2 * q + 1) for t in range(4)]
            chunks = sorted((frag_words(True, wx, kk, g, t)[2 * grp][0] % 16) >> 1 for g, t in lanes)
            assert chunks == list(range(8)), (wx, kk, grp, q, chunks)
        # k-contiguous: LDS.64, a half-warp (16 lanes x 8 bytes) per wavefront
        for idx, hw in itertools.product(range(4), range(2)):
            lanes = [(g, t) for g in range(4 * hw, 4 * hw + 4) for t in range(4)]
            words = sorted(frag_words(False, wx, kk, g, t)[idx][0] % 16 for g, t in lanes)
            assert words == list(range(16)), (wx, kk, idx, hw, words)


def test_epilogue_coordinates_are_what_the_dmmas_accumulate():
    """mma.m8n8k4: lane (g, t) holds A[g][t], B[t][g] and C[g][2t], C[g][2t+1].  With tile i's row g = x_A and tile
    j's column q = x_B, acc[i][j][e] of lane (g, t) is C[x_A(i, g)][x_B(j, 2t + e)]: row_of / col_of of the kernel."""
    rng = np.random.default_rng(0)
    for ak, bk in itertools.product((True, False), repeat=2):
        A = rng.standard_normal((32, 4))          # [x, k] of one k-step, natural order
        B = rng.standard_normal((32, 4))
        C = np.zeros((32, 32))
        for i, j, g, t, e in itertools.product(range(4), range(4), range(8), range(4), range(2)):
            # what the DMMA of tile (i, j) puts into lane (g, t)'s c[e]: sum_k A_tile[g][k] B_tile[k][2t+e]
            xa = 16 * (i >> 1) + 2 * perm(ak, g) + (i & 1)
            xb = 16 * (j >> 1) + 2 * perm(bk, 2 * t + e) + (j & 1)
            val = A[xa].dot(B[xb])
            row = 16 * (i >> 1) + 2 * perm(ak, g) + (i & 1)                     # row_of(i)
            col = 16 * (j >> 1) + 2 * perm(bk, 2 * t + e) + (j & 1)             # col_of(j >> 1, e) + (j & 1)
            C[row, col] += val
        assert np.allclose(C, A.dot(B.T), rtol=0, atol=1e-13)
        # the two accumulators a lane stores together (tiles 2jp and 2jp+1, same e) are adjacent columns: double2
        for t, e in itertools.product(range(4), range(2)):
            c0 = 2 * perm(bk, 2 * t + e)
            assert c0 % 2 == 0 and c0 + 1 < 16
