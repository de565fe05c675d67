"""GPU parity tests: every stage of the CUDA path against the CPU oracle,
called through the C-ABI (pyscf_isdf_b200._lib -> libisdf_b200.so)."""
import numpy as np
import pytest

import oracle
from oracle import c_oracle
from pyscf_isdf_b200._lib import pad8
from util import diamond_c8, he_c_cell, oracle_phi, to_dev, rand_dm

pytestmark = pytest.mark.gpu


def _gemm_case(handle, M, N, K, ak, bk, neg=0, epi=0, cin=False, seed=0, pad=0):
    import torch
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((K, M) if ak else (M, K))
    B = rng.standard_normal((K, N) if bk else (N, K))
    Cin = rng.standard_normal((M, N)) if cin else None
    E = rng.standard_normal((M, N)) if epi == 2 else None
    dA = to_dev(handle, A, A.shape[1] + pad)
    dB = to_dev(handle, B, B.shape[1] + pad)
    dC = torch.zeros((M, N + pad), dtype=torch.float64, device=handle.torch_device)
    dCin = to_dev(handle, Cin, N + pad) if cin else None
    dE = to_dev(handle, E, N + pad) if E is not None else None
    handle.dgemm(M, N, K, dA, dA.shape[1], ak, dB, dB.shape[1], bk, dC, N + pad,
                 Cin=dCin, ldcin=N + pad, E=dE, lde=N + pad, alpha=0.75, neg=neg, epi=epi)
    torch.cuda.synchronize()
    got = dC.cpu().numpy()[:, :N]
    ref = c_oracle.gemm_chain(A, ak, B, bk, Cin=Cin, E=E, alpha=0.75, neg=neg, epi=epi)
    return got, ref


@pytest.mark.parametrize('ak,bk', [(1, 1), (0, 1), (1, 0), (0, 0)])
def test_gemm_layouts_bit_exact(handle, ak, bk):
    """All four operand layouts; result is the oracle's FMA chain bit for bit."""
    for (M, N, K) in [(128, 128, 64), (200, 333, 77), (64, 64, 16), (1, 5, 3), (300, 40, 1000)]:
        got, ref = _gemm_case(handle, M, N, K, ak, bk)
        assert np.array_equal(got, ref), (M, N, K, abs(got - ref).max())


def test_gemm_big_tile_and_epilogues(handle):
    # >= 148 tiles of 128x128 selects the 8x8 micro-tile kernel
    for epi, neg, cin in [(0, 0, False), (1, 0, False), (2, 0, False), (0, 1, True)]:
        got, ref = _gemm_case(handle, 1700, 1500, 130, 1, 1, neg=neg, epi=epi, cin=cin, seed=epi + neg)
        assert np.array_equal(got, ref), (epi, neg, cin, abs(got - ref).max())
    got, ref = _gemm_case(handle, 1601, 1409, 75, 0, 0, seed=5)          # ragged, k-contiguous
    assert np.array_equal(got, ref)
    got, ref = _gemm_case(handle, 1601, 1409, 75, 1, 1, seed=6, pad=1)   # odd leading dimension -> scalar path
    assert np.array_equal(got, ref)


@pytest.mark.parametrize('form', [7, 8, 9, 10, 11])
def test_gemm_tma_forms_bit_exact(handle, form):
    """The warp-specialised GEMM fed by 2-D tensor-map TMA (csrc/gemm_tma.cu; CTA tiles 64x128, 128x64,
    128x128, 64x64, 32x128): the oracle's FMA chain bit for bit, every layout, ragged M / N / K (the TMA unit's
    out-of-bounds zero fill), accumulator init, epilogues; leading dimensions even so that TMA can address them."""
    handle.set_gemm_tile(form)
    try:
        # even sizes + pad 2: every leading dimension is even (TMA-addressable) and differs from the width;
        # none of the sizes is a multiple of the CTA tiles or of the 16-deep k tile
        for (M, N, K) in [(128, 300, 64), (70, 132, 38), (258, 66, 16), (34, 1000, 130), (6, 4, 2)]:
            for ak, bk in [(1, 1), (0, 1), (1, 0), (0, 0)]:
                got, ref = _gemm_case(handle, M, N, K, ak, bk, neg=ak, cin=bool(bk), seed=M + form, pad=2)
                assert np.array_equal(got, ref), (form, M, N, K, ak, bk, abs(got - ref).max())
        # odd K: k-major operands have K rows, any K works
        got, ref = _gemm_case(handle, 130, 258, 77, 1, 1, neg=1, cin=True, seed=9)
        assert np.array_equal(got, ref)
        for epi in (1, 2):
            got, ref = _gemm_case(handle, 200, 150, 90, 1, 1, epi=epi, seed=epi)
            assert np.array_equal(got, ref), (form, epi)
    finally:
        handle.set_gemm_tile(0)


def test_gemm_tma_many_ctas_per_sm(handle):
    """Flush-shaped products (C and B are rows of one matrix, in place) wide enough for three CTAs per SM and
    several waves: the default kernel (TMA-fed where it applies) bit-equal to the LDGSTS kernel, ten times each, and to
    the oracle's FMA chain once.  This is the case a missing generic->async proxy fence corrupts (gemm_tma.cu)."""
    import torch
    dev = handle.torch_device
    g = torch.Generator(device=dev); g.manual_seed(3)
    N, ld = 42875, 42880
    for M, K, ak in [(64, 192, 0), (52, 194, 0), (128, 400, 1)]:
        A = torch.rand((K, M + 2) if ak else (M, K + 2), dtype=torch.float64, device=dev, generator=g) - 0.5
        Lt = torch.rand((K + M, ld), dtype=torch.float64, device=dev, generator=g) - 0.5
        C0 = Lt[K:].clone()
        ref = None
        try:
            for rep, form in enumerate([4] + [0] * 10):
                handle.set_gemm_tile(form)
                Lt[K:].copy_(C0)
                torch.cuda.synchronize()
                handle.dgemm(M, N, K, A, A.shape[1], ak, Lt, ld, 1, Lt[K:], ld, Cin=Lt[K:], ldcin=ld, neg=1)
                torch.cuda.synchronize()
                if form == 4:
                    ref = Lt[K:, :N].clone()
                else:
                    assert torch.equal(Lt[K:, :N], ref), (M, K, rep, int((Lt[K:, :N] != ref).sum()))
        finally:
            handle.set_gemm_tile(0)
        if M == 52:
            a = A[:, :M].cpu().numpy() if ak else A[:, :K].cpu().numpy()
            chain = c_oracle.gemm_chain(a, ak, Lt[:K, :N].cpu().numpy(), 1, Cin=C0[:, :N].cpu().numpy(), alpha=1.0, neg=1)
            assert np.array_equal(ref.cpu().numpy(), chain)


def test_gemm_empty(handle):
    import torch
    z = torch.zeros((4, 4), dtype=torch.float64, device=handle.torch_device)
    handle.dgemm(0, 4, 4, z, 4, 1, z, 4, 1, z, 4)       # M = 0: no-op
    handle.dgemm(4, 4, 0, z, 4, 1, z, 4, 1, z, 4)       # K = 0: C = 0
    torch.cuda.synchronize()
    assert float(z.abs().max()) == 0.0


def _gpu_phi(handle, cell, g0=0, g1=None):
    import torch
    ng = int(np.prod(cell.mesh))
    g1 = ng if g1 is None else g1
    ld = pad8(g1 - g0)
    phi = torch.full((cell.nao, ld), float('nan'), dtype=torch.float64, device=handle.torch_device)
    rcut, Ls = cell.ao_eval_params()
    handle.eval_ao(cell._atm, cell._bas, cell._env, cell.lattice_vectors(), cell.mesh, Ls, rcut,
                   g0, g1, phi, ld)
    torch.cuda.synchronize()
    return phi, ld


def test_eval_ao_matches_oracle(handle):
    """Stage 1, s/p/d shells, lattice sum; tolerance 1e-12 absolute."""
    cell = diamond_c8([15, 15, 15], jitter=0.2)
    ref = oracle_phi(cell)
    phi, ld = _gpu_phi(handle, cell)
    got = phi.cpu().numpy()[:, :ref.shape[1]]
    assert np.isfinite(got).all()
    assert abs(got - ref).max() < 1e-12
    assert abs(oracle.fp(got.T) - oracle.fp(ref.T)) < 1e-10


def test_eval_ao_golden_cell_and_slab(handle):
    """He + C/gth-szv fixture of pyscf/pbc/df/test/test_fft.py:545-554; also a
    slab [g0, g1) that is not aligned to the 56-point screening blocks."""
    cell = he_c_cell()
    ref = oracle_phi(cell)
    phi, _ = _gpu_phi(handle, cell)
    got = phi.cpu().numpy()[:, :ref.shape[1]]
    assert abs(got - ref).max() < 1e-12
    # exact-grid J/K from the GPU AO values reproduces the reference's golden numbers
    vj, vk = oracle.fftdf_get_jk(got.T, np.eye(cell.nao, dtype=int), cell.lattice_vectors(), cell.mesh)
    assert abs(oracle.fp(vj) - 3.7955873127283377) < 1e-8
    assert abs(oracle.fp(vk) - 4.290076429522121) < 1e-8
    g0, g1 = 1001, 5003
    slab, _ = _gpu_phi(handle, cell, g0, g1)
    assert abs(slab.cpu().numpy()[:, :g1 - g0] - ref[:, g0:g1]).max() < 1e-12


def _select(handle, phi_np, npts, tol=-1.0):
    import torch
    nao, ng = phi_np.shape
    ld = pad8(ng)
    phi = to_dev(handle, phi_np, ld)
    Lt = torch.zeros((npts, ld), dtype=torch.float64, device=handle.torch_device)
    ip, resid = handle.select_ip(phi, nao, ng, ld, npts, tol, Lt, ld)
    torch.cuda.synchronize()
    return ip, resid, Lt.cpu().numpy()[:len(ip), :ng]


def test_select_ip_bit_exact_random(handle):
    """Stage 2 on a generic matrix: pivots, residuals and factor bit-identical
    to the unblocked FMA-chain oracle."""
    rng = np.random.default_rng(3)
    phi = rng.standard_normal((24, 5000))
    ip_ref, resid_ref, Lt_ref = c_oracle.select_ip_chain(phi, 300, return_factor=True)
    ip, resid, Lt = _select(handle, phi, 300)
    assert np.array_equal(ip, ip_ref)
    assert np.array_equal(resid, resid_ref)
    assert np.array_equal(Lt, Lt_ref)
    # numpy (BLAS-order) oracle picks the same points on generic data
    ip_np, _ = oracle.select_ip(phi, 300)
    assert np.array_equal(ip, ip_np)


def test_select_ip_symmetric_cell_and_rank_stop(handle):
    """Perfect diamond (exact ties by symmetry): tie-break = lowest grid index,
    still bit-exact.  Also the early stop when the residual falls below tol."""
    cell = diamond_c8([12, 12, 12], basis='gth-szv')
    phi = oracle_phi(cell)
    ip_ref, resid_ref = c_oracle.select_ip_chain(phi, 200)
    ip, resid, _ = _select(handle, phi, 200)
    assert np.array_equal(ip, ip_ref)
    assert np.array_equal(resid, resid_ref)
    # rank-deficient: 3 AOs -> pair space of dimension 6
    rng = np.random.default_rng(5)
    small = rng.standard_normal((3, 700))
    ip_ref, resid_ref = c_oracle.select_ip_chain(small, 50)
    ip, resid, _ = _select(handle, small, 50)
    assert len(ip) == len(ip_ref) <= 7
    assert np.array_equal(ip, ip_ref)


def test_fit_aux_matches_oracle(handle):
    """Stage 3: zeta = A^-1 B (Cholesky path) and the factor-reuse path agree
    with the oracle on the fitted pair densities."""
    import torch
    cell = diamond_c8([12, 12, 12], basis='gth-szv', jitter=0.2)
    phi_np = oracle_phi(cell)
    nao, ng = phi_np.shape
    npts = 5 * nao
    ip_ref, _ = c_oracle.select_ip_chain(phi_np, npts)
    zeta_ref = oracle.fit_aux(phi_np, ip_ref)
    ld = pad8(ng)
    phi = to_dev(handle, phi_np, ld)
    zeta = torch.zeros((npts, ld), dtype=torch.float64, device=handle.torch_device)
    handle.fit_aux(phi, nao, ng, ld, ip_ref, zeta, ld)
    torch.cuda.synchronize()
    got = zeta.cpu().numpy()[:, :ng]
    X = phi_np[:, ip_ref]
    pair = lambda z: np.einsum('mp,np,pg->mng', X, X, z)       # fitted pair densities
    scale = abs(pair(zeta_ref)).max()
    assert abs(pair(got) - pair(zeta_ref)).max() < 1e-9 * scale
    assert abs(got - zeta_ref).max() < 1e-6 * abs(zeta_ref).max()
    # interpolation property: zeta_P(r_Q) = delta_PQ
    assert abs(got[:, ip_ref] - np.eye(npts)).max() < 1e-7
    # factor-reuse path
    Lt = torch.zeros((npts, ld), dtype=torch.float64, device=handle.torch_device)
    ip, _ = handle.select_ip(phi, nao, ng, ld, npts, -1.0, Lt, ld)
    assert np.array_equal(ip, ip_ref)
    handle.fit_aux_from_factor(Lt, ld, ng, ip)
    torch.cuda.synchronize()
    got2 = Lt.cpu().numpy()[:, :ng]
    assert abs(pair(got2) - pair(zeta_ref)).max() < 1e-9 * scale


@pytest.mark.parametrize('mesh', [[12, 10, 9], [7, 11, 13], [16, 8, 5], [35, 6, 4],
                                  [71, 5, 4], [6, 23, 47], [4, 22, 71]])
def test_coulomb_apply_matches_numpy_fft(handle, mesh):
    """Stage 4 FFT + fused 4 pi/G^2 on even/odd/prime meshes (primes > 7 go through
    Rader's algorithm, 22 = 2*11 through the direct large-radix stage), triclinic lattice."""
    import torch
    rng = np.random.default_rng(11)
    a = np.array([[5.0, 0.3, 0.1], [0.2, 4.5, -0.4], [0.0, 0.6, 6.0]])
    ng = int(np.prod(mesh))
    nvec = 5
    z = rng.random((nvec, ng)) - .5
    ref = oracle.ifft(oracle.fft(z, mesh) * oracle.get_coulG(a, mesh), mesh).real
    ld = pad8(ng)
    dz = to_dev(handle, z, ld)
    out = torch.zeros((nvec, ld), dtype=torch.float64, device=handle.torch_device)
    handle.coulomb_apply(dz, ld, nvec, mesh, a, out, ld)
    torch.cuda.synchronize()
    got = out.cpu().numpy()[:, :ng]
    assert abs(got - ref).max() < 1e-10 * abs(ref).max()


def test_build_W_matches_oracle(handle):
    import torch
    rng = np.random.default_rng(12)
    mesh = [12, 9, 10]
    a = np.array([[5.0, 0.3, 0.1], [0.2, 4.5, -0.4], [0.0, 0.6, 6.0]])
    ng = int(np.prod(mesh))
    npts = 37
    z = rng.random((npts, ng)) - .5
    Wref = oracle.build_W(z, a, mesh)
    Vref = oracle.build_V(z, a, mesh)
    ld = pad8(ng)
    dz = to_dev(handle, z, ld)
    W = torch.zeros((npts, npts), dtype=torch.float64, device=handle.torch_device)
    V = torch.zeros((npts, ld), dtype=torch.float64, device=handle.torch_device)
    handle.build_W(dz, npts, ld, mesh, a, W, V=V, ldv=ld)
    torch.cuda.synchronize()
    assert abs(W.cpu().numpy() - Wref).max() < 1e-10 * abs(Wref).max()
    assert abs(V.cpu().numpy()[:, :ng] - Vref).max() < 1e-10 * abs(Vref).max()


def test_get_jk_matches_oracle(handle):
    import torch
    rng = np.random.default_rng(13)
    nao, npts = 26, 130
    X = rng.standard_normal((nao, npts))
    W = rng.standard_normal((npts, npts)); W = W + W.T
    dms = np.stack([rand_dm(nao, 1), rng.standard_normal((nao, nao))])     # symmetric and general
    vj_ref, vk_ref = oracle.isdf_get_jk(X, W, dms)
    dX, dW = to_dev(handle, X), to_dev(handle, W)
    ddm = torch.from_numpy(dms).to(handle.torch_device)
    vj = torch.zeros_like(ddm); vk = torch.zeros_like(ddm)
    handle.get_jk(dX, dW, ddm, 2, nao, npts, True, True, vj, vk)
    torch.cuda.synchronize()
    assert abs(vj.cpu().numpy() - vj_ref).max() < 1e-10 * abs(vj_ref).max()
    assert abs(vk.cpu().numpy() - vk_ref).max() < 1e-10 * abs(vk_ref).max()
    vk2 = torch.zeros_like(ddm)
    handle.get_jk(dX, dW, ddm, 2, nao, npts, False, True, None, vk2)
    torch.cuda.synchronize()
    assert np.array_equal(vk2.cpu().numpy(), vk.cpu().numpy())


def test_three_threads_launch_x_passes_of_different_widths(handle):
    """Several handles driven by threads of one process (SURVEY 8b's threading row): kernel attributes are
    per (function, device), so launches of different shared-memory sizes must not disturb each other.
    Three threads run the fused x pass on column slabs of different widths 200 times each (round 1's
    driver run failed exactly here: fft.cu set the attribute to each launch's own size)."""
    import threading
    import torch
    from pyscf_isdf_b200._lib import Handle
    mesh = [15, 14, 12]
    a = np.array([[5.0, 0.3, 0.1], [0.2, 4.5, -0.4], [0.0, 0.6, 6.0]])
    nx, ncols = mesh[0], mesh[1] * (mesh[2] // 2 + 1)
    widths = [(0, 33), (33, 33), (66, 32)]            # the 33/33/32 split of three ranks on this mesh
    assert widths[-1][0] + widths[-1][1] == ncols
    nvec = 3
    rng = np.random.default_rng(21)
    spec = rng.standard_normal((nvec, nx, ncols)) + 1j * rng.standard_normal((nvec, nx, ncols))
    # the kernel's weight at a Nyquist index is the mean of coulG(G) and coulG(-G) (same quadratic form on real fields)
    G2f = oracle.get_coulG(a, mesh).reshape(mesh)
    G2m = G2f[(-np.arange(mesh[0])) % mesh[0]][:, (-np.arange(mesh[1])) % mesh[1]][:, :, (-np.arange(mesh[2])) % mesh[2]]
    G2 = (0.5 * (G2f + G2m))[:, :, :mesh[2] // 2 + 1].reshape(nx, ncols)
    wz = np.full(mesh[2] // 2 + 1, 2.0); wz[0] = 1.0
    if mesh[2] % 2 == 0:
        wz[-1] = 1.0
    wgt = np.sqrt(G2 * np.tile(wz, mesh[1])[None, :])
    ref = np.fft.fft(spec, axis=1) * wgt[None]
    errs, outs = [], [None] * 3
    start = threading.Barrier(3)

    def work(r):
        try:
            c0, ncl = widths[r]
            with torch.cuda.stream(torch.cuda.Stream()):
                h2 = Handle(0)                                        # bound to this thread's stream
                src = torch.from_numpy(np.ascontiguousarray(spec[:, :, c0:c0 + ncl])).to(h2.torch_device)
                start.wait()
                for it in range(200):
                    Y = src.clone()
                    h2.fft_x_coulomb(Y, nx * ncl, nvec, mesh, a, c0, ncl)
                torch.cuda.current_stream().synchronize()
                outs[r] = Y.cpu().numpy()
        except Exception as e:            # noqa: BLE001
            errs.append(e)
            start.abort()

    ths = [threading.Thread(target=work, args=(r,)) for r in range(3)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(timeout=300)
    if errs:
        raise errs[0]
    for r, (c0, ncl) in enumerate(widths):
        got, want = outs[r], ref[:, :, c0:c0 + ncl]
        nz = abs(want) > 0
        assert abs(got - want)[nz].max() < 1e-10 * abs(want).max(), r


def _coul_weight(a, mesh):
    """sqrt(w coulG) on the half spectrum as the x pass applies it: [nx, ny * (nz/2+1)]"""
    nx, ny, nz = mesh
    nzh = nz // 2 + 1
    G2f = oracle.get_coulG(a, mesh).reshape(mesh)
    G2m = G2f[(-np.arange(nx)) % nx][:, (-np.arange(ny)) % ny][:, :, (-np.arange(nz)) % nz]
    G2 = (0.5 * (G2f + G2m))[:, :, :nzh].reshape(nx, ny * nzh)
    wz = np.full(nzh, 2.0); wz[0] = 1.0
    if nz % 2 == 0:
        wz[-1] = 1.0
    return np.sqrt(G2 * np.tile(wz, ny)[None, :])


@pytest.mark.parametrize('mesh', [[12, 10, 9], [7, 11, 13], [16, 8, 5], [35, 6, 4], [71, 5, 4], [6, 23, 47],
                                  [4, 22, 71], [2, 3, 2], [128, 3, 4], [5, 128, 6], [3, 4, 128], [17, 18, 20],
                                  [9, 130, 4]])
def test_forward_passes_match_numpy_fft(handle, mesh):
    """The slab-decomposed forward transform (isdf_fft_zy on a range of x planes, isdf_fft_x_coulomb on column
    slabs) against numpy.fft: every axis length class -- even, odd, prime, 2, the largest the dense-DFT passes take
    (128) and one beyond it (130: Stockham / direct-radix path), ragged slabs."""
    import torch
    rng = np.random.default_rng(31)
    a = np.array([[5.0, 0.3, 0.1], [0.2, 4.5, -0.4], [0.0, 0.6, 6.0]])
    nx, ny, nz = mesh
    nzh = nz // 2 + 1
    ncols = ny * nzh
    nvec = 3
    z = rng.random((nvec, nx, ny, nz)) - .5
    ref_zy = np.fft.fft(np.fft.rfft(z, axis=3), axis=2).reshape(nvec, nx, ncols)
    dev = handle.torch_device
    # z,y passes on two ranges of x planes
    x_split = [0, max(1, nx // 3), nx]
    spec = np.zeros((nvec, nx, ncols), dtype=complex)
    for x0, x1 in zip(x_split[:-1], x_split[1:]):
        nxl = x1 - x0
        ldin = pad8(nxl * ny * nz)
        zin = torch.zeros((nvec, ldin), dtype=torch.float64, device=dev)
        zin[:, :nxl * ny * nz] = torch.from_numpy(z[:, x0:x1].reshape(nvec, -1)).to(dev)
        out = torch.zeros((nvec, nxl * ncols), dtype=torch.complex128, device=dev)
        handle.fft_zy(zin, ldin, nvec, nxl, mesh, out, nxl * ncols)
        torch.cuda.synchronize()
        spec[:, x0:x1] = out.cpu().numpy().reshape(nvec, nxl, ncols)
    assert abs(spec - ref_zy).max() < 1e-11 * max(1.0, abs(ref_zy).max())
    # x pass with the fused weight on three ragged column slabs
    ref = np.fft.fft(ref_zy, axis=1) * _coul_weight(a, mesh)[None]
    c_split = sorted({0, ncols // 3, (2 * ncols) // 3 + 1 if ncols > 3 else ncols, ncols})
    for c0, c1 in zip(c_split[:-1], c_split[1:]):
        ncl = c1 - c0
        Y = torch.from_numpy(np.ascontiguousarray(ref_zy[:, :, c0:c1])).to(dev)
        handle.fft_x_coulomb(Y, nx * ncl, nvec, mesh, a, c0, ncl)
        torch.cuda.synchronize()
        got = Y.cpu().numpy()
        assert abs(got - ref[:, :, c0:c1]).max() < 1e-11 * max(1.0, abs(ref).max()), (c0, c1)
