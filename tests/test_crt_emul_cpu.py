"""CPU: the residue arithmetic of the CRT form of the int8 metric product
(pyscf_isdf_b200/csrc/crt_core.cuh, compiled with g++ through tests/emul/crt_emul.cpp) against exact
Python integers, and the whole scheme (scale -> residues -> per-modulus integer products -> Garner ->
Horner) against a float128/integer reference on small matrices.  Integer steps must be exact; the
rebuilt dot products must agree with the exact integers to 2^-50 relative."""
import ctypes
import os
import subprocess
from math import gcd

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'emul', 'crt_emul.cpp')
OUT = os.path.join(HERE, 'emul', '_build', 'libcrt_emul.so')
DEPS = [SRC, os.path.join(HERE, '..', 'pyscf_isdf_b200', 'csrc', 'crt_core.cuh')]


@pytest.fixture(scope='module')
def crt():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in DEPS):
        subprocess.check_call(['g++', '-std=c++17', '-O1', '-fPIC', '-shared', '-o', OUT, SRC])
    lib = ctypes.CDLL(OUT)
    lib.crt_reconstruct.restype = ctypes.c_double
    lib.crt_reconstruct.argtypes = [ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    lib.crt_residues.argtypes = [ctypes.c_longlong, ctypes.c_char_p, ctypes.c_int]
    return lib


def _moduli(lib):
    return [lib.crt_modulus(k) for k in range(16)]


def test_moduli_and_input_bits(crt):
    p = _moduli(crt)
    assert all(2 <= q <= 256 for q in p)
    assert all(gcd(a, b) == 1 for i, a in enumerate(p) for b in p[i + 1:])
    for nm in range(6, 17):
        P = 1
        for q in p[:nm]:
            P *= q
        b = crt.crt_input_bits(nm)
        assert (1 << 15) * (1 << (2 * b)) <= P // 2 - 1                      # the dot products fit
        assert b == 53 or (1 << 15) * (1 << (2 * (b + 1))) > P // 2 - 1      # and b is the largest such
    assert crt.crt_input_bits(16) == 53


def test_residues_exact(crt):
    p = _moduli(crt)
    rng = np.random.default_rng(1)
    xs = [0, 1, -1, 2**53, -2**53, 127, 128, -128, -129] + [int(v) for v in rng.integers(-2**53, 2**53, 2000)]
    buf = ctypes.create_string_buffer(16)
    for x in xs:
        crt.crt_residues(x, buf, 16)
        r = np.frombuffer(buf.raw, dtype=np.int8)
        for k, q in enumerate(p):
            assert (int(r[k]) - x) % q == 0
            assert -(q // 2) <= int(r[k]) <= (q - 1) // 2
    # run-time modulus reduction of int32 accumulators (product-kernel epilogue)
    for v in [0, 1, -1, 2**31 - 1, -2**31 + 1, 536870912, -536870912] + [int(t) for t in rng.integers(-2**31 + 1, 2**31 - 1, 2000)]:
        for q in p:
            r = crt.crt_centred_mod_rt(v, q)
            assert (r - v) % q == 0 and -(q // 2) <= r <= (q - 1) // 2


@pytest.mark.parametrize('nm', [8, 12, 14, 16])
def test_reconstruct_against_big_integers(crt, nm):
    p = _moduli(crt)[:nm]
    P = 1
    for q in p:
        P *= q
    rng = np.random.default_rng(nm)
    vals = [0, 1, -1, P // 2 - 1, -(P // 2 - 1), 12345678901234567890 % (P // 2)]
    for bits in (3, 20, 52, 53, 60, 90, 120):
        if (1 << bits) < P // 2:
            vals += [int(rng.integers(0, 2**62)) % (1 << bits) * (1 if rng.random() < .5 else -1) for _ in range(50)]
            vals += [((1 << bits) - 1), -((1 << bits) - 1)]
    arr = (ctypes.c_int * nm)()
    for x in vals:
        for k, q in enumerate(p):
            arr[k] = x % q                       # any representative is allowed
        got = crt.crt_reconstruct(arr, nm)
        if abs(x) < 2**53:
            assert got == float(x), (x, got)
        else:
            assert abs(got - float(x)) <= 2.0**-50 * abs(float(x)), (x, got)


def test_whole_scheme_matches_exact_product(crt):
    """Rows scaled by powers of two, B-bit integers, residue matrices, exact integer products per modulus
    (what the int8 tensor cores compute), reduction, reconstruction -- against the exact integer product
    and against the FP64 product of the original matrix."""
    p = _moduli(crt)
    nm = 16
    B = crt.crt_input_bits(nm)
    rng = np.random.default_rng(7)
    n, K = 9, 300
    A = rng.standard_normal((n, K)) * np.exp(3 * rng.standard_normal((n, 1)))      # rows of very different size
    e = np.ceil(np.log2(abs(A).max(axis=1)))
    sc = 2.0 ** e
    X = np.rint(A / sc[:, None] * 2.0**B).astype(np.int64)                        # |X| <= 2^B
    Xo = X.astype(object)
    exact = Xo.dot(Xo.T)
    buf = ctypes.create_string_buffer(16)
    R = np.zeros((nm, n, K), dtype=np.int64)
    for i in range(n):
        for k in range(K):
            crt.crt_residues(int(X[i, k]), buf, nm)
            R[:, i, k] = np.frombuffer(buf.raw, dtype=np.int8)[:nm]
    arr = (ctypes.c_int * nm)()
    C = np.zeros((n, n))
    for i in range(n):
        for j in range(n):
            for m in range(nm):
                acc = int(R[m, i].dot(R[m, j]))                  # int32-exact on the tensor cores
                assert abs(acc) < 2**31
                arr[m] = crt.crt_centred_mod_rt(acc, p[m])
            C[i, j] = crt.crt_reconstruct(arr, nm) * sc[i] * sc[j] * 2.0 ** (-2 * B)
            ex = float(exact[i, j])
            assert abs(crt.crt_reconstruct(arr, nm) - ex) <= 2.0**-50 * max(abs(ex), 1.0)
    ref = (A.astype(np.longdouble)).dot(A.astype(np.longdouble).T).astype(np.float64)
    norms = np.sqrt((A * A).sum(axis=1))
    assert (abs(C - ref) / np.outer(norms, norms)).max() < 1e-15


def _read_plane(planes, m, plane_stride, rb_stride, nrows_pad, kcpad):
    """Undo the pre-tiled SWIZZLE_128B layout documented in crt_core.cuh / syrk_i8.cu: block (i/128, k/128)
    is 16 KiB contiguous, row r at r*128 bytes, 16-byte chunk c at position c ^ (r & 7)."""
    out = np.zeros((nrows_pad, kcpad), dtype=np.int64)
    base = m * plane_stride
    for i in range(nrows_pad):
        r = i & 127
        for kg in range(0, kcpad, 16):
            c = (kg & 127) >> 4
            off = base + (i >> 7) * rb_stride + (kg >> 7) * 16384 + r * 128 + ((c ^ (r & 7)) << 4)
            out[i, kg:kg + 16] = planes[off:off + 16]
    return out


@pytest.mark.parametrize('nm', [16, 12])
def test_kernel_item_routines_end_to_end(crt, nm):
    """i8_split_crt_kernel -> (numpy in place of the tensor cores: exact integer products of the residue
    planes read back through the documented layout, reduced like the epilogue) -> crt_reconstruct_kernel,
    on a ragged case (n, kc not multiples of the block sizes, k0 > 0, padded leading dimensions)."""
    lib = crt
    p = _moduli(lib)
    rng = np.random.default_rng(3 + nm)
    n, K, k0, kc = 150, 700, 64, 333
    lda = K + 5
    A = np.zeros((n, lda))
    A[:, :K] = rng.standard_normal((n, K)) * np.exp(2 * rng.standard_normal((n, 1)))
    A[7] = 0.0                                                     # an all-zero row
    chunk = A[:, k0:k0 + kc]
    mx = abs(chunk).max(axis=1)
    sc = np.where(mx > 0, 2.0 ** (np.floor(np.log2(np.where(mx > 0, mx, 1.0))) + 1), 1.0)   # max = m 2^e, m in [0.5, 1)
    bits = lib.crt_input_bits(nm)
    nrows_pad, kcpad = 512, 384                                    # multiples of 512 rows / 128 bytes
    rb_stride = (kcpad // 128 | 1) * 16384
    plane_stride = rb_stride * (nrows_pad // 128)
    planes = np.full(nm * plane_stride, 77, dtype=np.int8)
    lib.crt_split_all.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_longlong,
                                  ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong,
                                  ctypes.c_longlong, ctypes.c_int, ctypes.c_int]
    lib.crt_split_all(A.ctypes.data, lda, n, nrows_pad, k0, kc, kcpad, sc.ctypes.data, planes.ctypes.data,
                      rb_stride, plane_stride, nm, bits)
    ldr = (n + 15) // 16 * 16
    res_stride = n * ldr
    R = np.zeros(nm * res_stride, dtype=np.int8)
    X = np.rint(chunk / sc[:, None] * 2.0 ** bits).astype(np.int64)
    for m in range(nm):
        Pm = _read_plane(planes, m, plane_stride, rb_stride, nrows_pad, kcpad)
        assert not Pm[n:].any() and not Pm[:, kc:].any()           # padding rows / columns are zero
        assert ((Pm[:n, :kc] - X) % p[m] == 0).all() and abs(Pm).max() <= 128
        acc = Pm[:n].dot(Pm[:n].T)                                   # what the int8 MMAs accumulate (exact in int32)
        assert abs(acc).max() < 2**31
        red = np.array([[lib.crt_centred_mod_rt(int(v), p[m]) for v in row] for row in acc], dtype=np.int8)
        R[m * res_stride:(m + 1) * res_stride].reshape(n, ldr)[:, :n] = red
    ldc = n + 3
    C = np.zeros((n, ldc))
    C[:, n:] = 7.0                                                 # padding columns must stay untouched
    alpha = 0.37
    lib.crt_reconstruct_all.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p, ctypes.c_longlong]
    lib.crt_reconstruct_all(R.ctypes.data, ldr, res_stride, nm, n, sc.ctypes.data,
                            alpha * 2.0 ** (-2 * bits), C.ctypes.data, ldc)
    ref = alpha * chunk.astype(np.longdouble).dot(chunk.astype(np.longdouble).T).astype(np.float64)
    got = C[:, :n]
    iu = np.triu_indices(n)
    norms = np.sqrt((chunk * chunk).sum(axis=1))
    scale = np.outer(norms, norms)[iu] + 1e-300
    tol = 1e-15 if nm == 16 else 2.0 ** (-bits + 2)
    assert (abs(got[iu] - ref[iu]) / scale).max() < tol
    assert (C[:, n:] == 7.0).all() and (np.tril(got, -1) == 0).all()      # only the upper triangle is touched
