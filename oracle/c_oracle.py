"""ctypes access to the C oracle (oracle/c/isdf_oracle.c) -- test infrastructure only."""
import ctypes
import os
import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'c', 'libisdf_oracle.so')
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            raise RuntimeError('C oracle not built: run `make -C oracle/c` (or __graft_entry__.build())')
        _lib = ctypes.CDLL(_PATH)
        _lib.oracle_select_ip.restype = ctypes.c_long
        _lib.oracle_select_ip_exact.restype = ctypes.c_long
        _lib.oracle_select_ip_blocked.restype = ctypes.c_long
        _lib.oracle_recon_parts.restype = ctypes.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def select_ip_chain(aoR, npts, tol=-1.0, return_factor=False):
    """Bit-exact FMA-chain pivoted Cholesky; aoR [nao, ng] (C-contiguous)."""
    aoR = np.ascontiguousarray(aoR, dtype=np.float64)
    nao, ng = aoR.shape
    Lt = np.zeros((npts, ng))
    ip = np.zeros(npts, dtype=np.int64)
    resid = np.zeros(npts)
    rank = lib().oracle_select_ip(_p(aoR), ctypes.c_int(nao), ctypes.c_long(ng), ctypes.c_long(ng),
                                  ctypes.c_int(npts), ctypes.c_double(tol), _p(Lt), ctypes.c_long(ng),
                                  _p(ip), _p(resid))
    if return_factor:
        return ip[:rank].copy(), resid[:rank].copy(), Lt[:rank]
    return ip[:rank].copy(), resid[:rank].copy()


def select_ip_blocked(aoR, npts, tol=-1.0, return_factor=False):
    """The same selection as select_ip_chain, bit for bit (tests/test_oracle_blocked_cpu.py), evaluated in
    panels with OpenMP threads (oracle/c/isdf_oracle_blocked.c): the form that finishes at the BASELINE sizes."""
    aoR = np.ascontiguousarray(aoR, dtype=np.float64)
    nao, ng = aoR.shape
    Lt = np.zeros((npts, ng))
    ip = np.zeros(npts, dtype=np.int64)
    resid = np.zeros(npts)
    rank = lib().oracle_select_ip_blocked(_p(aoR), ctypes.c_int(nao), ctypes.c_long(ng), ctypes.c_long(ng),
                                          ctypes.c_int(npts), ctypes.c_double(tol), _p(Lt), ctypes.c_long(ng),
                                          _p(ip), _p(resid))
    if return_factor:
        return ip[:rank].copy(), resid[:rank].copy(), Lt[:rank]
    return ip[:rank].copy(), resid[:rank].copy()


def select_ip_exact(aoR, npts, tol=-1.0, return_factor=False):
    """PROTOTYPE: the selection with exact fixed-point long sums
    (oracle/c/isdf_oracle_exact.c); same interface as select_ip_chain."""
    aoR = np.ascontiguousarray(aoR, dtype=np.float64)
    nao, ng = aoR.shape
    Lt = np.zeros((npts, ng))
    ip = np.zeros(npts, dtype=np.int64)
    resid = np.zeros(npts)
    rank = lib().oracle_select_ip_exact(_p(aoR), ctypes.c_int(nao), ctypes.c_long(ng), ctypes.c_long(ng),
                                        ctypes.c_int(npts), ctypes.c_double(tol), _p(Lt), ctypes.c_long(ng),
                                        _p(ip), _p(resid))
    if return_factor:
        return ip[:rank].copy(), resid[:rank].copy(), Lt[:rank]
    return ip[:rank].copy(), resid[:rank].copy()


def recon_int(value):
    """R(S) of the exact-arithmetic prototype for a Python integer |S| < 2^124."""
    hi, lo = divmod(int(value), 1 << 64)
    return lib().oracle_recon_parts(ctypes.c_longlong(hi), ctypes.c_ulonglong(lo))


def gemm_chain(A, a_kmajor, B, b_kmajor, Cin=None, E=None, alpha=1.0, neg=0, epi=0):
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    K, M = A.shape if a_kmajor else A.shape[::-1]
    Kb, N = B.shape if b_kmajor else B.shape[::-1]
    assert K == Kb
    C = np.zeros((M, N))
    if Cin is not None:
        Cin = np.ascontiguousarray(Cin, dtype=np.float64)
    if E is not None:
        E = np.ascontiguousarray(E, dtype=np.float64)
    lib().oracle_gemm_chain(ctypes.c_int(M), ctypes.c_int(N), ctypes.c_int(K),
                            _p(A), ctypes.c_long(A.shape[1]), ctypes.c_int(a_kmajor),
                            _p(B), ctypes.c_long(B.shape[1]), ctypes.c_int(b_kmajor),
                            _p(C), ctypes.c_long(N),
                            _p(Cin) if Cin is not None else None, ctypes.c_long(N),
                            _p(E) if E is not None else None, ctypes.c_long(N),
                            ctypes.c_double(alpha), ctypes.c_int(neg), ctypes.c_int(epi))
    return C
