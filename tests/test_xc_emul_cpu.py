"""CPU: the device exchange-correlation functionals (pyscf_isdf_b200/csrc/xc_core.cuh, forward-mode dual
numbers, compiled with g++ through tests/emul/xc_emul.cpp) point by point against oracle/xc.py (the same
published formulas differentiated by the complex-step rule and pinned to the reference's BLYP numbers)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
from oracle import xc

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'emul', 'xc_emul.cpp')
OUT = os.path.join(HERE, 'emul', '_build', 'libxc_emul.so')
DEPS = [SRC, os.path.join(HERE, '..', 'pyscf_isdf_b200', 'csrc', 'xc_core.cuh')]
CODES = {'lda,': 1, 'blyp': 2, 'pbe': 3, 'lda,vwn': 4, 'pbe0': 5, 'b3lyp': 6}


@pytest.fixture(scope='module')
def lib():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in DEPS):
        subprocess.check_call(['g++', '-std=c++17', '-O1', '-ffp-contract=off', '-fPIC', '-shared', '-o', OUT, SRC])
    return ctypes.CDLL(OUT)


def _eval(lib, code, rho, sigma):
    n = len(rho)
    out = [np.zeros(n) for _ in range(3)]
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    rc = lib.xc_eval(ctypes.c_int(code), ctypes.c_longlong(n), p(rho), p(sigma), *[p(o) for o in out])
    return rc, out


@pytest.mark.parametrize('name', ['lda,', 'blyp', 'pbe', 'lda,vwn', 'pbe0', 'b3lyp'])
def test_dual_number_functionals_match_the_oracle(lib, name):
    rng = np.random.default_rng(11)
    n = 4000
    rho = 10.0**rng.uniform(-8, 2, n)
    sigma = rho**(8 / 3.) * 10.0**rng.uniform(-6, 3, n)
    rho[:5] = [0.0, 1e-16, 1e-15, 2e-15, 1.0]                 # around the density threshold
    sigma[:5] = [0.0, 0.0, 1.0, 0.0, 0.0]                     # zero gradient: the sigma floor
    rc, (exc, vr, vs) = _eval(lib, CODES[name], rho, sigma)
    assert rc == 0
    e0, vr0, vs0 = xc.eval_xc(name, rho, sigma)
    assert np.isfinite(exc).all() and np.isfinite(vr).all() and np.isfinite(vs).all()
    assert (abs(exc - e0) <= 1e-12 * abs(e0) + 1e-300).all()
    assert (abs(vr - vr0) <= 1e-11 * abs(vr0) + 1e-300).all()
    # vsigma of b88 + lyp nearly cancels in places: compare against the size of the terms, not the sum
    scale = abs(vs0) + abs(e0) * rho / np.maximum(sigma, 1e-40) * 1e-3
    assert (abs(vs - vs0) <= 1e-10 * scale + 1e-300).all()
    assert (exc[:3] == 0).all() and (vr[:3] == 0).all() and (vs[:3] == 0).all()


def test_unknown_code_is_an_error(lib):
    rc, _ = _eval(lib, 9, np.ones(2), np.ones(2))
    assert rc == 1


@pytest.mark.parametrize('name', ['pbe', 'blyp'])
def test_streaming_kernel_bodies_give_nr_rks(lib, name):
    """rho / functional / aow item routines of xc.cu on a real cell, NumPy products in place of the two
    GEMMs: (nelec, exc, vmat) equal the oracle's nr_rks; padding columns stay untouched."""
    import oracle
    import sys
    sys.path.insert(0, HERE)
    from util import diamond_c8, rand_dm
    cell = diamond_c8([9, 10, 8], basis='gth-dzvp', jitter=0.15)
    rcut, Ls = cell.ao_eval_params()
    ao4 = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, deriv=1)   # [4, ng, nao]
    ng, nao = ao4.shape[1:]
    ld = (ng + 7) // 8 * 8 + 8
    phi4 = np.zeros((4, nao, ld))
    phi4[:, :, :ng] = ao4.transpose(0, 2, 1)
    dm = rand_dm(nao) * 0.05
    weight = cell.vol / ng
    T = np.ascontiguousarray(dm.dot(phi4[0]))
    rho4 = np.full((4, ld), 7.0)
    wv4 = np.full((4, ld), 7.0)
    aow = np.full((nao, ld), 7.0)
    sums = np.zeros(2)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    lib.xc_stream_items(ctypes.c_int(CODES[name]), ctypes.c_int(nao), ctypes.c_longlong(ng), ctypes.c_longlong(ld),
                        p(phi4), p(T), ctypes.c_double(weight), p(rho4), p(wv4), p(aow), p(sums), ctypes.c_int(5))
    m = phi4[0, :, :ng].dot(aow[:, :ng].T)
    ne0, exc0, vmat0 = xc.nr_rks(ao4, weight, name, dm)
    assert abs(sums[0] - ne0) < 1e-12 * abs(ne0) and abs(sums[1] - exc0) < 1e-12 * abs(exc0)
    assert abs(m + m.T - vmat0).max() < 1e-12 * abs(vmat0).max()
    assert (rho4[:, ng:] == 7.0).all() and (wv4[:, ng:] == 7.0).all() and (aow[:, ng:] == 7.0).all()
