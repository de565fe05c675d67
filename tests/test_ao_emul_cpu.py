"""CPU: the angular and radial pieces of the AO gradient kernel (pyscf_isdf_b200/csrc/ao_core.cuh,
compiled with g++ through tests/emul/ao_emul.cpp) against the oracle's cartesian -> spherical route
(oracle/eval_ao.py:shell_values_deriv1, pinned to the reference's fingerprint of the AO gradients)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
from oracle.eval_ao import FAC_SP, shell_values_deriv1

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'emul', 'ao_emul.cpp')
OUT = os.path.join(HERE, 'emul', '_build', 'libao_emul.so')
DEPS = [SRC, os.path.join(HERE, '..', 'pyscf_isdf_b200', 'csrc', 'ao_core.cuh')]


@pytest.fixture(scope='module')
def lib():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in DEPS):
        subprocess.check_call(['g++', '-std=c++17', '-O1', '-fPIC', '-shared', '-o', OUT, SRC])
    return ctypes.CDLL(OUT)


@pytest.mark.parametrize('l', [0, 1, 2, 3])
def test_shell_values_and_gradients(lib, l):
    rng = np.random.default_rng(l)
    n = 500
    dr = rng.normal(size=(n, 3)) * 1.5
    alpha = np.array([2.1, 0.7, 0.15])
    coeff = np.array([[0.3, -0.8, 1.1]])
    ref = shell_values_deriv1(l, dr, np.einsum('gx,gx->g', dr, dr), alpha, coeff)       # [4, 2l+1, n]
    out = np.zeros((4, 2 * l + 1, n))
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    lib.ao_shell_deriv1(ctypes.c_int(l), ctypes.c_int(3), p(alpha), p(np.ascontiguousarray(coeff[0])),
                        ctypes.c_double(FAC_SP.get(l, 1.0)), ctypes.c_longlong(n), p(np.ascontiguousarray(dr)), p(out))
    assert abs(ref).max() > 0.1
    assert abs(out - ref).max() < 1e-13 * max(1.0, abs(ref).max())


@pytest.mark.parametrize('which', ['he_c', 'c8'])
def test_whole_grid_with_image_screening(lib, which):
    """The kernel's per-point accumulation (shared routine) with the kernel's block / image-list structure
    on real cells -- lattice sums, two contractions per shell, s/p/d shells -- against the oracle."""
    import oracle
    from util import diamond_c8, he_c_cell
    cell = he_c_cell() if which == 'he_c' else diamond_c8([9, 10, 8], basis='gth-dzvp', jitter=0.15)
    rcut, Ls = cell.ao_eval_params()
    coords = np.ascontiguousarray(cell.get_uniform_grids())
    ref = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords, Ls, rcut, deriv=1).transpose(0, 2, 1)   # [4, nao, ng]
    out = np.zeros(ref.shape)
    atm = np.ascontiguousarray(cell._atm, dtype=np.int32)
    bas = np.ascontiguousarray(cell._bas, dtype=np.int32)
    env = np.ascontiguousarray(cell._env, dtype=np.float64)
    rc = np.ascontiguousarray(rcut, dtype=np.float64)
    Lc = np.ascontiguousarray(Ls, dtype=np.float64)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    lib.ao_grid_deriv1(p(atm), ctypes.c_int(len(atm)), p(bas), ctypes.c_int(len(bas)), p(env), p(rc), p(Lc),
                       ctypes.c_int(len(Lc)), ctypes.c_longlong(len(coords)), p(coords), p(out))
    assert abs(ref[1:]).max() > 0.1
    assert abs(out - ref).max() < 1e-12
