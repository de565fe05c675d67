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
