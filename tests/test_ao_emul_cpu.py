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


KSRC = os.path.join(HERE, 'emul', 'ao_kernel_emul.cpp')
KOUT = os.path.join(HERE, 'emul', '_build', 'libao_kernel_emul.so')
KDEPS = [KSRC] + [os.path.join(HERE, '..', 'pyscf_isdf_b200', 'csrc', f) for f in ('ao_core.cuh', 'eval_ao_deriv1_kernel.cuh')]


@pytest.fixture(scope='module')
def klib():
    os.makedirs(os.path.dirname(KOUT), exist_ok=True)
    if not os.path.exists(KOUT) or any(os.path.getmtime(d) > os.path.getmtime(KOUT) for d in KDEPS):
        subprocess.check_call(['g++', '-std=c++20', '-O1', '-fPIC', '-shared', '-pthread', '-Wno-unknown-pragmas',
                               '-o', KOUT, KSRC])
    return ctypes.CDLL(KOUT)


def _run_kernel(klib, cell, g0, g1, ny, fill=7.0):
    rcut, Ls = cell.ao_eval_params()
    nao = cell.nao
    ld = (g1 - g0 + 7) // 8 * 8 + 8
    out = np.full((4, nao, ld), fill)
    arrs = [np.ascontiguousarray(cell._atm, dtype=np.int32), np.ascontiguousarray(cell._bas, dtype=np.int32),
            np.ascontiguousarray(cell._env, dtype=np.float64), np.ascontiguousarray(cell.lattice_vectors(), dtype=np.float64),
            np.ascontiguousarray(cell.mesh, dtype=np.int32), np.ascontiguousarray(Ls, dtype=np.float64),
            np.ascontiguousarray(rcut, dtype=np.float64)]
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    atm, bas, env, a, mesh, Lc, rc = arrs
    rcode = klib.emul_eval_ao_deriv1(p(atm), ctypes.c_int(len(atm)), p(bas), ctypes.c_int(len(bas)), p(env),
                                     ctypes.c_int(len(env)), p(a), p(mesh), p(Lc), ctypes.c_int(len(Lc)), p(rc),
                                     ctypes.c_longlong(g0), ctypes.c_longlong(g1), p(out), ctypes.c_longlong(ld),
                                     ctypes.c_longlong(nao * ld), ctypes.c_int(ny))
    assert rcode == 0
    return out


def test_the_kernel_itself_block_by_block(klib):
    """eval_ao_deriv1_kernel compiled with g++ and run CTA by CTA (64 threads, barriers, emulated warp
    shuffles / ballots): whole grid, a slab that starts and ends inside 56-point blocks, atoms split over
    grid.y -- values and gradients equal the oracle, nothing outside [g0, g1) is written."""
    import oracle
    from util import diamond_c8, he_c_cell
    for cell, ny in ((he_c_cell(), 1), (diamond_c8([7, 8, 9], basis='gth-dzvp', jitter=0.15), 4)):
        rcut, Ls = cell.ao_eval_params()
        ref = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, deriv=1).transpose(0, 2, 1)
        ng = ref.shape[2]
        out = _run_kernel(klib, cell, 0, ng, ny)
        assert abs(out[:, :, :ng] - ref).max() < 1e-12
        assert (out[:, :, ng:] == 7.0).all()
        g0, g1 = 100, 100 + 333
        part = _run_kernel(klib, cell, g0, g1, 1)
        assert abs(part[:, :, :g1 - g0] - ref[:, :, g0:g1]).max() < 1e-12
        assert (part[:, :, g1 - g0:] == 7.0).all()
