"""GPU: the ISDF plugin end to end (build + get_jk through host buffers)."""
import numpy as np
import pytest
import oracle
from oracle import c_oracle
from util import diamond_c8, he_c_cell, oracle_phi, rand_dm
from pyscf_isdf_b200.isdf import ISDF

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('fit', ['cholesky', 'pivot_factor'])
def test_isdf_plugin_matches_cpu_isdf(fit):
    """IP indices bit-exact; J/K within 1e-8 of the CPU ISDF oracle (north_star tolerance)."""
    cell = diamond_c8([14, 14, 14], basis='gth-dzvp', jitter=0.15)
    mydf = ISDF(cell, c=6, fit=fit)
    dm = rand_dm(cell.nao)
    vj, vk = mydf.get_jk(dm, exxdiv=None)                 # lazy build, like SCF would trigger it
    phi = oracle_phi(cell)
    assert abs(mydf.aoR_host() - phi).max() < 1e-12
    ip, resid = c_oracle.select_ip_chain(mydf.aoR_host(), 6 * cell.nao)
    assert np.array_equal(mydf.ip_index, ip)
    assert np.array_equal(mydf.ip_resid, resid)
    ref = oracle.isdf_build(phi, cell.lattice_vectors(), cell.mesh, 6 * cell.nao)
    assert np.array_equal(ref['ip'], ip)
    vj0, vk0 = oracle.isdf_get_jk(ref['X'], ref['W'], dm)
    assert abs(vj - vj0).max() < 1e-8
    assert abs(vk - vk0).max() < 1e-8
    # J-only / K-only and shapes (pyscf/pbc/df/fft.py:298-315 convention)
    j_only, none = mydf.get_jk(dm, with_k=False)
    assert none is None and np.array_equal(j_only, vj)
    dms = np.stack([dm, 2 * dm])
    vj2, vk2 = mydf.get_jk(dms)
    assert vj2.shape == dms.shape and abs(vk2[1] - 2 * vk).max() < 1e-9


def test_isdf_exact_limit_reproduces_fftdf_goldens():
    """Np = rank of the pair space -> ISDF equals exact FFTDF J/K: reproduces the
    reference's golden fingerprints (pyscf/pbc/df/test/test_fft.py:642-645)."""
    cell = he_c_cell()
    mydf = ISDF(cell, n_ip=60)                       # pair space has dimension 21
    dm = np.eye(cell.nao, dtype=int)                 # integer DM is cast (df_jk.py:1415-1416)
    vj, vk = mydf.get_jk(dm, exxdiv=None)
    assert mydf.n_ip <= 22
    assert abs(oracle.fp(vj) - 3.7955873127283377) < 1e-7
    assert abs(oracle.fp(vk) - 4.290076429522121) < 1e-7


def test_isdf_ewald_and_convergence_with_c():
    cell = diamond_c8([14, 14, 14], basis='gth-szv', jitter=0.15)
    phi = oracle_phi(cell)
    dm = rand_dm(cell.nao)
    a, mesh = cell.lattice_vectors(), cell.mesh
    ovlp = (cell.vol / phi.shape[1]) * phi.dot(phi.T)
    from pyscf_isdf_b200.cell import madelung
    vj_x, vk_x = oracle.fftdf_get_jk(phi.T, dm, a, mesh, exxdiv='ewald', ovlp=ovlp, madelung=madelung(cell))
    errs = []
    for c in (4, 8, 16):
        mydf = ISDF(cell, c=c)
        vj, vk = mydf.get_jk(dm, exxdiv='ewald')
        errs.append(max(abs(vj - vj_x).max(), abs(vk - vk_x).max()))
    assert errs[0] > errs[1] > errs[2], errs          # tier C: error decreases with c
    assert errs[2] < 1e-3 * abs(vk_x).max(), errs


def test_isdf_save_load(tmp_path):
    cell = diamond_c8([12, 12, 12], basis='gth-szv', jitter=0.1)
    dm = rand_dm(cell.nao)
    a = ISDF(cell, c=5).build()
    vj, vk = a.get_jk(dm)
    a.save(str(tmp_path / 'isdf.npz'))
    b = ISDF(cell, c=5).load(str(tmp_path / 'isdf.npz'))
    vj2, vk2 = b.get_jk(dm)
    assert np.array_equal(vj, vj2) and np.array_equal(vk, vk2)


def test_isdf_matches_committed_golden_fixture():
    """Pivots bit-exact and W/J/K within 1e-8 of tests/golden/isdf_small.npz
    (made by tests/golden/make_golden.py; no oracle call here)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'isdf_small.npz'))
    cell = diamond_c8(list(z['mesh']), basis='gth-dzvp', jitter=float(z['jitter']))
    for fit in ('cholesky', 'pivot_factor'):
        mydf = ISDF(cell, c=int(z['c']), fit=fit).build()
        assert np.array_equal(mydf.ip_index, z['ip'])
        assert np.allclose(mydf.ip_resid, z["resid"], rtol=1e-9, atol=0)   # AO values differ in the last bits
        W = mydf.W.cpu().numpy()
        assert abs(np.diag(W) - z['W_diag']).max() < 1e-8
        assert abs(W[:16, :16] - z['W_block']).max() < 1e-8
        vj, vk = mydf.get_jk(rand_dm(cell.nao))
        assert abs(vj - z['vj']).max() < 1e-8 and abs(vk - z['vk']).max() < 1e-8


@pytest.mark.parametrize('wgemm', ['i8', 'i8:7'])
def test_isdf_int8_coulomb_metric(wgemm):
    """W through the int8 tensor-core split product (csrc/syrk_i8.cu): same pivots, W equal to the
    FP64-GEMM route to ~1e-13 relative, J/K within the 1e-8 north_star tolerance of the CPU oracle."""
    cell = diamond_c8([14, 14, 14], basis='gth-dzvp', jitter=0.15)
    dm = rand_dm(cell.nao)
    ref = ISDF(cell, c=6, fit='implicit', wgemm='fp64')
    vj0, vk0 = ref.get_jk(dm, exxdiv=None)
    mydf = ISDF(cell, c=6, fit='implicit', wgemm=wgemm)
    vj, vk = mydf.get_jk(dm, exxdiv=None)
    assert np.array_equal(mydf.ip_index, ref.ip_index)
    W0, W1 = ref.W.cpu().numpy(), mydf.W.cpu().numpy()
    assert abs(W1 - W0).max() < (1e-12 if wgemm == 'i8' else 1e-10) * abs(W0).max()
    assert abs(vj - vj0).max() < 1e-10 and abs(vk - vk0).max() < 1e-10
    orc = oracle.isdf_build(oracle_phi(cell), cell.lattice_vectors(), cell.mesh, 6 * cell.nao)
    vjo, vko = oracle.isdf_get_jk(orc['X'], orc['W'], dm)
    assert abs(vj - vjo).max() < 1e-8 and abs(vk - vko).max() < 1e-8


def test_exact_density_j_matches_reference_fftdf():
    """exact_j=True: J by the reference's FFTDF route (pyscf/pbc/df/fft_jk.py:33-109) on the device;
    reproduces the reference golden fp(vj) of pyscf/pbc/df/test/test_fft.py:642-645 and the
    oracle's FFTDF J on a second cell, for one and two density matrices; K stays ISDF."""
    cell = he_c_cell()
    mydf = ISDF(cell, n_ip=60, exact_j=True)
    dm = np.eye(cell.nao, dtype=int)
    vj, vk = mydf.get_jk(dm, exxdiv=None)
    assert abs(oracle.fp(vj) - 3.7955873127283377) < 1e-9
    assert abs(oracle.fp(vk) - 4.290076429522121) < 1e-7
    cell = diamond_c8([14, 14, 14], basis='gth-dzvp', jitter=0.15)
    mydf = ISDF(cell, c=6, exact_j=True)
    dms = np.stack([rand_dm(cell.nao), rand_dm(cell.nao)[::-1, ::-1].copy()])
    vj = mydf.get_j(dms)
    ref = oracle.fftdf_get_j(oracle_phi(cell).T, dms, cell.lattice_vectors(), cell.mesh)
    assert vj.shape == dms.shape and abs(vj - ref).max() < 1e-10
    # J-only call does not need (or touch) the ISDF contraction
    j_only, none = mydf.get_jk(dms[0], with_k=False)
    assert none is None and abs(j_only - ref[0]).max() < 1e-10


def test_tagged_density_matrix_takes_the_occupied_orbital_path():
    """A density matrix carrying mo_coeff / mo_occ (pyscf.lib.tag_array, read by the reference's K build at
    pyscf/pbc/df/fft_jk.py:206-210) goes through isdf_get_jk_block_occ: same J/K as the plain matrix to 1e-10,
    also with unoccupied columns present, fractional occupations, and exxdiv='ewald'."""
    class Tagged(np.ndarray):
        pass
    cell = diamond_c8([14, 14, 14], basis='gth-dzvp', jitter=0.15)
    nao = cell.nao
    rng = np.random.default_rng(5)
    C = np.linalg.qr(rng.standard_normal((nao, nao)))[0]
    occ = np.zeros(nao)
    occ[:16] = 2.0
    occ[16:19] = [1.3, 0.4, 0.05]
    dm = (C * occ).dot(C.T)
    mydf = ISDF(cell, c=6)
    vj0, vk0 = mydf.get_jk(dm, exxdiv='ewald')
    tagged = dm.view(Tagged)
    tagged.mo_coeff, tagged.mo_occ = C, occ
    assert mydf.last_jk_form == 'density-matrix'
    vj1, vk1 = mydf.get_jk(tagged, exxdiv='ewald')
    assert mydf.last_jk_form == 'occupied-orbital'
    assert abs(vj1 - vj0).max() < 1e-10 and abs(vk1 - vk0).max() < 1e-10
    # the oracle on the same density matrix
    orc = oracle.isdf_build(oracle_phi(cell), cell.lattice_vectors(), cell.mesh, 6 * nao)
    vjo, vko = oracle.isdf_get_jk(orc['X'], orc['W'], dm)
    vj2, vk2 = mydf.get_jk(tagged, exxdiv=None)
    assert abs(vj2 - vjo).max() < 1e-8 and abs(vk2 - vko).max() < 1e-8
    # a stack of matrices or complex orbitals falls back to the density-matrix path
    vj3, vk3 = mydf.get_jk(np.stack([dm, dm]))
    assert mydf.last_jk_form == 'density-matrix'
    assert abs(vk3[1] - vk2).max() < 1e-10
