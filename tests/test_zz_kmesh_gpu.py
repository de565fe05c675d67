"""GPU: J and K on a uniform k-point mesh, served by the Gamma-point ISDF of the supercell (pyscf_isdf_b200/kmesh.py).

Checked against the k-point FFTDF oracle (oracle/fftdf_jk.py, pinned to pyscf/pbc/df/test/test_fft.py:664-676 in
tests/test_oracle_golden.py) in the exact limit: number of interpolation points >= rank of the supercell's pair space.
"""
import numpy as np
import pytest
import oracle
from oracle.fftdf_jk import fftdf_get_j_kpts, fftdf_get_k_kpts
from util import he_c_cell
from pyscf_isdf_b200.isdf import ISDF
from pyscf_isdf_b200.cell import madelung

pytestmark = pytest.mark.gpu


def _bloch_ao(cell, kpts):
    rcut, Ls = cell.ao_eval_params()
    return oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts)


def test_kmesh_get_jk_exact_limit_matches_kpoint_fftdf():
    cell = he_c_cell()                                  # mesh 21^3, nao 6
    nao = cell.nao
    kpts = cell.make_kpts([3, 1, 1])                    # supercell: nao 18, mesh 63 x 21 x 21 (all odd)
    # 180 points offered, the pair space has dimension <= 171.  The default stop (ng eps max d, the reference's
    # pyscf/lib/scipy_helper.py:88-90) drops pair products of norm ~1e-5 and leaves 1e-6 in J; tol = 1e-15 keeps them
    # (tools/kmesh_tol_scan.py: 1.4e-6 / 4e-8 / 5e-9 at the default / 1e-14 / 1e-15)
    mydf = ISDF(cell, kpts=kpts, n_ip=60, tol=1e-15)
    rng = np.random.default_rng(11)
    dms = rng.random((3, nao, nao)) + 1j * rng.random((3, nao, nao))
    dms = dms + dms.conj().transpose(0, 2, 1)           # hermitian, not time-reversal symmetric
    vj, vk = mydf.get_jk(dms, kpts=kpts, exxdiv=None)
    assert vj.shape == vk.shape == (3, nao, nao) and vj.dtype == vk.dtype == np.complex128   # test_fft.py:654-655
    assert mydf._sc.n_ip < 180                          # stopped by the residual, not by the count

    ao_k = _bloch_ao(cell, kpts)
    a, mesh, coords = cell.lattice_vectors(), cell.mesh, cell.get_uniform_grids()
    vj0 = fftdf_get_j_kpts(ao_k, dms, a, mesh)
    vk0 = fftdf_get_k_kpts(ao_k, dms, a, mesh, kpts, coords)
    assert abs(vj - vj0).max() < 1e-7, abs(vj - vj0).max()
    assert abs(vk - vk0).max() < 1e-7, abs(vk - vk0).max()

    # exxdiv='ewald': + madelung(cell, kpts) S_k D_k S_k with the Madelung constant of the supercell
    # (pyscf/pbc/df/df_jk.py:1425-1455, pyscf/pbc/tools/pbc.py:483-493)
    vk_e = mydf.get_jk(dms, kpts=kpts, with_j=False, exxdiv='ewald')[1]
    w = cell.vol / len(coords)
    mad = madelung(mydf._kmesh.supercell)
    for k in range(3):
        s = w * ao_k[k].conj().T.dot(ao_k[k])
        assert abs(vk_e[k] - vk0[k] - mad * s.dot(dms[k]).dot(s)).max() < 1e-7

    # sets of density matrices, J only, and the order of the k-points
    perm = [2, 0, 1]
    df2 = ISDF(cell, kpts=kpts[perm], n_ip=60, tol=1e-15)
    vj2 = df2.get_j(np.stack([dms[perm], 2 * dms[perm]]), kpts=kpts[perm])
    assert vj2.shape == (2, 3, nao, nao)
    assert abs(vj2[0] - vj[perm]).max() < 1e-9 and abs(vj2[1] - 2 * vj[perm]).max() < 1e-9


def test_kmesh_rejects_what_the_supercell_cannot_serve():
    cell = he_c_cell()
    with pytest.raises(NotImplementedError):
        ISDF(cell, kpts=np.array([[0.1, 0.2, 0.3]])).build()                 # not a uniform mesh
    kpts = cell.make_kpts([3, 1, 1])
    mydf = ISDF(cell, kpts=kpts, n_ip=60)
    with pytest.raises(NotImplementedError):
        mydf.get_jk(np.zeros((2, cell.nao, cell.nao)), kpts=kpts[:2])       # a different k-point list
    with pytest.raises(NotImplementedError):
        mydf.get_jk(np.zeros((3, cell.nao, cell.nao)), kpts=kpts, kpts_band=kpts[:1])
    with pytest.raises(NotImplementedError):
        mydf.get_nuc(kpts[:2])


def test_kmesh_hcore_and_xc_match_the_kpoint_oracle():
    """get_nuc / get_pp / nr_rks on the k-mesh: the supercell's device results folded to the k-points, against the
    k-point oracle (pinned to the reference's get_nuc / get_pp fingerprints and nr_rks numbers in the CPU tests)."""
    from oracle import pseudo as opp, xc
    cell = he_c_cell()
    nao = cell.nao
    kpts = cell.make_kpts([3, 1, 1])
    mydf = ISDF(cell, kpts=kpts, n_ip=60)
    vnuc, vpp = mydf.get_nuc(kpts), mydf.get_pp(kpts)
    assert vnuc.shape == vpp.shape == (3, nao, nao)

    rcut, Ls = cell.ao_eval_params()
    coords_g = cell.get_uniform_grids()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords_g, Ls, rcut, kpts=kpts, deriv=1)
    a, mesh = cell.lattice_vectors(), cell.mesh
    coords = cell.atom_coords()
    charges = [float(z) for z in cell.atom_charges()]
    pps = [cell._pseudo_for(cell.atom_symbol(i)) for i in range(cell.natm)]
    w = cell.vol / len(coords_g)
    vne = opp.nuc_on_grid(a, mesh, coords, charges)
    vloc = opp.pp_loc_on_grid(a, mesh, coords, charges, pps)
    for k in range(3):
        A = ao[k][0]
        assert abs(vnuc[k] - (A.conj().T * vne).dot(A)).max() < 1e-8          # fft.py:55-56: no quadrature weight
        ref = (A.conj().T * vloc).dot(A) + opp.pp_nl_matrix(a, mesh, cell._atm, cell._bas, cell._env, coords, pps, kpts[k])
        assert abs(vpp[k] - ref).max() < 1e-8

    rng = np.random.default_rng(5)
    dms = rng.random((3, nao, nao)) + 1j * rng.random((3, nao, nao))
    dms = 0.2 * (dms + dms.conj().transpose(0, 2, 1)) + np.eye(nao)          # positive density
    ne, exc, vmat = mydf.nr_rks('pbe', dms, kpts)
    rho = sum(xc.eval_rho(ao[k], dms[k]) for k in range(3)) / 3
    assert rho[0].min() > 0
    sigma = np.einsum('xg,xg->g', rho[1:], rho[1:])
    e, vrho, vsigma = xc.eval_xc('pbe', rho[0], sigma)
    wv = np.empty_like(rho)
    wv[0] = w * vrho * .5
    wv[1:] = w * 2 * vsigma * rho[1:]
    assert abs(ne - w * rho[0].sum()) < 1e-8 and abs(exc - w * rho[0].dot(e)) < 1e-8
    for k in range(3):
        mat = ao[k][0].conj().T.dot(np.einsum('npi,np->pi', ao[k], wv))
        assert abs(vmat[k] - (mat + mat.conj().T)).max() < 1e-8


def test_kmesh_lda_energy_on_the_device_reproduces_the_reference():
    """KRKS 'lda,vwn' on the reference's 2 x 2 x 1 mesh (pyscf/pbc/dft/test/test_krks.py:83-88: -0.3846075202893169 Ha)
    with every plugin call on the device path: ISDF.get_nuc, get_j (exact-density route) and nr_rks at the k-points,
    each served by the supercell and folded back; kinetic and overlap matrices from the oracle's integrals folded
    the same way (they belong to the SCF driver's side).  The CPU twin with oracle stand-ins is
    tests/test_scf_cpu.py::test_kmesh_lda_energy_through_the_supercell_reproduces_the_reference."""
    from oracle import int1e
    from scf_harness import krks_kmesh
    from test_scf_cpu import h2_kmesh_fixture, E_KLDA_H2_KPTS_221
    cell, kpts = h2_kmesh_fixture()
    mydf = ISDF(cell, kpts=kpts, exact_j=True)
    mydf.check_sanity()
    km = mydf._kmesh
    sc = km.supercell
    S, T = int1e.overlap_kinetic(sc.lattice_vectors(), sc._atm, sc._bas, sc._env)
    hcore_k = km.mat_to_kpts(T) + mydf.get_nuc(kpts)
    e, dm_k, ncyc = krks_kmesh(hcore_k, km.mat_to_kpts(S), cell.nelectron, cell.ewald(),
                               lambda dm: mydf.get_j(dm, kpts=kpts),
                               lambda dm: mydf.nr_rks('lda,vwn', dm, kpts))
    assert abs(e - E_KLDA_H2_KPTS_221) < 1e-7, e - E_KLDA_H2_KPTS_221
