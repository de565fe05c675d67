"""CPU: pin the oracle to the reference's own known-answer tests."""
import numpy as np
import oracle
from pyscf_isdf_b200.cell import Cell, cartesian_prod, madelung
from util import he_c_cell, oracle_phi


def test_fft_matches_numpy():
    # pyscf/pbc/tools/test/test_pbc.py:185-217
    rng = np.random.default_rng(0)
    for mesh in ([31, 31, 31], [31, 31, 8], [8, 31, 8]):
        f = rng.random((2, int(np.prod(mesh)))) + 0j
        ref = np.fft.fftn(f.reshape(-1, *mesh), axes=(1, 2, 3)).reshape(2, -1)
        assert abs(oracle.fft(f, mesh) - ref).max() < 1e-10
        ref = np.fft.ifftn(f.reshape(-1, *mesh), axes=(1, 2, 3)).reshape(2, -1)
        assert abs(oracle.ifft(f, mesh) - ref).max() < 1e-10


def test_coulG_goldens():
    # pyscf/pbc/tools/test/test_pbc.py:54-80
    np.random.seed(19)
    kpt = np.random.random(3)
    a = (np.array(((0., 1.7834, 1.7834), (1.7834, 0., 1.7834), (1.7834, 1.7834, 0.)))
         + np.random.random((3, 3)).T) / 0.52917721092
    assert abs(oracle.fp(oracle.get_coulG(a, [11, 9, 7], kpt)) - 62.75448804333378) < 1e-9
    a = np.eye(3)
    k = np.array([0, np.pi, 0])
    assert abs(oracle.fp(oracle.get_coulG(a, [11, 9, 7], k)) - 4.6737453679713905) < 1e-9
    assert abs(oracle.fp(oracle.get_coulG(a, [11, 9, 7], k, wrap_around=False)) - 4.5757877990664744) < 1e-9
    c2 = Cell(atom='C 0. 0. 0.; C 0.8917 0.8917 0.8917', a=np.eye(3), unit='B',
              basis='gth-szv', pseudo='gth-pade', mesh=[11, 9, 7]).build()
    g = oracle.get_coulG(a, [11, 9, 7], exx='ewald', madelung=madelung(c2))
    assert abs(oracle.fp(g) - 4.888843468914021) < 1e-9


def test_eval_ao_d_shell_golden():
    # pyscf/pbc/gto/test/test_cell.py:500-504 (d shell, 27 k-points)
    cell = Cell(a=np.eye(3) * 4, atom='He 1 1 1', basis=[[2, (1, .5), (.5, .5)]],
                pseudo=None, precision=1e-10).build()
    coords = cell.get_uniform_grids([10] * 3, wrap_around=False)
    kpts = cartesian_prod([np.arange(3) / 3.] * 3).dot(cell.reciprocal_vectors())
    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords, Ls, rcut, kpts=kpts)
    assert abs(oracle.fp(ao) - (-0.27594803231989179 + 0.0064644591759109114j)) < 1e-9


def test_fftdf_jk_golden():
    # pyscf/pbc/df/test/test_fft.py:642-645 (dm = eye(int), exxdiv=None, Gamma point)
    cell = he_c_cell()
    ao = oracle_phi(cell).T
    dm = np.eye(cell.nao, dtype=int)
    vj, vk = oracle.fftdf_get_jk(ao, dm, cell.lattice_vectors(), cell.mesh)
    assert abs(oracle.fp(vj) - 3.7955873127283377) < 1e-8
    assert abs(oracle.fp(vk) - 4.290076429522121) < 1e-8


def test_ewald_goldens():
    # pyscf/pbc/gto/test/test_cell.py:117-146: the result does not depend on the splitting parameter
    c = Cell(unit='B', a=np.diag([5., 5., 5.]), mesh=[41] * 3, atom=[['He', (2, 2.5, 2.5)], ['He', (3, 2.5, 2.5)]],
             basis={'He': [[0, (1.0, 1.0)]]}, pseudo=None).build()
    for eta, cut in ((0.2, 30), (1, 30), (None, None)):
        assert abs(c.ewald(eta, cut) - (-0.468640671931)) < 1e-9
    rs = np.random.RandomState(10)
    a = rs.random_sample((3, 3)) * 2 + np.eye(3) * 2
    c = Cell(a=a, mesh=[41] * 3, atom=[['He', (1, 1, 2)], ['He', (3, 2, 1)]], basis={'He': [[0, (1.0, 1.0)]]},
             pseudo=None).build()
    for eta, cut in ((1, 20), (2, 10), (2, 5), (None, None)):
        assert abs(c.ewald(eta, cut) - (-2.3711356723457615)) < 1e-9


def test_cell_goldens_gv_si_lattice_images_and_overlap():
    """Host-side Cell pieces against the reference's own numbers (pyscf/pbc/gto/test/test_cell.py:
    :40-51 fixture, :65-76 get_Gv / get_SI, :93-114 number of lattice images, :280-292 Gamma-point overlap)."""
    from pyscf_isdf_b200 import int1e as pint
    from test_xc_cpu import HE_CCPVDZ
    rs = np.random.RandomState(1)
    cl1 = Cell(a=rs.random_sample((3, 3)).T, precision=1e-9, mesh=[41] * 3, pseudo=None, basis={'He': HE_CCPVDZ},
               atom='He .1 .0 .0; He .5 .1 .0; He .0 .5 .0; He .1 .3 .2').build()
    Gv = cl1.get_Gv()
    assert abs(oracle.fp(Gv) - (-99.791927068519939)) < 1e-9
    assert abs(oracle.fp(cl1.get_SI(Gv)) - (16.506917823339265 + 1.6393578329869585j)) < 1e-9
    G5 = np.random.RandomState(2).random_sample((5, 3))
    assert abs(oracle.fp(cl1.get_SI(G5)) - (0.65237631847195221 - 1.5736011413431059j)) < 1e-12
    # analytic lattice-summed overlap of the reference (it asserts 4 places) against the reciprocal-space sum
    S, T = pint.overlap_kinetic(cl1)
    assert abs(oracle.fp(S) - 492.30658304804126) < 1e-4
    c2 = Cell(atom='C 0 0 0; C 1.685068664391 1.685068664391 1.685068664391', unit='B', basis='gth-dzvp',
              pseudo='gth-pade', mesh=[15] * 3,
              a=[[0, 3.370137329, 3.370137329], [3.370137329, 0, 3.370137329], [3.370137329, 3.370137329, 0]]).build()
    assert c2.get_lattice_Ls(rcut=c2.rcut).shape == (1439, 3)


def test_c2s_orthonormal():
    """The restated libcint cart->sph matrices give orthonormal real harmonics."""
    from oracle.eval_ao import C2S, cart_powers
    x, w = np.polynomial.legendre.leggauss(24)
    phi = (np.arange(48) + .5) * 2 * np.pi / 48
    ct, ph = np.meshgrid(x, phi, indexing='ij')
    st = np.sqrt(1 - ct**2)
    xyz = np.stack([st * np.cos(ph), st * np.sin(ph), ct]).reshape(3, -1)
    wt = (w[:, None] * np.ones_like(ph) * 2 * np.pi / 48).ravel()
    for l, c2s in C2S.items():
        cart = np.stack([xyz[0]**a * xyz[1]**b * xyz[2]**c for a, b, c in cart_powers(l)])
        Y = c2s.dot(cart)
        S = (Y * wt).dot(Y.T)
        assert abs(S - np.eye(2 * l + 1)).max() < 1e-12


def test_isdf_oracle_exact_limit():
    """All grid points as interpolation points reproduce the FFTDF J/K."""
    cell = Cell(atom='He 1. .5 .5; C .1 1.3 2.1',
                basis={'He': [(0, (2.5, 1)), (0, (1., 1))], 'C': 'gth-szv'},
                pseudo={'C': 'gth-pade'}, a=np.eye(3) * 2.5, mesh=[7] * 3).build()
    phi = oracle_phi(cell)
    a, mesh = cell.lattice_vectors(), cell.mesh
    nao, ng = phi.shape
    rng = np.random.default_rng(1)
    dm = rng.random((nao, nao)); dm = dm + dm.T
    vj0, vk0 = oracle.fftdf_get_jk(phi.T, dm, a, mesh)
    # rank of the pair-product space is nao(nao+1)/2 = 21 << ng: ISDF is exact once Np >= 21
    ip, resid = oracle.select_ip(phi, 40)
    assert len(ip) <= 21 + 1
    zeta = oracle.fit_aux(phi, ip)
    W = oracle.build_W(zeta, a, mesh)
    vj, vk = oracle.isdf_get_jk(phi[:, ip], W, dm)
    assert abs(vj - vj0).max() < 1e-9
    assert abs(vk - vk0).max() < 1e-9


def test_golden_files_consistent():
    """tests/golden: the reference KAT table is the one this file checks, and the
    self-generated ISDF fixture is reproduced by the oracle (regeneration check)."""
    import json, os
    from oracle import c_oracle
    from util import diamond_c8
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
    kats = json.load(open(os.path.join(here, 'reference_kats.json')))
    assert kats['fftdf_vj_eye']['value'] == 3.7955873127283377
    assert kats['coulG_ewald']['value'] == 4.888843468914021
    z = np.load(os.path.join(here, 'isdf_small.npz'))
    cell = diamond_c8(list(z['mesh']), basis='gth-dzvp', jitter=float(z['jitter']))
    phi = oracle_phi(cell)
    assert abs(oracle.fp(phi) - float(z['phi_fp'])) < 1e-9
    ip, resid = c_oracle.select_ip_chain(phi, int(z['c']) * cell.nao)
    assert np.array_equal(ip, z['ip']) and np.array_equal(resid, z['resid'])


def _bloch_ao(cell, kpts):
    rcut, Ls = cell.ao_eval_params()
    return oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts)


def test_fftdf_k_kpts_golden():
    # pyscf/pbc/df/test/test_fft.py:536-544 (kpts), :664-676 (kpts_band, occupied-orbital tags): the k-point
    # oracle K is pinned to the reference's known answer, both through the tags and through the plain matrices
    from oracle.fftdf_jk import fftdf_get_k_kpts
    cell = he_c_cell()
    nao = cell.nao
    np.random.seed(1)
    kpts = np.random.random((4, 3))
    kpts[3] = kpts[0] - kpts[1] + kpts[2]
    np.random.seed(1)
    kpts_band = np.random.random((2, 3))
    mo_coeff = np.random.random((4, nao, nao))
    mo_occ = np.array(np.random.random((4, nao)) > .6, dtype=np.double)
    dms = np.einsum('kpi,ki,kqi->kpq', mo_coeff, mo_occ, mo_coeff)
    ao_k, ao_b = _bloch_ao(cell, kpts), _bloch_ao(cell, kpts_band)
    a, mesh, coords = cell.lattice_vectors(), cell.mesh, cell.get_uniform_grids()
    gold = 10.239828255099447 + 2.1190549216896182j
    vk = fftdf_get_k_kpts(ao_k, dms, a, mesh, kpts, coords, ao_band=ao_b, kpts_band=kpts_band, mo=(mo_coeff, mo_occ))
    assert abs(oracle.fp(vk) - gold) < 1e-8
    vk = fftdf_get_k_kpts(ao_k, dms, a, mesh, kpts, coords, ao_band=ao_b, kpts_band=kpts_band)
    assert abs(oracle.fp(vk) - gold) < 1e-8


def test_kmesh_supercell_equivalence():
    # pyscf/pbc/tools/k2gamma.py: J and K on a uniform k-point mesh are the Bloch transforms of the supercell's
    # Gamma-point J and K (the product's route to k-points, pyscf_isdf_b200/kmesh.py), here with both sides exact
    from oracle.fftdf_jk import fftdf_get_j_kpts, fftdf_get_k_kpts
    from pyscf_isdf_b200.kmesh import KMesh, detect_kmesh
    cell = Cell(atom='He 1. .5 .5; He .1 1.3 2.1', basis={'He': [(0, (2.5, 1)), (0, (1., 1))]}, pseudo=None,
                a=np.eye(3) * 2.5 + 0.1, mesh=[7, 9, 7], unit='B').build()
    nao = cell.nao
    assert detect_kmesh(cell, cell.make_kpts([3, 1, 2]))[0] == (3, 1, 2)
    kpts = cell.make_kpts([3, 1, 3])
    perm = np.array([2, 0, 5, 1, 4, 3, 8, 7, 6])
    kpts = kpts[perm]                                  # any order
    assert detect_kmesh(cell, kpts)[0] == (3, 1, 3)
    assert detect_kmesh(cell, kpts[:5]) is None and detect_kmesh(cell, kpts + 0.01) is None
    km = KMesh(cell, kpts)
    sc = km.supercell
    assert list(sc.mesh) == [21, 9, 21] and sc.nao == 9 * nao      # all odd: exact up to the AO lattice-sum cutoff (cell.precision)
    rng = np.random.default_rng(3)
    dms = rng.random((9, nao, nao)) + 1j * rng.random((9, nao, nao))
    dms = dms + dms.conj().transpose(0, 2, 1)          # hermitian, no time-reversal symmetry: D_sc is complex
    ao_k = _bloch_ao(cell, kpts)
    a, mesh, coords = cell.lattice_vectors(), cell.mesh, cell.get_uniform_grids()
    vj_ref = fftdf_get_j_kpts(ao_k, dms, a, mesh)
    vk_ref = fftdf_get_k_kpts(ao_k, dms, a, mesh, kpts, coords)

    D = km.dm_to_supercell(dms)
    assert abs(D - D.conj().T).max() < 1e-12 and abs(D.imag).max() > 1e-3
    ao_sc = oracle_phi(sc).T

    class ExactSupercellDF:                    # stands where the supercell's ISDF object stands in the product
        def get_jk(self, dm, hermi=1, with_j=True, with_k=True, exxdiv=None):
            assert dm.dtype == np.float64 and dm.shape == (2, 9 * nao, 9 * nao) and hermi == 0
            return oracle.fftdf_get_jk(ao_sc, dm, sc.lattice_vectors(), sc.mesh, with_j, with_k)

    class Stub:
        _kmesh, _sc = km, ExactSupercellDF()

    from pyscf_isdf_b200.kmesh import get_jk_kmesh
    vj, vk = get_jk_kmesh(Stub(), dms)
    assert vj.shape == vk.shape == dms.shape
    assert get_jk_kmesh(Stub(), dms, with_k=False)[1] is None
    assert abs(vj - vj_ref).max() < 1e-8, abs(vj - vj_ref).max()
    assert abs(vk - vk_ref).max() < 1e-8, abs(vk - vk_ref).max()
