"""CPU: host-side local potentials (nuclear attraction, local GTH part) against the oracle's
literal restatement of the reference (pyscf/pbc/df/fft.py:39-84) and against analytic properties."""
import numpy as np
from oracle import pseudo as opp
from util import diamond_c8, he_c_cell
from pyscf_isdf_b200 import pseudo
from pyscf_isdf_b200.cell import Cell


def _atoms(cell):
    coords = cell.atom_coords()
    charges = [float(z) for z in cell.atom_charges()]
    pps = [cell._pseudo_for(cell.atom_symbol(i)) for i in range(cell.natm)]
    return coords, charges, pps


def test_pp_loc_and_nuc_match_the_per_atom_restatement():
    for cell in (diamond_c8([12, 14, 15], jitter=0.2), he_c_cell()):      # he_c: He carries no pseudopotential
        coords, charges, pps = _atoms(cell)
        a, mesh = cell.lattice_vectors(), cell.mesh
        v0 = opp.pp_loc_on_grid(a, mesh, coords, charges, pps)
        v1 = pseudo.pp_loc_on_grid(cell)
        assert v1.shape == v0.shape and abs(v1 - v0).max() < 1e-12 * max(1.0, abs(v0).max())
        n0 = opp.nuc_on_grid(a, mesh, coords, charges)
        n1 = pseudo.nuc_on_grid(cell)
        assert abs(n1 - n0).max() < 1e-12 * max(1.0, abs(n0).max())


def test_polynomial_part_equals_the_reference_test_helper_form():
    """full local part - erf part == the C1..C4 part written as in test_pp.py:28-64."""
    cell = diamond_c8([12, 12, 12], jitter=0.1)
    coords, charges, pps = _atoms(cell)
    a, mesh = cell.lattice_vectors(), cell.mesh
    full = opp.pp_loc_on_grid(a, mesh, coords, charges, pps)
    erf_only = opp.pp_loc_on_grid(a, mesh, coords, charges, [dict(rloc=p['rloc'], cloc=[]) for p in pps])
    part2 = opp.pp_loc_part2_on_grid(a, mesh, coords, pps)
    assert abs(full - erf_only - part2).max() < 1e-12


def test_form_factor_limits():
    """G -> 0 of the non-divergent part; rloc -> 0 recovers the point charge (pp_int.py:65-71)."""
    pp = dict(rloc=0.35, cloc=[-8.5, 1.2])
    G2 = np.array([0.0, 1e-8, 4.0])
    v = pseudo.gth_form_factor(pp, 4.0, G2)
    c0 = -(2 * np.pi)**1.5 * pp['rloc']**3 * (pp['cloc'][0] + 3 * pp['cloc'][1])
    assert abs(v[0] - (-2 * np.pi * 4.0 * pp['rloc']**2 + c0)) < 1e-12
    # just off G=0 the erf part is Z (4 pi/G^2 - 2 pi rloc^2) + O(G^2): continuous with the G=0 value once 4 pi Z/G^2 is removed
    assert abs((v[1] - 4.0 * 4 * np.pi / G2[1]) - v[0]) < 1e-6
    tiny = pseudo.gth_form_factor(dict(rloc=1e-9, cloc=[]), 3.0, G2[2:])
    assert abs(tiny[0] - 3.0 * 4 * np.pi / 4.0) < 1e-12


def test_translation_and_neutral_sum():
    """Moving every atom by one grid step rolls the potential; its grid sum is ng * vG[0] / ng."""
    cell = diamond_c8([10, 10, 10], jitter=0.2)
    v = pseudo.pp_loc_on_grid(cell).reshape(10, 10, 10)
    shifted = Cell(atom=[('C', tuple((r + cell.lattice_vectors()[0] / 10) * 0.52917721092)) for r in cell.atom_coords()],
                   a=cell.lattice_vectors() * 0.52917721092, basis='gth-dzvp', pseudo='gth-pade', mesh=[10] * 3).build()
    v2 = pseudo.pp_loc_on_grid(shifted).reshape(10, 10, 10)
    assert abs(np.roll(v, 1, axis=0) - v2).max() < 1e-9
    # sum over the grid = G=0 component: -sum_a (-2 pi Z rloc^2 - (2pi)^1.5 rloc^3 (C1 + 3 C2))
    pp = cell._pseudo_for('C')
    g0 = -2 * np.pi * 4 * pp['rloc']**2 - (2 * np.pi)**1.5 * pp['rloc']**3 * (pp['cloc'][0] + 3 * pp['cloc'][1])
    assert abs(v.sum() - (-8 * g0)) < 1e-9
    assert abs(pseudo.nuc_on_grid(cell).sum()) < 1e-9          # coulG(0) = 0


def test_has_nonlocal():
    assert pseudo.has_nonlocal(diamond_c8([8, 8, 8]))
    he = Cell(atom='He 0 0 0; He 1 1 1', a=np.eye(3) * 3, basis='gth-szv', pseudo='gth-pade', mesh=[9] * 3).build()
    assert not pseudo.has_nonlocal(he)


# ---------------------------------------------------------------------------
# reference fingerprints: pin the oracle (k-point AO values, structure factors, local form
# factor, non-local projectors, analytic AO transforms) to numbers the reference's own tests hold
# ---------------------------------------------------------------------------
def _fp(a):
    """pyscf.lib.fp: dot(cos(arange(size)), a.ravel())."""
    return np.dot(np.cos(np.arange(a.size)), a.ravel())


# pyscf/pbc/df/test/test_fft.py:589-611 (fixture :545-558: np.random.seed(1); kpts = random((4,3)); kpts[3] = k0-k1+k2)
GOLD_NUC = [(-5.7646608099493841 + 0.19126294430138713j), (-5.6567258309199193 + 0.86813371243952175j),
            (-6.1528952645454895 + 0.09517054428060109j), (-5.7445962879770942 + 0.24611951427601772j)]
GOLD_PP = [(-5.6240249083785869 + 0.22094834302524968j), (-5.5387702576467603 + 1.0439333717227581j),
           (-6.0530899866313366 + 0.2817289667029651j), (-5.6011543542444446 + 0.27597306418805201j)]


def test_oracle_reproduces_the_reference_fingerprints_of_get_nuc_and_get_pp():
    import oracle
    cell = he_c_cell()
    rs = np.random.RandomState(1)
    kpts = rs.random_sample((4, 3))
    kpts[3] = kpts[0] - kpts[1] + kpts[2]
    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts)
    coords, charges, pps = _atoms(cell)
    a, mesh = cell.lattice_vectors(), cell.mesh
    vne = opp.nuc_on_grid(a, mesh, coords, charges)
    vloc = opp.pp_loc_on_grid(a, mesh, coords, charges, pps)
    for k in range(4):
        A = ao[k]
        assert abs(_fp((A.conj().T * vne).dot(A)) - GOLD_NUC[k]) < 2e-8          # assertAlmostEqual(..., 8)
        vpp = (A.conj().T * vloc).dot(A) + opp.pp_nl_matrix(a, mesh, cell._atm, cell._bas, cell._env, coords, pps, kpts[k])
        assert abs(_fp(vpp) - GOLD_PP[k]) < 2e-8


def _si_cell(mesh=(14, 15, 16)):
    """Si: two s projectors with an off-diagonal h and one p projector."""
    return Cell(atom='Si 0 0 0; Si 1.3 1.4 1.2', a=np.eye(3) * 4.0 + 0.1, basis='gth-dzvp',
                pseudo='gth-pade', mesh=list(mesh)).build()


def test_projector_part_matches_the_per_atom_restatement():
    for cell in (he_c_cell(), diamond_c8([12, 14, 15], jitter=0.2), _si_cell()):
        coords, charges, pps = _atoms(cell)
        v0 = opp.pp_nl_matrix(cell.lattice_vectors(), cell.mesh, cell._atm, cell._bas, cell._env, coords, pps)
        v1 = pseudo.pp_nl_matrix(cell)
        assert abs(v0).max() > 0.1 and abs(v1 - v0.real).max() < 1e-12
        assert abs(v1 - v1.T).max() < 1e-13
        # disjoint ranges of reciprocal vectors add up (how ranks share the sum); small batches
        Z = pseudo.pp_nl_overlaps(cell, 0, 1000, max_bytes=1 << 20) + pseudo.pp_nl_overlaps(cell, 1000, None)
        assert abs(pseudo.pp_nl_from_overlaps(cell, Z) - v1).max() < 1e-13
    assert pseudo.has_nonlocal(he_c_cell()) and not pseudo.has_nonlocal(
        Cell(atom='He 0 0 0', a=np.eye(3) * 3, basis='gth-szv', pseudo='gth-pade', mesh=[8] * 3).build())


def test_projector_overlaps_equal_the_real_space_integrals():
    """<p_i^lm | chi> from the reciprocal-space route against a grid quadrature of the projectors'
    real-space definition  p_i^l(r) = sqrt(2) r^(l+2(i-1)) exp(-r^2/2r_l^2) / (r_l^(l+(4i-1)/2) sqrt(Gamma(l+(4i-1)/2)))
    (the Mathematica line of pyscf/pbc/gto/pseudo/pp.py:153) in a box where both are band limited:
    checks the q_li table, the analytic AO transforms and every normalisation independently."""
    import oracle
    from math import gamma
    bohr = 0.52917721092
    cell = Cell(atom=[('Si', (4.1 * bohr, 5.2 * bohr, 4.7 * bohr))], a=np.eye(3) * 10 * bohr, basis='gth-dzvp',
                pseudo='gth-pade', mesh=[40] * 3).build()
    rcut, Ls = cell.ao_eval_params()
    r = cell.get_uniform_grids()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, r, Ls, rcut)
    d = r - cell.atom_coords()[0]
    d -= np.round(d / 10.0) * 10.0
    rr = np.einsum('gx,gx->g', d, d)
    rows, ls = [], []
    for l, (rl, nl, _) in enumerate(cell._pseudo_for('Si')['nonlocal_']):
        for i in range(1, nl + 1):
            rad = np.sqrt(2.) * rr**(i - 1) * np.exp(-rr / (2 * rl * rl)) / (rl**(l + (4 * i - 1) / 2.) * np.sqrt(gamma(l + (4 * i - 1) / 2.)))
            rows.append(pseudo.solid_harmonics(l, d) * rad)
            ls += [l] * (2 * l + 1)
    ov_real = np.concatenate(rows).dot(ao) * cell.vol / len(r)
    Z = pseudo.pp_nl_overlaps(cell) / cell.vol
    ov_G = np.array([(1j)**l * z for l, z in zip(ls, Z)])      # the reference drops the projector's (-i)^l: it cancels in Z^H h Z
    assert abs(ov_real).max() > 0.1
    assert abs(ov_G.imag).max() < 1e-12 and abs(ov_G.real - ov_real).max() < 1e-8


def test_kmesh_fold_of_hcore_matches_the_kpoint_oracle():
    """get_nuc / get_pp on a uniform k-point mesh are served by the supercell (ISDF._folded, kmesh.py): the
    supercell's Gamma-point matrices -- here from the product's host routines and a NumPy contraction where the
    device contraction stands in the product -- folded to the k-points equal the k-point oracle, which reproduces
    the reference's fingerprints (test above)."""
    import oracle
    from pyscf_isdf_b200.kmesh import KMesh
    # The local parts fold exactly on any mesh (same real-space points).  The projector part is a sum over the
    # mesh's plane waves k + G (pyscf/pbc/df/fft.py:100-141): the union of those sets over the k-mesh is the
    # supercell's set shifted by less than one primitive reciprocal vector at the edge, so the two agree to the
    # projectors' weight at the plane-wave cutoff -- 1e-3 on an 11-point mesh, below 1e-8 on this one (21 points)
    cell = he_c_cell()
    kpts = cell.make_kpts([3, 1, 1])
    km = KMesh(cell, kpts)
    sc = km.supercell
    rcut, Ls = sc.ao_eval_params()
    ao_sc = oracle.eval_ao(sc._atm, sc._bas, sc._env, sc.get_uniform_grids(), Ls, rcut)
    contract = lambda v: (ao_sc.T * v).dot(ao_sc)          # fft.py:55-56, :79-81: the grid potentials carry the weight
    nuc_k = km.mat_to_kpts(contract(pseudo.nuc_on_grid(sc)))
    pp_k = km.mat_to_kpts(contract(pseudo.pp_loc_on_grid(sc)) + pseudo.pp_nl_matrix(sc))

    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts)
    coords, charges, pps = _atoms(cell)
    a, mesh = cell.lattice_vectors(), cell.mesh
    vne = opp.nuc_on_grid(a, mesh, coords, charges)
    vloc = opp.pp_loc_on_grid(a, mesh, coords, charges, pps)
    for k in range(3):
        A = ao[k]
        ref_nuc = (A.conj().T * vne).dot(A)
        ref_pp = (A.conj().T * vloc).dot(A) + opp.pp_nl_matrix(a, mesh, cell._atm, cell._bas, cell._env, coords, pps, kpts[k])
        assert abs(nuc_k[k] - ref_nuc).max() < 1e-8, abs(nuc_k[k] - ref_nuc).max()
        assert abs(pp_k[k] - ref_pp).max() < 1e-8, abs(pp_k[k] - ref_pp).max()
