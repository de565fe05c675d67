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
