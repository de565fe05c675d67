"""GPU: hcore pieces on the device-resident AO values ("next" row 2 of SURVEY 8f):
ISDF.get_nuc / get_pp_loc / get_pp against the oracle's restatement of FFTDF.get_nuc / get_pp
(pyscf/pbc/df/fft.py:39-152; the oracle itself is pinned to the reference's fingerprints of both
matrices, tests/test_pseudo_cpu.py).  Tolerance 1e-10 abs (one FP64 contraction over the grid).

(File name sorts last: this row was added after the last GPU session of round 1.)"""
import numpy as np
import pytest
from oracle import pseudo as opp
from util import diamond_c8, he_c_cell, oracle_phi
from pyscf_isdf_b200.cell import Cell
from pyscf_isdf_b200.isdf import ISDF

pytestmark = pytest.mark.gpu


def _ref(cell):
    coords = cell.atom_coords()
    charges = [float(z) for z in cell.atom_charges()]
    pps = [cell._pseudo_for(cell.atom_symbol(i)) for i in range(cell.natm)]
    a, mesh = cell.lattice_vectors(), cell.mesh
    ao = oracle_phi(cell).T
    vpp = opp.potential_matrix(ao, opp.pp_loc_on_grid(a, mesh, coords, charges, pps))
    vne = opp.potential_matrix(ao, opp.nuc_on_grid(a, mesh, coords, charges))
    return vpp, vne


def _ref_nl(cell):
    coords = cell.atom_coords()
    pps = [cell._pseudo_for(cell.atom_symbol(i)) for i in range(cell.natm)]
    return opp.pp_nl_matrix(cell.lattice_vectors(), cell.mesh, cell._atm, cell._bas, cell._env, coords, pps).real


@pytest.mark.parametrize('which', ['c8', 'he_c'])
def test_local_potentials_match_oracle(which):
    cell = diamond_c8([15, 14, 12], basis='gth-dzvp', jitter=0.15) if which == 'c8' else he_c_cell()
    vpp0, vne0 = _ref(cell)
    mydf = ISDF(cell, c=4)
    vne = mydf.get_nuc()                      # before any build: evaluates the AO values itself
    assert mydf.W is None
    assert vne.shape == vne0.shape and abs(vne - vne0).max() < 1e-10
    mydf.build()
    vpp = mydf.get_pp_loc(np.zeros(3))
    assert abs(vpp - vpp0).max() < 1e-10
    assert abs(vpp - vpp.T).max() < 1e-12
    vnl0 = _ref_nl(cell)                      # C carries non-local projectors
    assert abs(vnl0).max() > 1e-3
    assert abs(mydf.get_pp() - (vpp0 + vnl0)).max() < 1e-10
    with pytest.raises(NotImplementedError):
        mydf.get_nuc(np.array([0.1, 0, 0]))


def test_get_pp_without_projectors_is_complete():
    cell = Cell(atom='He 0.3 0.2 0.1; He 1.6 1.4 1.5', a=np.eye(3) * 3, basis='gth-dzvp',
                pseudo='gth-pade', mesh=[15] * 3).build()
    vpp0, _ = _ref(cell)
    assert abs(ISDF(cell, c=4).get_pp() - vpp0).max() < 1e-10


def test_sharded_local_potential():
    """Grid sharded over 3 ranks (threads on one GPU): slab contractions summed."""
    import threading
    from test_zzy_dist_gpu import ThreadComm
    cell = diamond_c8([13, 10, 12], basis='gth-szv', jitter=0.2)
    vpp0, _ = _ref(cell)
    n = 3
    shared = ThreadComm.Shared(n)
    out, full, err = [None] * n, [None] * n, []

    def work(r):
        try:
            mydf = ISDF(cell, c=6, comm=ThreadComm(shared, r))
            mydf.build()
            out[r] = mydf.get_pp_loc()
            full[r] = mydf.get_pp()
        except Exception as e:        # noqa: BLE001
            err.append(e)
            shared.barrier.abort()

    ths = [threading.Thread(target=work, args=(r,)) for r in range(n)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(timeout=600)
    if err:
        raise err[0]
    vnl0 = _ref_nl(cell)
    for r in range(n):
        assert abs(out[r] - vpp0).max() < 1e-10
        assert abs(full[r] - (vpp0 + vnl0)).max() < 1e-10


def test_nonlocal_projectors_device_equals_host_with_p_channels_and_batches():
    """Si carries two s projectors and a p projector: the device form (isdf_phase_expand + isdf_dgemm, several
    batches of G vectors, a sub-range as a rank would take it) against the host form, and the matrix against the oracle."""
    from pyscf_isdf_b200 import pseudo
    a = 5.431
    frac = np.array([[0, 0, 0], [.25, .25, .25], [.5, .5, 0], [.75, .75, .25]])
    cell = Cell(atom=[('Si', tuple(r)) for r in frac * a] + [('C', (0.5 * a, 0.1 * a, 0.6 * a))],
                a=np.eye(3) * a, basis='gth-dzvp', pseudo='gth-pade', mesh=[13, 12, 11]).build()
    mydf = ISDF(cell, c=4)
    h = mydf.handle
    ng = int(np.prod(cell.mesh))
    Z0 = pseudo.pp_nl_overlaps(cell, 0, ng, cell.mesh)
    zre, zim = pseudo.pp_nl_overlaps_device(cell, h, 0, ng, cell.mesh, batch=257)
    Z1 = zre.cpu().numpy() + 1j * zim.cpu().numpy()
    assert Z1.shape == Z0.shape
    assert abs(Z1 - Z0).max() < 1e-11 * max(1.0, abs(Z0).max())
    # ranges as ranks take them: the partial sums add up to the whole (the host form pairs +-G across range
    # boundaries, so only the totals are comparable)
    cuts = [0, ng // 3, 2 * ng // 3 + 5, ng]
    Zs = 0
    for g0, g1 in zip(cuts[:-1], cuts[1:]):
        zre, zim = pseudo.pp_nl_overlaps_device(cell, h, g0, g1, cell.mesh)
        Zs = Zs + zre.cpu().numpy() + 1j * zim.cpu().numpy()
    assert abs(Zs - Z0).max() < 1e-11 * max(1.0, abs(Z0).max())
    vnl0 = _ref_nl(cell)
    assert abs(mydf.get_pp_nl() - vnl0).max() < 1e-10
    assert abs(mydf.get_pp_nl(on_device=False) - vnl0).max() < 1e-10
