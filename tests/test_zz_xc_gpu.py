"""GPU: GGA exchange-correlation on the device ("next" row 1 of SURVEY 8f): AO gradients
(isdf_eval_ao_deriv1) and ISDF.nr_rks against the oracle (oracle/eval_ao.py deriv=1, oracle/xc.py -- both
pinned to the reference's fingerprints in pyscf/pbc/dft/test/test_numint.py, see tests/test_xc_cpu.py).
Tolerances: AO values and gradients 1e-12 abs on O(1) numbers; nelec, exc and vmat 1e-10 relative (FP64
sums over the grid in a different order).

First run on a B200 at the end of round 1: all passed (profiles/r01_optin_forms.md)."""
import os
import numpy as np
import pytest
import oracle
from oracle import xc
from util import diamond_c8, he_c_cell, rand_dm
from pyscf_isdf_b200 import _lib
from pyscf_isdf_b200.isdf import ISDF

pytestmark = pytest.mark.gpu


def _oracle_ao4(cell):
    rcut, Ls = cell.ao_eval_params()
    return oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, deriv=1)


def _cells():
    return {'he_c': he_c_cell(), 'c8': diamond_c8([15, 14, 12], basis='gth-dzvp', jitter=0.15)}


@pytest.mark.parametrize('which', ['he_c', 'c8'])
def test_ao_gradients_match_oracle_and_value_plane_matches_eval_ao(which):
    cell = _cells()[which]
    mydf = ISDF(cell, c=4)
    phi4 = mydf._ao_deriv1_on_grid().cpu().numpy()            # [4, nao, ld]
    ref = _oracle_ao4(cell).transpose(0, 2, 1)                # [4, nao, ng]
    ng = ref.shape[2]
    assert abs(phi4[:, :, :ng] - ref).max() < 1e-12
    assert (phi4[:, :, ng:] == 0).all()
    phi, ld = mydf._ao_on_grid()
    assert abs(phi.cpu().numpy()[:, :ng] - phi4[0, :, :ng]).max() < 1e-14
    # a slab that does not start on a 56-point block boundary
    h = mydf.handle
    rcut, Ls = cell.ao_eval_params()
    g0, g1 = 100, 100 + 777
    ldp = _lib.pad8(g1 - g0)
    part = h.zeros((4, cell.nao, ldp))
    h.eval_ao_deriv1(cell._atm, cell._bas, cell._env, np.ascontiguousarray(cell.lattice_vectors()),
                     np.asarray(cell.mesh, dtype=np.int32), Ls, rcut, g0, g1, part, ldp, cell.nao * ldp)
    assert abs(part.cpu().numpy()[:, :, :g1 - g0] - ref[:, :, g0:g1]).max() < 1e-12


@pytest.mark.parametrize('name', ['pbe', 'blyp', 'lda,', 'lda,vwn', 'pbe0', 'b3lyp'])
def test_nr_rks_matches_oracle(name):
    cell = _cells()['c8']
    dm = rand_dm(cell.nao) * 0.05
    ao4 = _oracle_ao4(cell)
    ne0, exc0, vmat0 = xc.nr_rks(ao4, cell.vol / ao4.shape[1], name, dm)
    ne, exc, vmat = ISDF(cell, c=4).nr_rks(name, dm)
    assert abs(ne - ne0) < 1e-10 * abs(ne0)
    assert abs(exc - exc0) < 1e-10 * abs(exc0)
    assert abs(vmat - vmat0).max() < 1e-10 * abs(vmat0).max()
    assert abs(vmat - vmat.T).max() == 0.0
    if name == 'pbe':       # a small memory budget walks the grid in chunks and re-evaluates the AO gradients
        ne2, exc2, vmat2 = ISDF(cell, c=4).nr_rks(name, dm, max_bytes=32 * cell.nao * 500)
        assert abs(ne2 - ne0) < 1e-10 * abs(ne0) and abs(exc2 - exc0) < 1e-10 * abs(exc0)
        assert abs(vmat2 - vmat0).max() < 1e-10 * abs(vmat0).max()


def test_nr_rks_errors():
    cell = _cells()['he_c']
    mydf = ISDF(cell, c=4)
    dm = rand_dm(cell.nao)
    with pytest.raises(NotImplementedError):
        mydf.nr_rks('scan', dm)
    with pytest.raises(NotImplementedError):
        mydf.nr_rks('pbe', dm, kpts=np.array([0.1, 0.0, 0.0]))
    with pytest.raises(ValueError):
        mydf.nr_rks('pbe', dm + np.triu(np.ones_like(dm), 1))


def test_nr_rks_sharded_grid():
    """Grid sharded over 3 ranks (threads on one GPU): slab results summed."""
    import threading
    from test_zzy_dist_gpu import ThreadComm
    cell = diamond_c8([13, 10, 12], basis='gth-szv', jitter=0.2)
    dm = rand_dm(cell.nao) * 0.1
    ao4 = _oracle_ao4(cell)
    ne0, exc0, vmat0 = xc.nr_rks(ao4, cell.vol / ao4.shape[1], 'pbe', dm)
    n = 3
    shared = ThreadComm.Shared(n)
    out, err = [None] * n, []

    def work(r):
        try:
            mydf = ISDF(cell, c=6, comm=ThreadComm(shared, r))
            mydf.build()
            out[r] = mydf.nr_rks('pbe', dm)
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
    for ne, exc, vmat in out:
        assert abs(ne - ne0) < 1e-10 * abs(ne0) and abs(exc - exc0) < 1e-10 * abs(exc0)
        assert abs(vmat - vmat0).max() < 1e-10 * abs(vmat0).max()
