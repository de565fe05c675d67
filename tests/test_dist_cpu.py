"""CPU: the multi-rank host logic of pyscf_isdf_b200/dist.py under gloo with
world_size 2 (and in-process with one rank), driven through the NumPy test
double tests/fake_ops.py, checked against the CPU ISDF oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def _cell():
    from util import diamond_c8
    return diamond_c8([9, 9, 9], basis='gth-szv', jitter=0.2)


def _reference(cell, npts, dm):
    import oracle
    from util import oracle_phi
    phi = oracle_phi(cell)
    ref = oracle.isdf_build(phi, cell.lattice_vectors(), cell.mesh, npts)
    vj, vk = oracle.isdf_get_jk(ref['X'], ref['W'], dm)
    return ref, vj, vk


def _run_rank(comm, out=None, fit='implicit'):
    from fake_ops import FakeOps
    from pyscf_isdf_b200 import dist
    from pyscf_isdf_b200.isdf import ISDF
    from util import rand_dm
    cell = _cell()
    mydf = ISDF(cell, c=5, fit=fit, comm=comm)
    mydf._handle = FakeOps()
    mydf.build()
    dm = rand_dm(cell.nao)
    vj, vk = dist.get_jk_device(mydf, comm, torch.from_numpy(dm[None].copy()))
    vpp = mydf.get_pp_loc()                    # hcore local part on the rank's grid slab, summed over ranks
    vne = mydf.get_nuc()
    vnl = mydf.get_pp_nl()                     # projector part: every rank sums its own range of G vectors
    xc_out = {name: mydf.nr_rks(name, dm) for name in ('pbe', 'blyp')}      # GGA exchange-correlation on the rank's slab
    # a memory budget below the slab size walks the slab in chunks of whole 56-point blocks, same result
    chunked = mydf.nr_rks('pbe', dm, max_bytes=32 * cell.nao * 200)
    assert abs(chunked[0] - xc_out['pbe'][0]) < 1e-12 * abs(chunked[0]) and abs(chunked[1] - xc_out['pbe'][1]) < 1e-12 * abs(chunked[1])
    assert abs(chunked[2] - xc_out['pbe'][2]).max() < 1e-12 * abs(chunked[2]).max()
    # the NumInt-shaped adapter: positional call of pyscf/pbc/dft/rks.py:88-89, one matrix and a set of two
    ni = mydf.numint()
    one = ni.nr_rks(cell, None, 'pbe', dm, 0, 1, np.zeros(3), None, max_memory=100)
    two = ni.nr_rks(cell, None, 'pbe', np.stack([dm, dm]), 0, 1, None, None)
    assert one[0] == xc_out['pbe'][0] and np.array_equal(one[2], xc_out['pbe'][2])
    assert two[0].shape == (2,) and two[2].shape == (2,) + dm.shape and np.array_equal(two[2][1], one[2])
    assert ni.rsh_and_hybrid_coeff('pbe0') == (0.0, 0.0, 0.25) and mydf.xc_hybrid_fraction('pbe') == 0.0
    xc_out['b3lyp'] = mydf.nr_rks('b3lyp', dm)          # semilocal part of a hybrid
    res = dict(xc=xc_out, ip=mydf.ip_index, W=mydf.W.numpy().copy(), X=mydf.aoR_ip.numpy().copy(), vpp=vpp, vne=vne, vnl=vnl,
               p0=int(mydf._part.ip_off[comm.rank]),
               vj=vj.numpy()[0].copy(), vk=vk.numpy()[0].copy(), slab=mydf._slab,
               panels=mydf.timings['select_panels'])
    if out is not None:
        out.put((comm.rank, res))
    return res


def _check(res, tol=1e-9):
    from util import rand_dm
    cell = _cell()
    ref, vj, vk = _reference(cell, 5 * cell.nao, rand_dm(cell.nao))
    assert np.array_equal(res['ip'], ref['ip'])
    assert abs(res['X'] - ref['X']).max() < 1e-12
    Wref = ref['W'][res['p0']:res['p0'] + len(res['W'])]          # this rank's row block
    assert abs(res['W'] - Wref).max() < 1e-8 * abs(ref['W']).max()
    assert abs(res['vj'] - vj).max() < tol * abs(vj).max()
    assert abs(res['vk'] - vk).max() < tol * abs(vk).max()
    # local one-electron potentials (pyscf/pbc/df/fft.py:39-84) against the per-atom restatement
    from oracle import pseudo as opp
    from util import oracle_phi
    coords = cell.atom_coords()
    charges = [float(z) for z in cell.atom_charges()]
    pps = [cell._pseudo_for(cell.atom_symbol(i)) for i in range(cell.natm)]
    ao = oracle_phi(cell).T
    a, mesh = cell.lattice_vectors(), cell.mesh
    vpp0 = opp.potential_matrix(ao, opp.pp_loc_on_grid(a, mesh, coords, charges, pps))
    vne0 = opp.potential_matrix(ao, opp.nuc_on_grid(a, mesh, coords, charges))
    assert abs(res['vpp'] - vpp0).max() < 1e-10
    assert abs(res['vne'] - vne0).max() < 1e-10
    # exchange-correlation (pyscf/pbc/dft/numint.py:283-381) against the oracle pinned to the reference's BLYP numbers
    import oracle
    from oracle import xc
    rcut, Ls = cell.ao_eval_params()
    ao4 = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, deriv=1)
    for name, (ne, exc, vmat) in res['xc'].items():
        ne0, exc0, vmat0 = xc.nr_rks(ao4, cell.vol / ao4.shape[1], name, rand_dm(cell.nao))
        assert abs(ne - ne0) < 1e-10 * abs(ne0) and abs(exc - exc0) < 1e-10 * abs(exc0)
        assert abs(vmat - vmat0).max() < 1e-10 * abs(vmat0).max()
    vnl0 = opp.pp_nl_matrix(a, mesh, cell._atm, cell._bas, cell._env, coords, pps).real      # fft.py:86-145
    assert abs(vnl0).max() > 1e-3 and abs(res['vnl'] - vnl0).max() < 1e-12


def test_partition():
    from pyscf_isdf_b200.dist import SlabPartition
    p = SlabPartition([71, 71, 71], 8, 12480)
    assert p.x_off[0] == 0 and p.x_off[-1] == 71 and (np.diff(p.x_off) >= 8).all()
    assert p.col_off[-1] == 71 * 36 and p.ip_off[-1] == 12480
    assert all(o % 8 == 0 for o in p.ip_off[:-1])
    assert p.grid_range(0)[0] == 0 and p.grid_range(7)[1] == 71 ** 3
    assert sum(p.nx_local(r) for r in range(8)) == 71
    p = SlabPartition([9, 9, 9], 3, 50)           # ragged: 50 points over 3 ranks
    assert list(p.ip_off) == [0, 16, 32, 50]
    with pytest.raises(ValueError):
        SlabPartition([4, 4, 4], 8, 10)


@pytest.mark.parametrize('fit', ['implicit', 'pivot_factor'])
def test_single_rank_driver_matches_oracle(fit):
    from pyscf_isdf_b200.dist import SerialComm
    _check(_run_rank(SerialComm(), fit=fit))


def test_single_rank_driver_with_aliased_build_buffers(monkeypatch):
    """ISDF_ALIAS_ARENAS=1: the spectra overwrite the factor rows as they are consumed (what a build takes when
    two arenas do not fit the device); same result."""
    from pyscf_isdf_b200.dist import SerialComm
    monkeypatch.setenv('ISDF_ALIAS_ARENAS', '1')
    _check(_run_rank(SerialComm(), fit='implicit'))


def _worker(rank, world, port, q, alias=False):
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    if alias:
        os.environ['ISDF_ALIAS_ARENAS'] = '1'
    # several small all-to-all batches: their count must not depend on the rank's slab size
    os.environ['ISDF_A2A_BATCH_BYTES'] = '20000'
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from pyscf_isdf_b200.dist import Comm
        _run_rank(Comm(), q)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world,alias', [(2, False), (3, False), (2, True)])
def test_gloo_ranks_match_oracle(world, alias):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, alias)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        r, res = q.get(timeout=300)
        got[r] = res
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # every rank ends with the same replicated ip / X / W / J / K
    for r in range(1, world):
        assert np.array_equal(got[r]['ip'], got[0]['ip'])
        assert np.array_equal(got[r]['vk'], got[0]['vk'])
    slabs = [got[r]['slab'] for r in range(world)]
    assert slabs[0][0] == 0 and slabs[-1][1] == 9 ** 3
    assert all(slabs[r][1] == slabs[r + 1][0] for r in range(world - 1))
    _check(got[0])
