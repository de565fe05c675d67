"""Regenerates the fixtures in tests/golden/.

Two kinds of vectors live here:

* ``reference_kats.json`` -- known-answer constants copied (as numbers, with
  file:line) from the reference's own test-suite.  They pin stages 1 and 4 and
  the exact-grid J/K limit.  The reference itself cannot be imported in the
  build container (no h5py, no libcint), so these constants --
  not freshly generated reference outputs -- are the pin.
* ``isdf_small.npz`` -- outputs of THIS repository's CPU oracle (bit-exact C
  chain for the pivots, NumPy/SciPy for the rest) on a small jittered diamond
  cell.  The mounted reference has no ISDF code: PARITY UNPINNED for these.
  They let the GPU tests check pivots / W / J / K without running the oracle.

Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

KATS = {
    'coulG_random_k': {'value': 62.75448804333378, 'src': 'pyscf/pbc/tools/test/test_pbc.py:54-63'},
    'coulG_cubic_k': {'value': 4.6737453679713905, 'src': 'pyscf/pbc/tools/test/test_pbc.py:65-70'},
    'coulG_cubic_k_nowrap': {'value': 4.5757877990664744, 'src': 'pyscf/pbc/tools/test/test_pbc.py:71-72'},
    'coulG_ewald': {'value': 4.888843468914021, 'src': 'pyscf/pbc/tools/test/test_pbc.py:73-80'},
    'fftdf_vj_eye': {'value': 3.7955873127283377, 'src': 'pyscf/pbc/df/test/test_fft.py:642-644'},
    'fftdf_vk_eye': {'value': 4.290076429522121, 'src': 'pyscf/pbc/df/test/test_fft.py:642-645'},
    'eval_gto_d_shell_27k': {'value': [-0.27594803231989179, 0.0064644591759109114],
                             'src': 'pyscf/pbc/gto/test/test_cell.py:500-504'},
}


def main():
    import oracle
    from oracle import c_oracle
    from util import diamond_c8, oracle_phi, rand_dm
    with open(os.path.join(HERE, 'reference_kats.json'), 'w') as f:
        json.dump(KATS, f, indent=1, sort_keys=True)
    cell = diamond_c8([12, 12, 12], basis='gth-dzvp', jitter=0.15)
    phi = oracle_phi(cell)
    npts = 4 * cell.nao
    ip, resid = c_oracle.select_ip_chain(phi, npts)
    ref = oracle.isdf_build(phi, cell.lattice_vectors(), cell.mesh, npts)
    assert np.array_equal(ref['ip'], ip)
    dm = rand_dm(cell.nao)
    vj, vk = oracle.isdf_get_jk(ref['X'], ref['W'], dm)
    np.savez_compressed(os.path.join(HERE, 'isdf_small.npz'),
                        mesh=np.asarray(cell.mesh), jitter=0.15, c=4, ip=ip, resid=resid,
                        phi_fp=oracle.fp(phi), W_fp=oracle.fp(ref['W']), W_diag=np.diag(ref['W']).copy(),
                        W_block=ref['W'][:16, :16].copy(), vj=vj, vk=vk)
    print('wrote', os.listdir(HERE))


if __name__ == '__main__':
    main()
