"""Exact-limit error of the k-mesh route against the k-point FFTDF oracle as a function of the selection tolerance."""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import oracle
from oracle.fftdf_jk import fftdf_get_j_kpts, fftdf_get_k_kpts
from util import he_c_cell
from pyscf_isdf_b200.isdf import ISDF

cell = he_c_cell()
nao = cell.nao
kpts = cell.make_kpts([3, 1, 1])
rng = np.random.default_rng(11)
dms = rng.random((3, nao, nao)) + 1j * rng.random((3, nao, nao))
dms = dms + dms.conj().transpose(0, 2, 1)
rcut, Ls = cell.ao_eval_params()
coords = cell.get_uniform_grids()
ao_k = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords, Ls, rcut, kpts=kpts)
a, mesh = cell.lattice_vectors(), cell.mesh
vj0 = fftdf_get_j_kpts(ao_k, dms, a, mesh)
vk0 = fftdf_get_k_kpts(ao_k, dms, a, mesh, kpts, coords)
for tol in (-1.0, 1e-12, 1e-13, 1e-14, 1e-15, 1e-16):
    for fit in ('implicit', 'cholesky'):
        mydf = ISDF(cell, kpts=kpts, n_ip=60, tol=tol, fit=fit)
        vj, vk = mydf.get_jk(dms, kpts=kpts)
        print(json.dumps({'tol': tol, 'fit': fit, 'n_ip': int(mydf._sc.n_ip), 'ej': float(abs(vj - vj0).max()),
                          'ek': float(abs(vk - vk0).max())}), flush=True)
