"""Shared builders for the test-suite (deterministic synthetic cells)."""
import numpy as np
from pyscf_isdf_b200.cell import Cell

C8_FRAC = np.array([[0, 0, 0], [.25, .25, .25], [.5, .5, 0], [.75, .75, .25],
                    [.5, 0, .5], [.75, .25, .75], [0, .5, .5], [.25, .75, .75]])


def diamond_c8(mesh, basis='gth-dzvp', jitter=0.0, seed=7, a=3.5668):
    """Cubic 8-atom diamond cell of examples/pbc/10-gamma_point_scf.py:20-27;
    ``jitter`` (Angstrom) breaks the point-group symmetry."""
    rng = np.random.default_rng(seed)
    xyz = C8_FRAC * a + jitter * (rng.random((8, 3)) - .5)
    atoms = [('C', tuple(r)) for r in xyz]
    return Cell(atom=atoms, a=np.eye(3) * a, basis=basis, pseudo='gth-pade', mesh=list(mesh)).build()


def he_c_cell():
    """Fixture of pyscf/pbc/df/test/test_fft.py:545-554."""
    return Cell(atom='He 1. .5 .5; C .1 1.3 2.1',
                basis={'He': [(0, (2.5, 1)), (0, (1., 1))], 'C': 'gth-szv'},
                pseudo={'C': 'gth-pade'}, a=np.eye(3) * 2.5, mesh=[21] * 3).build()


def oracle_phi(cell):
    """[nao, ng] AO matrix from the CPU oracle."""
    import oracle
    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut)
    return np.ascontiguousarray(ao.T)


def to_dev(handle, arr, ld=None):
    """numpy [r, c] -> torch cuda [r, ld] (zero padded)"""
    import torch
    arr = np.asarray(arr, dtype=np.float64)
    r, c = arr.shape
    ld = c if ld is None else ld
    out = torch.zeros((r, ld), dtype=torch.float64, device=handle.torch_device)
    out[:, :c] = torch.from_numpy(np.ascontiguousarray(arr)).to(handle.torch_device)
    return out


def rand_dm(nao, seed=2):
    rng = np.random.default_rng(seed)
    d = rng.random((nao, nao))
    return d + d.T
