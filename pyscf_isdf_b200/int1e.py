"""Overlap and kinetic-energy matrices of the lattice-summed AOs at the Gamma point, on the host.

In the reference these come from libcint's analytic lattice sums (``cell.pbc_intor('int1e_ovlp')``,
``'int1e_kin'``, pyscf/pbc/scf/hf.py:56-75) and belong to the SCF driver, not to the density-fitting
plugin; they are provided here so that the package can be exercised end to end (examples/) and so that
``exxdiv='ewald'`` can use the analytic overlap where PySCF is absent.  By Poisson summation the lattice
sum of a two-centre integral is  (1/vol) sum_G conj(chi_mu(G)) chi_nu(G)  over all reciprocal vectors
(times |G|^2/2 for the kinetic energy); the closed-form transforms of pyscf_isdf_b200/pseudo.py decay like
Gaussians, so a sphere of G vectors out to  exp(-G^2/(2 alpha_max)) ~ tol  gives both to rounding.
"""
import numpy as np
from . import pseudo


def _g_sphere(cell, gmax):
    """All reciprocal lattice vectors with |G| <= gmax."""
    a = cell.lattice_vectors()
    b = cell.reciprocal_vectors()
    n = [int(np.ceil(gmax * np.linalg.norm(a[i]) / (2 * np.pi))) for i in range(3)]      # |n_i| <= gmax |a_i| / 2 pi
    idx = np.stack(np.meshgrid(*[np.arange(-m, m + 1) for m in n], indexing='ij'), axis=-1).reshape(-1, 3)
    Gv = idx.dot(b)
    G2 = np.einsum('gx,gx->g', Gv, Gv)
    keep = G2 <= gmax * gmax
    return Gv[keep], G2[keep]


def overlap_kinetic(cell, tol=1e-18, chunk=16384):
    """(S, T), real symmetric [nao, nao]."""
    bas, env = cell._bas, cell._env
    amax = max(env[bas[ib, 5]:bas[ib, 5] + bas[ib, 2]].max() for ib in range(len(bas)))
    lmax = int(bas[:, 1].max())
    gmax = np.sqrt(-2 * amax * np.log(tol)) * (1 + 0.15 * lmax)
    Gv, G2 = _g_sphere(cell, gmax)
    coords = cell.atom_coords()
    nao = cell.nao_nr()
    first = {}
    for ia in range(cell.natm):
        first.setdefault(cell.atom_symbol(ia), ia)
    S = np.zeros((nao, nao), dtype=np.complex128)
    T = np.zeros((nao, nao), dtype=np.complex128)
    for i in range(0, len(Gv), chunk):
        q, qq = Gv[i:i + chunk], G2[i:i + chunk]
        SI = np.exp(-1j * np.dot(coords, q.T))
        ft = {s: pseudo.species_ft_shells(cell, ia, q, qq) for s, ia in first.items()}
        F = np.concatenate([ft[cell.atom_symbol(ia)] * SI[ia] for ia in range(cell.natm)])       # [nao, n]
        S += F.conj().dot(F.T)
        T += (F.conj() * (0.5 * qq)).dot(F.T)
    return S.real / cell.vol, T.real / cell.vol
