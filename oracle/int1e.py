"""CPU oracle (test infrastructure): overlap and kinetic-energy matrices of the lattice-summed AOs at
the Gamma point -- what the reference's SCF gets from ``cell.pbc_intor('int1e_ovlp')`` and
``'int1e_kin'`` (pyscf/pbc/scf/hf.py:56-75, analytic lattice sums in libcint/libpbc, not on the ISDF
path).  Restated in reciprocal space: by Poisson summation the lattice sum of a two-centre integral
equals  (1/vol) sum_G conj(chi_mu(G)) chi_nu(G)  (times |G|^2/2 for the kinetic energy) over ALL
reciprocal vectors, and the analytic transforms of oracle/pseudo.py:ft_ao decay like Gaussians, so a
cube of G vectors past  exp(-G^2 / (4 alpha_max)) ~ 1e-18  gives the integrals to rounding.

Pinning: with these two matrices, oracle.pseudo.nuc_on_grid, oracle.fftdf_get_jk and Cell.ewald, a
plain Roothaan loop reproduces the reference's Gamma-point RHF energies of
pyscf/pbc/scf/test/test_hf.py:94-95 and :129-132 (tests/test_scf_cpu.py)."""
import numpy as np
from .pbc_tools import get_Gv
from .pseudo import ft_ao


def overlap_kinetic(a, atm, bas, env, tol=1e-18, chunk=20000):
    a = np.asarray(a, dtype=float)
    vol = abs(np.linalg.det(a))
    b = 2 * np.pi * np.linalg.inv(a.T)
    bas = np.asarray(bas)
    env = np.asarray(env)
    amax = max(env[bas[ib, 5]:bas[ib, 5] + bas[ib, 2]].max() for ib in range(len(bas)))
    lmax = int(bas[:, 1].max())
    gmax = np.sqrt(-2 * amax * np.log(tol)) * (1 + 0.15 * lmax)        # exp(-G^2/(2 amax)) for a product of two
    mesh = [2 * int(np.ceil(gmax / np.linalg.norm(b[i]) * 1.0)) + 1 for i in range(3)]
    # the cube must contain the sphere of radius gmax: use the heights of the reciprocal cell
    hinv = np.linalg.norm(a, axis=1) / (2 * np.pi)                     # 1 / height of b along each axis
    mesh = [2 * int(np.ceil(gmax * hinv[i])) + 1 for i in range(3)]
    Gv = get_Gv(b, mesh)
    G2 = np.einsum('gx,gx->g', Gv, Gv)
    keep = G2 <= gmax * gmax
    Gv, G2 = Gv[keep], G2[keep]
    nao = int(((bas[:, 1] * 2 + 1) * bas[:, 3]).sum())
    S = np.zeros((nao, nao), dtype=np.complex128)
    T = np.zeros((nao, nao), dtype=np.complex128)
    for i in range(0, len(Gv), chunk):
        F = ft_ao(atm, bas, env, Gv[i:i + chunk])
        S += F.conj().T.dot(F)
        T += (F.conj().T * (0.5 * G2[i:i + chunk])).dot(F)
    return S.real / vol, T.real / vol
