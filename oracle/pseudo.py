"""CPU oracle (test infrastructure): local one-electron potentials of FFTDF --
``get_nuc`` (pyscf/pbc/df/fft.py:39-62) and the local part of ``get_pp``
(pyscf/pbc/df/fft.py:64-84), written the way the reference writes them: dense
per-atom structure factor ``SI[natm, ng]`` (pyscf/pbc/gto/cell.py:645), per-atom
form factor ``vlocG[natm, ng]`` (pyscf/pbc/gto/pseudo/pp.py:58-94 on top of
pp_int.py:51-71), ``-einsum('ij,ij->j', SI, vlocG)``, one inverse FFT, and
``ao.T * vR @ ao``.

Pinning: the reference holds no fixed numbers for these matrices
(pyscf/pbc/df/test/test_fft.py:589-611 compares two of its own code paths), so
this is a restatement checked against the second formulation in the reference's
tests (``get_pp_loc_part2``, pyscf/pbc/gto/pseudo/test/test_pp.py:28-64, the
polynomial part alone) and against analytic limits -- **parity unpinned**.
"""
import numpy as np
from .pbc_tools import get_Gv, get_coulG, ifft


def get_SI(coords, Gv):
    return np.exp(-1j * np.dot(coords, Gv.T))


def get_gth_vlocG(a, mesh, charges, pseudos):
    """[natm, ng]; ``pseudos[ia]`` is None or a dict(rloc=, cloc=[...])."""
    a = np.asarray(a, dtype=float)
    b = 2 * np.pi * np.linalg.inv(a.T)
    Gv = get_Gv(b, mesh)
    coulG = get_coulG(a, mesh)
    G2 = np.einsum('ix,ix->i', Gv, Gv)
    G0idx = np.where(G2 == 0)[0]
    natm = len(charges)
    vlocG = np.zeros((natm, len(G2)))
    for ia in range(natm):                       # pp_int.py:61-71
        Zia = charges[ia]
        vlocG[ia] = Zia * coulG
        if pseudos[ia] is not None:
            rloc = pseudos[ia]['rloc']
            vlocG[ia] *= np.exp(-0.5 * rloc**2 * G2)
            vlocG[ia, G0idx] = -2 * np.pi * Zia * rloc**2
    for ia in range(natm):                       # pp.py:73-92
        if pseudos[ia] is None:
            continue
        rloc, cexp = pseudos[ia]['rloc'], pseudos[ia]['cloc']
        nexp = len(cexp)
        G2_red = G2 * rloc**2
        cfacs = 0
        if nexp >= 1:
            cfacs += cexp[0]
        if nexp >= 2:
            cfacs += cexp[1] * (3 - G2_red)
        if nexp >= 3:
            cfacs += cexp[2] * (15 - 10 * G2_red + G2_red**2)
        if nexp >= 4:
            cfacs += cexp[3] * (105 - 105 * G2_red + 21 * G2_red**2 - G2_red**3)
        vlocG[ia, :] -= (2 * np.pi)**(3 / 2.) * rloc**3 * np.exp(-0.5 * G2_red) * cfacs
    return vlocG


def pp_loc_on_grid(a, mesh, coords, charges, pseudos):
    b = 2 * np.pi * np.linalg.inv(np.asarray(a, dtype=float).T)
    SI = get_SI(coords, get_Gv(b, mesh))
    vpplocG = -np.einsum('ij,ij->j', SI, get_gth_vlocG(a, mesh, charges, pseudos))
    return ifft(vpplocG, mesh).real


def pp_loc_part2_on_grid(a, mesh, coords, pseudos):
    """The polynomial (C1..C4) part alone, in the form of the reference's own test helper
    (pyscf/pbc/gto/pseudo/test/test_pp.py:28-64): powers of G_red = |G| rloc, dot(cexp, cfacs)."""
    b = 2 * np.pi * np.linalg.inv(np.asarray(a, dtype=float).T)
    Gv = get_Gv(b, mesh)
    SI = get_SI(coords, Gv)
    G = np.linalg.norm(Gv, axis=1)
    vlocG = np.zeros((len(coords), len(G)))
    for ia in range(len(coords)):
        if pseudos[ia] is None:
            continue
        rloc, cexp = pseudos[ia]['rloc'], pseudos[ia]['cloc']
        G_red = G * rloc
        cfacs = np.array([1 * G_red**0, 3 - G_red**2, 15 - 10 * G_red**2 + G_red**4,
                          105 - 105 * G_red**2 + 21 * G_red**4 - G_red**6])
        vlocG[ia, :] = -(2 * np.pi)**(3 / 2.) * rloc**3 * np.exp(-0.5 * G_red**2) * np.dot(cexp, cfacs[:len(cexp)])
    return ifft(-np.sum(SI * vlocG, axis=0), mesh).real


def nuc_on_grid(a, mesh, coords, charges):
    b = 2 * np.pi * np.linalg.inv(np.asarray(a, dtype=float).T)
    SI = get_SI(coords, get_Gv(b, mesh))
    rhoG = np.dot(-np.asarray(charges, dtype=float), SI)
    return ifft(rhoG * get_coulG(a, mesh), mesh).real


def potential_matrix(ao, vR):
    """ao[ng, nao]: lib.dot(ao.T.conj()*vR, ao) (pyscf/pbc/df/fft.py:55-56, :79-81)."""
    return np.dot(ao.T * vR, ao)
