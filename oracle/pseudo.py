"""CPU oracle (test infrastructure): one-electron potentials of FFTDF --
``get_nuc`` (pyscf/pbc/df/fft.py:39-62) and ``get_pp`` (pyscf/pbc/df/fft.py:64-152),
written the way the reference writes them: dense per-atom structure factor
``SI[natm, ng]`` (pyscf/pbc/gto/cell.py:645), per-atom form factor
``vlocG[natm, ng]`` (pyscf/pbc/gto/pseudo/pp.py:58-94 on top of pp_int.py:51-71),
``-einsum('ij,ij->j', SI, vlocG)``, one inverse FFT, ``ao.T * vR @ ao``; then the
non-local projectors atom by atom and channel by channel in reciprocal space
(fft.py:86-141) against analytic Fourier transforms of the AOs.

Pinning: **pinned to the reference's own fixed numbers** -- the eight fingerprints
``lib.fp(get_nuc(kpts)[k])`` and ``lib.fp(get_pp(kpts)[k])``, k = 0..3, of
pyscf/pbc/df/test/test_fft.py:589-611 on the He/C fixture of :545-558 are
reproduced to 3e-10 (tests/test_pseudo_cpu.py), with the k-point AO values of
oracle/eval_ao.py.  That fixture has one s projector (C); the second s projector,
its off-diagonal h and the p channel (Si) are checked against a grid quadrature of
the projectors' real-space definition (same test file), and the polynomial part of
the local potential against the second formulation in the reference's tests
(``get_pp_loc_part2``, pyscf/pbc/gto/pseudo/test/test_pp.py:28-64).
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


# ---------------------------------------------------------------------------
# non-local GTH projectors (pyscf/pbc/df/fft.py:86-141)
# ---------------------------------------------------------------------------
def h_matrix(rows):
    """Symmetric h_ij from the upper-triangular rows of the GTH table
    (pyscf/gto/basis/parse_cp2k_pp.py:86-89)."""
    n = len(rows)
    h = np.zeros((n, n))
    for i, row in enumerate(rows):
        h[i, i:] = row
    return h + h.T - np.diag(h.diagonal())


def qli(x, l, i):
    """Radial prefactor polynomials of the projectors in reciprocal space
    (pyscf/pbc/gto/pseudo/pp.py:150-194, i counted from 0), l <= 2 and i <= 2."""
    s = np.sqrt
    table = {
        (0, 0): lambda x: 4 * s(2.) + 0 * x,
        (0, 1): lambda x: 8 * s(2 / 15.) * (3 - x**2),
        (0, 2): lambda x: 16 / 3. * s(2 / 105.) * (15 - 10 * x**2 + x**4),
        (1, 0): lambda x: 8 * s(1 / 3.) + 0 * x,
        (1, 1): lambda x: 16 * s(1 / 105.) * (5 - x**2),
        (1, 2): lambda x: 32 / 3. * s(1 / 1155.) * (35 - 14 * x**2 + x**4),
        (2, 0): lambda x: 8 * s(2 / 15.) + 0 * x,
        (2, 1): lambda x: 16 / 3. * s(2 / 105.) * (7 - x**2),
        (2, 2): lambda x: 32 / 3. * s(2 / 15015.) * (63 - 18 * x**2 + x**4),
    }
    return table[(l, i)](np.asarray(x, dtype=float))


def ft_ao(atm, bas, env, q):
    """Analytic Fourier transform of the AOs, ``int chi(r) exp(-i q.r) dr`` -> complex [nq, nao]
    (what pyscf/pbc/df/ft_ao.py:ft_ao returns at q = Gv + kpt; the lattice sum of a periodic AO
    leaves exactly the transform of the single Gaussian at those q).  For a real solid harmonic
    S_lm times a Gaussian,  FT[S_lm(r) exp(-a r^2)](q) = (-i)^l (pi/a)^(3/2) (2a)^(-l) S_lm(q) exp(-q^2/4a),
    so the angular code of the real-space evaluator is reused at the points q."""
    from .eval_ao import shell_values
    atm = np.asarray(atm); bas = np.asarray(bas); env = np.asarray(env)
    q = np.asarray(q, dtype=float)
    qq = np.einsum('gx,gx->g', q, q)
    dims = (bas[:, 1] * 2 + 1) * bas[:, 3]
    ao_loc = np.concatenate([[0], np.cumsum(dims)])
    out = np.zeros((len(q), int(ao_loc[-1])), dtype=np.complex128)
    for ib in range(len(bas)):
        ia, l, npr, nc = bas[ib, 0], bas[ib, 1], bas[ib, 2], bas[ib, 3]
        alpha = env[bas[ib, 5]:bas[ib, 5] + npr]
        coeff = env[bas[ib, 6]:bas[ib, 6] + npr * nc].reshape(nc, npr)
        ratm = env[atm[ia, 1]:atm[ia, 1] + 3]
        cft = coeff * ((np.pi / alpha)**1.5 * (2 * alpha)**(-float(l)))
        val = shell_values(l, q, qq, 0.25 / alpha, cft)                  # [dim, nq]
        out[:, ao_loc[ib]:ao_loc[ib + 1]] = ((-1j)**l * val * np.exp(-1j * q.dot(ratm))).T
    return out


def pp_nl_matrix(a, mesh, atm, bas, env, coords, pseudos, kpt=None):
    """Non-local part of FFTDF.get_pp at one k-point, atom by atom and channel by channel as
    pyscf/pbc/df/fft.py:99-141 writes it: projector values  rl^(l+3/2) pi^(5/4) S_lm(q) exp(-rl^2 q^2/2) qli(|q| rl)
    at q = G + k (the reference gets the first three factors from eval_gto on a fake one-primitive
    shell, fft.py:113-116), times conj(SI), contracted with ft_ao / sqrt(vol), then h_ij, then 1/vol."""
    from .eval_ao import shell_values
    a = np.asarray(a, dtype=float)
    vol = abs(np.linalg.det(a))
    b = 2 * np.pi * np.linalg.inv(a.T)
    Gv = get_Gv(b, mesh)
    kpt = np.zeros(3) if kpt is None else np.asarray(kpt, dtype=float).reshape(3)
    Gk = Gv + kpt
    G_rad = np.linalg.norm(Gk, axis=1)
    G2 = G_rad**2
    SI = get_SI(coords, Gv)
    aokG = ft_ao(atm, bas, env, Gk) * (1 / vol)**.5
    nao = aokG.shape[1]
    vppnl = np.zeros((nao, nao), dtype=np.complex128)
    for ia in range(len(coords)):
        if pseudos[ia] is None:
            continue
        for l, (rl, nl, hrows) in enumerate(pseudos[ia]['nonlocal_']):
            if nl == 0:
                continue
            part = shell_values(l, Gk, G2, np.array([.5 * rl**2]), np.array([[rl**(l + 1.5) * np.pi**1.25]]))
            pYlm = np.stack([part * qli(G_rad * rl, l, k) for k in range(nl)])     # [nl, 2l+1, ng]
            SPG_lmi = pYlm * SI[ia].conj()
            SPG_lm_aoG = np.einsum('nmg,gp->nmp', SPG_lmi, aokG)
            tmp = np.einsum('ij,jmp->imp', h_matrix(hrows), SPG_lm_aoG)
            vppnl += np.einsum('imp,imq->pq', SPG_lm_aoG.conj(), tmp)
    return vppnl * (1. / vol)
