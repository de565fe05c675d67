"""TEST INFRASTRUCTURE: a plain Gamma-point restricted Hartree-Fock loop around a ``get_jk`` callable,
so that SCF energies -- the last item of the parity bar -- can be compared with the reference's own
fixed numbers.  The SCF driver is the CALLER of the ISDF plugin (pyscf/pbc/scf/hf.py in the reference)
and is not part of the product; this stands in for it in tests only: Roothaan iterations with DIIS on
the commutator, energy  E = 1/2 Tr[D (h + F)] + E_nuc  (pyscf/scf/hf.py:energy_elec)."""
import numpy as np
import scipy.linalg


def rhf_gamma(hcore, ovlp, nelec, enuc, get_jk, conv_tol=1e-11, max_cycle=60):
    """get_jk(dm) -> (vj, vk); returns (e_tot, dm, ncycle)."""
    nocc = nelec // 2

    def density(f):
        e, c = scipy.linalg.eigh(f, ovlp)
        co = c[:, :nocc]
        return 2 * co.dot(co.T)

    dm = density(hcore)
    e_last = None
    errs, focks = [], []
    for it in range(max_cycle):
        vj, vk = get_jk(dm)
        f = hcore + vj - 0.5 * vk
        e_tot = 0.5 * np.einsum('ij,ji->', dm, hcore + f) + enuc
        err = f.dot(dm).dot(ovlp)
        err = err - err.T
        focks.append(f)
        errs.append(err)
        focks, errs = focks[-8:], errs[-8:]
        if e_last is not None and abs(e_tot - e_last) < conv_tol and abs(err).max() < 1e-7:
            return e_tot, dm, it + 1
        e_last = e_tot
        n = len(errs)
        if n > 1:
            B = -np.ones((n + 1, n + 1))
            B[n, n] = 0
            for i in range(n):
                for j in range(n):
                    B[i, j] = np.vdot(errs[i], errs[j])
            rhs = np.zeros(n + 1)
            rhs[n] = -1
            try:
                c = np.linalg.solve(B, rhs)[:n]
                f = sum(ci * fi for ci, fi in zip(c, focks))
            except np.linalg.LinAlgError:
                pass
        dm = density(f)
    raise RuntimeError('SCF not converged')


def rks_gamma(hcore, ovlp, nelec, enuc, get_j, nr_rks, conv_tol=1e-10, max_cycle=80, get_k=None, hyb=0.0):
    """Restricted Kohn-Sham: get_j(dm) -> vj, nr_rks(dm) -> (nelec, exc, vxc) of the semilocal part; for a
    hybrid also get_k(dm) -> vk and its exact-exchange fraction hyb.
    E = Tr[D h] + 1/2 Tr[D J] + Exc - hyb/4 Tr[D K] + E_nuc,  F = h + J + Vxc - hyb/2 K
    (pyscf/pbc/dft/rks.py:44-125).  Returns (e_tot, dm, ncycle)."""
    nocc = nelec // 2

    def density(f):
        e, c = scipy.linalg.eigh(f, ovlp)
        co = c[:, :nocc]
        return 2 * co.dot(co.T)

    dm = density(hcore)
    e_last = None
    errs, focks = [], []
    for it in range(max_cycle):
        vj = get_j(dm)
        _, exc, vxc = nr_rks(dm)
        f = hcore + vj + vxc
        e_tot = np.einsum('ij,ji->', dm, hcore) + 0.5 * np.einsum('ij,ji->', dm, vj) + exc + enuc
        if hyb:
            vk = get_k(dm)
            f = f - 0.5 * hyb * vk
            e_tot -= 0.25 * hyb * np.einsum('ij,ji->', dm, vk)
        err = f.dot(dm).dot(ovlp)
        err = err - err.T
        focks.append(f)
        errs.append(err)
        focks, errs = focks[-8:], errs[-8:]
        if e_last is not None and abs(e_tot - e_last) < conv_tol and abs(err).max() < 1e-6:
            return e_tot, dm, it + 1
        e_last = e_tot
        n = len(errs)
        if n > 1:
            B = -np.ones((n + 1, n + 1))
            B[n, n] = 0
            for i in range(n):
                for j in range(n):
                    B[i, j] = np.vdot(errs[i], errs[j])
            rhs = np.zeros(n + 1)
            rhs[n] = -1
            try:
                c = np.linalg.solve(B, rhs)[:n]
                f = sum(ci * fi for ci, fi in zip(c, focks))
            except np.linalg.LinAlgError:
                pass
        dm = density(f)
    raise RuntimeError('SCF not converged')


def krks_kmesh(hcore_k, ovlp_k, nelec_cell, enuc_cell, get_j, nr_rks, conv_tol=1e-10, max_cycle=40):
    """Restricted Kohn-Sham on a k-point mesh (the caller of the plugin in pyscf/pbc/dft/krks.py, pyscf/pbc/scf/khf.py:
    get_occ fills the lowest nelec/2 * nk orbitals over all k-points; energies per unit cell, k-averaged).
    hcore_k, ovlp_k: [nk, nao, nao] hermitian; get_j(dm_k) -> vj_k; nr_rks(dm_k) -> (nelec, exc, vxc_k) per cell.
    Plain Roothaan iteration with 50 % damping after the first cycle (small symmetric test systems).
    Returns (e_tot per cell, dm_k, ncycle)."""
    nk = len(hcore_k)
    nocc_total = nelec_cell * nk // 2

    def density(f_k):
        es, cs = [], []
        for k in range(nk):
            e, c = scipy.linalg.eigh(f_k[k], ovlp_k[k])
            es.append(e)
            cs.append(c)
        order = np.argsort(np.concatenate(es), kind='stable')[:nocc_total]
        nao = hcore_k.shape[1]
        dm = np.zeros_like(hcore_k, dtype=np.complex128)
        for idx in order:
            k, i = divmod(int(idx), nao)
            dm[k] += 2 * np.outer(cs[k][:, i], cs[k][:, i].conj())
        return dm

    dm = density(hcore_k)
    e_last = None
    for it in range(max_cycle):
        vj = get_j(dm)
        _, exc, vxc = nr_rks(dm)
        f = hcore_k + vj + vxc
        e_tot = (np.einsum('kij,kji->', dm, hcore_k) + 0.5 * np.einsum('kij,kji->', dm, vj)).real / nk + exc + enuc_cell
        if e_last is not None and abs(e_tot - e_last) < conv_tol:
            return e_tot, dm, it + 1
        e_last = e_tot
        new = density(f)
        dm = new if it == 0 else 0.5 * (dm + new)
    raise RuntimeError('k-mesh SCF not converged')
