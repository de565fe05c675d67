"""Oracle: exact-grid (FFTDF) J and K at the Gamma point (test infrastructure).

Restates pyscf/pbc/df/fft_jk.py:33-109 (get_j_kpts) and :177-302 (get_k_kpts)
for a single real k-point, plus the exxdiv='ewald' G=0 correction of
pyscf/pbc/df/df_jk.py:1446-1455.  ISDF must converge to these as the
interpolation points approach the full grid.

``ao`` is the [ngrids, nao] real AO matrix on the uniform grid.
"""
import numpy as np
from .pbc_tools import fft, ifft, get_coulG


def _fmt_dms(dm, nao):
    dm = np.asarray(dm)
    dms = dm.reshape(-1, nao, nao)
    if dms.dtype not in (np.double, np.complex128):     # df_jk.py:1415-1416
        dms = np.asarray(dms, dtype=np.double)
    return dm, dms


def fftdf_get_j(ao, dm, a, mesh):
    ngrids, nao = ao.shape
    dm0, dms = _fmt_dms(dm, nao)
    vol = abs(np.linalg.det(a))
    coulG = get_coulG(a, mesh)
    vj = np.zeros((len(dms), nao, nao))
    for i, d in enumerate(dms):
        rhoR = np.einsum('gi,ij,gj->g', ao, d, ao, optimize=True)
        vR = ifft(coulG * fft(rhoR, mesh), mesh).real * (vol / ngrids)
        vj[i] = ao.T.dot(ao * vR[:, None])
    return vj.reshape(dm0.shape)


def fftdf_get_k(ao, dm, a, mesh, exxdiv=None, ovlp=None, madelung=None, blksize=8):
    ngrids, nao = ao.shape
    dm0, dms = _fmt_dms(dm, nao)
    vol = abs(np.linalg.det(a))
    weight = vol / ngrids
    if exxdiv not in (None, 'ewald'):
        raise NotImplementedError('exxdiv=%s' % exxdiv)
    coulG = get_coulG(a, mesh)                 # G=0 -> 0 for both None and ewald (fft_jk.py:267-268)
    aoT = np.ascontiguousarray(ao.T)
    vk = np.zeros((len(dms), nao, nao))
    ao_dms = [d.dot(aoT) for d in dms]         # [nao, ngrids]
    vR_dm = np.empty((len(dms), nao, ngrids))
    for p0 in range(0, nao, blksize):
        p1 = min(nao, p0 + blksize)
        rho1 = np.einsum('ig,jg->ijg', aoT[p0:p1], aoT)
        vG = fft(rho1.reshape(-1, ngrids), mesh) * coulG
        vR = ifft(vG, mesh).real.reshape(p1 - p0, nao, ngrids)
        for i in range(len(dms)):
            vR_dm[i, p0:p1] = np.einsum('ijg,jg->ig', vR, ao_dms[i])
    for i in range(len(dms)):
        vk[i] = weight * vR_dm[i].dot(ao)
        if exxdiv == 'ewald':
            vk[i] += madelung * ovlp.dot(dms[i]).dot(ovlp)
    return vk.reshape(dm0.shape)


def fftdf_get_jk(ao, dm, a, mesh, with_j=True, with_k=True, exxdiv=None,
                 ovlp=None, madelung=None):
    vj = fftdf_get_j(ao, dm, a, mesh) if with_j else None
    vk = fftdf_get_k(ao, dm, a, mesh, exxdiv, ovlp, madelung) if with_k else None
    return vj, vk


# ---- k-point sampling (complex Bloch AOs): pyscf/pbc/df/fft_jk.py:33-109 and :177-302 in full ----------------
def fftdf_get_j_kpts(ao_kpts, dm_kpts, a, mesh, ao_band=None):
    """J at sampled k-points.  ao_kpts[k] is the [ngrids, nao] complex Bloch AO matrix at kpts[k]
    (phi_k = sum_T exp(i k.T) phi(r - T), pyscf/pbc/gto/eval_gto.py:94-165), dm_kpts [nkpts, nao, nao].
    rho = 1/nk sum_k sum_{mu nu} phi_k,mu D_k[mu,nu] conj(phi_k,nu)  (fft_jk.py:77-83; real for hermitian D),
    vj_k = phi_k^H diag(vR) phi_k (fft_jk.py:99-107).  ao_band: Bloch AOs of the band k-points (default: ao_kpts)."""
    nk = len(ao_kpts)
    ngrids, nao = ao_kpts[0].shape
    vol = abs(np.linalg.det(a))
    dms = np.asarray(dm_kpts).reshape(nk, nao, nao)
    rhoR = np.zeros(ngrids, dtype=np.complex128)
    for k in range(nk):
        rhoR += np.einsum('xi,xi->x', ao_kpts[k].dot(dms[k]), ao_kpts[k].conj())
    rhoR /= nk
    vR = ifft(get_coulG(a, mesh) * fft(rhoR, mesh), mesh) * (vol / ngrids)
    band = ao_kpts if ao_band is None else ao_band
    return np.stack([ao.conj().T.dot(ao * vR[:, None]) for ao in band])


def fftdf_get_k_kpts(ao_kpts, dm_kpts, a, mesh, kpts, coords, ao_band=None, kpts_band=None, mo=None):
    """K at sampled k-points, statement for statement pyscf/pbc/df/fft_jk.py:226-292 (exxdiv=None: the G=0 term
    of coulG(k2 - k1) is dropped where k1 = k2).  mo = (mo_coeff[nk, nao, nmo], mo_occ[nk, nmo]) takes the
    occupied-orbital branch of :235-238.  kpts_band / ao_band: "band" k-points at which K is evaluated."""
    nk = len(ao_kpts)
    ngrids, nao = ao_kpts[0].shape
    vol = abs(np.linalg.det(a))
    kpts = np.reshape(kpts, (-1, 3))
    dms = np.asarray(dm_kpts).reshape(nk, nao, nao)
    weight = 1.0 / nk * (vol / ngrids)
    ao2 = [np.ascontiguousarray(ao.T) for ao in ao_kpts]
    if ao_band is None:
        ao1, kpts_band = ao2, kpts
    else:
        ao1, kpts_band = [np.ascontiguousarray(ao.T) for ao in ao_band], np.reshape(kpts_band, (-1, 3))
    if mo is not None:
        mo_coeff, mo_occ = mo
        cs = [mo_coeff[k][:, occ > 0] * np.sqrt(occ[occ > 0]) for k, occ in enumerate(mo_occ)]
        ao2 = [cs[k].T.dot(ao) for k, ao in enumerate(ao2)]
    vk = np.zeros((len(ao1), nao, nao), dtype=np.complex128)
    for k2, ao2T in enumerate(ao2):
        ao_dm = ao2T.conj() if mo is not None else dms[k2].dot(ao2T.conj())
        for k1, ao1T in enumerate(ao1):
            dk = kpts[k2] - kpts_band[k1]
            coulG = get_coulG(a, mesh, dk)
            expmikr = np.exp(-1j * coords.dot(dk))
            rho1 = np.einsum('ig,jg->ijg', ao1T.conj() * expmikr, ao2T)
            vR = ifft(fft(rho1.reshape(-1, ngrids), mesh) * coulG, mesh).reshape(nao, ao2T.shape[0], ngrids)
            vR_dm = np.einsum('ijg,jg->ig', vR, ao_dm) * expmikr.conj()
            vk[k1] += weight * vR_dm.dot(ao1T.T)
    return vk
