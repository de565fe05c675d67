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
