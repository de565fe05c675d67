"""Oracle: CPU ISDF (test infrastructure only) -- PARITY UNPINNED.

The mounted reference tree has no ISDF implementation,
so these functions are this repository's own statement of the textbook
algorithm, normalised to the reference's FFTDF conventions:

  select_ip : greedy diagonally-pivoted Cholesky of the pair-density Gram
              matrix G[g,g'] = (sum_mu phi_mu(g) phi_mu(g'))^2 , using the
              pivot rule / update order of the one in-tree precedent,
              pyscf/lib/scipy_helper.py:84-105 (argmax = first maximum; column
              = (G[:,p] - L[:, :k] L[p, :k]) / sqrt(d_p); d -= L[:,k]^2).
  fit_aux   : A = (X^T X)^2, B = (X^T Phi)^2, zeta = A^-1 B by Cholesky then
              two triangular solves (pattern of pyscf/pbc/df/rsdf_builder.py:226,827).
  build_V/W : V_P = ifft(coulG * fft(zeta_P)).real ; W = (vol/Ng) zeta V^T
              (normalisation of pyscf/pbc/df/fft_jk.py:72-74, :93-94).
  isdf_get_jk : J = X diag(W rho) X^T, rho_P = (X^T D X)_PP ;
                K = X ((X^T D X) o W) X^T ; + madelung S D S for exxdiv='ewald'
                (pyscf/pbc/df/df_jk.py:1446-1455).

``aoR`` is always the [nao, ngrids] (AO-major) matrix, ``X = aoR[:, ip]``.
"""
import numpy as np
import scipy.linalg
from .pbc_tools import fft, ifft, get_coulG


def select_ip(aoR, npts, tol=-1.0, return_factor=False):
    """Greedy pivoted Cholesky on the implicit pair-density Gram matrix.

    Returns (ip[int64, rank], resid[rank]) where resid[k] is the residual
    diagonal d_p at the moment pivot k was chosen.  Selected points get
    d = -1 so they can never be picked again.
    """
    aoR = np.asarray(aoR, dtype=float)
    nao, ng = aoR.shape
    s = np.einsum('ig,ig->g', aoR, aoR)
    d = s * s
    if tol < 0:
        tol = ng * np.finfo(float).eps * d.max()
    Lt = np.zeros((npts, ng))
    ip, resid = [], []
    for k in range(npts):
        p = int(np.argmax(d))
        if d[p] <= tol:
            break
        resid.append(d[p])
        ip.append(p)
        pivot_step(aoR, Lt, d, k, p)
    ip = np.array(ip, dtype=np.int64)
    resid = np.array(resid)
    if return_factor:
        return ip, resid, Lt[:len(ip)]
    return ip, resid


def pivot_step(aoR, Lt, d, k, p):
    """Step k of the pivoted Cholesky with pivot p: new factor row, residual update
    (pyscf/lib/scipy_helper.py:97-104).  Also what bench.py times per pivot on a slab."""
    col = aoR.T.dot(aoR[:, p])
    col = col * col
    if k:
        col -= Lt[:k].T.dot(Lt[:k, p])
    piv = np.sqrt(d[p])
    Lt[k] = col / piv
    d -= Lt[k] ** 2
    d[p] = -1.0


def fit_aux(aoR, ip):
    """zeta[Np, ngrids] with A zeta = B (least-squares fit of pair densities)."""
    X = aoR[:, ip]
    A = X.T.dot(X) ** 2
    B = X.T.dot(aoR) ** 2
    c = scipy.linalg.cholesky(A, lower=True)
    y = scipy.linalg.solve_triangular(c, B, lower=True)
    return scipy.linalg.solve_triangular(c.T, y, lower=False)


def build_V(zeta, a, mesh):
    coulG = get_coulG(a, mesh)
    return ifft(fft(zeta, mesh) * coulG, mesh).real


def build_W(zeta, a, mesh, V=None):
    if V is None:
        V = build_V(zeta, a, mesh)
    vol = abs(np.linalg.det(a))
    return (vol / zeta.shape[1]) * zeta.dot(V.T)


def isdf_get_jk(X, W, dm, with_j=True, with_k=True, exxdiv=None,
                ovlp=None, madelung=None):
    nao = X.shape[0]
    dm0 = np.asarray(dm)
    dms = dm0.reshape(-1, nao, nao)
    if dms.dtype not in (np.double, np.complex128):
        dms = np.asarray(dms, dtype=np.double)
    if exxdiv not in (None, 'ewald'):
        raise NotImplementedError('exxdiv=%s' % exxdiv)
    vj = np.zeros_like(dms) if with_j else None
    vk = np.zeros_like(dms) if with_k else None
    for i, d in enumerate(dms):
        T = d.dot(X)                                   # [nao, Np]
        if with_j:
            rho = np.einsum('mp,mp->p', X, T)
            v = W.dot(rho)
            vj[i] = (X * v).dot(X.T)
        if with_k:
            M = X.T.dot(T)
            vk[i] = X.dot(M * W).dot(X.T)
            if exxdiv == 'ewald':
                vk[i] += madelung * ovlp.dot(d).dot(ovlp)
    if with_j:
        vj = vj.reshape(dm0.shape)
    if with_k:
        vk = vk.reshape(dm0.shape)
    return vj, vk


def isdf_build(aoR, a, mesh, npts, tol=-1.0):
    """Stages 2-4 on a given AO matrix -> dict(ip, X, W, zeta)."""
    ip, resid = select_ip(aoR, npts, tol)
    zeta = fit_aux(aoR, ip)
    W = build_W(zeta, a, mesh)
    return dict(ip=ip, resid=resid, X=np.ascontiguousarray(aoR[:, ip]), W=W, zeta=zeta)
