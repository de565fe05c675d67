"""Oracle: lattice-summed AO values on a grid (test infrastructure only).

NumPy restatement of the reference's stage-1 path
  pyscf/pbc/gto/eval_gto.py:94-165      (driver, layout, Gamma -> real)
  pyscf/lib/pbc/grid_ao.c:301-429       (PBCeval_sph_iter: per 56-point block,
                                         per shell, images screened by
                                         min|r - R - L| < rcut[shell], summed
                                         with the Bloch phase)
  pyscf/lib/pbc/grid_ao.c:153-179       (_fill_grid2atm: block-minimum distance)
  pyscf/lib/gto/deriv1.c:31-58          (GTOcontract_exp0)
  pyscf/lib/gto/deriv1.c:71-165         (GTOshell_eval_grid_cart, xx xy xz yy yz zz order)
and of the two libcint v6.1.1 functions on that path whose source is not in
the tree (call sites grid_ao.c:359, :397):
  CINTcommon_fac_sp(l)   = 1/(2 sqrt(pi)), sqrt(3/(4 pi)), 1 for l = 0, 1, >=2
  CINTc2s_ket_sph1       = cartesian -> real solid harmonics, m = -l..l,
                           l <= 1 left as (x, y, z).
"""
import numpy as np

BLKSIZE = 56   # pyscf/lib/gto/grid_ao_drv.h:35

FAC_SP = {0: 0.282094791773878143, 1: 0.488602511902919921}

# rows: real spherical m=-l..l ; columns: cartesians in the order of deriv1.c
C2S = {
    2: np.array([
        # xx   xy   xz   yy   yz   zz
        [0, 1.092548430592079070, 0, 0, 0, 0],
        [0, 0, 0, 0, 1.092548430592079070, 0],
        [-0.315391565252520002, 0, 0, -0.315391565252520002, 0, 0.630783130505040012],
        [0, 0, 1.092548430592079070, 0, 0, 0],
        [0.546274215296039535, 0, 0, -0.546274215296039535, 0, 0]]),
    3: np.array([
        # xxx xxy xxz xyy xyz xzz yyy yyz yzz zzz
        [0, 1.770130769779930531, 0, 0, 0, 0, -0.590043589926643510, 0, 0, 0],
        [0, 0, 0, 0, 2.890611442640554055, 0, 0, 0, 0, 0],
        [0, -0.457045799464465739, 0, 0, 0, 0, -0.457045799464465739, 0, 1.828183197857862944, 0],
        [0, 0, -1.119528997770346170, 0, 0, 0, 0, -1.119528997770346170, 0, 0.746352665180230782],
        [-0.457045799464465739, 0, 0, -0.457045799464465739, 0, 1.828183197857862944, 0, 0, 0, 0],
        [0, 0, 1.445305721320277020, 0, 0, 0, 0, -1.445305721320277020, 0, 0],
        [0.590043589926643510, 0, 0, -1.770130769779930530, 0, 0, 0, 0, 0, 0]]),
}


def cart_powers(l):
    out = []
    for lx in range(l, -1, -1):
        for ly in range(l - lx, -1, -1):
            out.append((lx, ly, l - lx - ly))
    return out


def shell_values(l, dr, rr, alpha, coeff):
    """Values of one contracted shell at displaced points.
    dr [n,3], rr [n]; alpha [np]; coeff [nc,np] -> [nc*(2l+1), n]"""
    fac = FAC_SP.get(l, 1.0)
    eprim = np.exp(-np.outer(alpha, rr)) * fac          # [np, n]
    ectr = coeff.dot(eprim)                              # [nc, n]
    if l == 0:
        return ectr
    x, y, z = dr[:, 0], dr[:, 1], dr[:, 2]
    if l == 1:
        ang = np.stack([x, y, z])
    else:
        cart = np.stack([x**a * y**b * z**c for a, b, c in cart_powers(l)])
        if l not in C2S:
            raise NotImplementedError('l=%d' % l)
        ang = C2S[l].dot(cart)
    nc = ectr.shape[0]
    return (ectr[:, None, :] * ang[None, :, :]).reshape(nc * ang.shape[0], -1)


def shell_values_deriv1(l, dr, rr, alpha, coeff):
    """Values and Cartesian gradient of one contracted shell at displaced points
    (GTOshell_eval_grid_cart_deriv1, pyscf/lib/gto/deriv1.c:167-204, followed by the same
    cartesian -> spherical step): [4, nc*(2l+1), n], components (value, d/dx, d/dy, d/dz).
    d/dx [x^a y^b z^c e(r^2)] = a x^(a-1) y^b z^c e + x^(a+1) y^b z^c e',  e' = sum c (-2 alpha) exp(-alpha r^2)."""
    fac = FAC_SP.get(l, 1.0)
    eprim = np.exp(-np.outer(alpha, rr)) * fac
    e0 = coeff.dot(eprim)                                            # [nc, n]
    e1 = coeff.dot(eprim * (-2 * alpha)[:, None])
    xyz = [dr[:, 0], dr[:, 1], dr[:, 2]]
    pw = cart_powers(l)
    cart = np.stack([xyz[0]**a * xyz[1]**b * xyz[2]**c for a, b, c in pw])
    dcart = []
    for ax in range(3):
        rows = []
        for a, b, c in pw:
            e = [a, b, c]
            if e[ax] == 0:
                rows.append(np.zeros_like(rr))
                continue
            k = e[ax]
            e[ax] -= 1
            rows.append(k * xyz[0]**e[0] * xyz[1]**e[1] * xyz[2]**e[2])
        dcart.append(np.stack(rows))
    if l >= 2:
        if l not in C2S:
            raise NotImplementedError('l=%d' % l)
        cart = C2S[l].dot(cart)
        dcart = [C2S[l].dot(d) for d in dcart]
    nc = e0.shape[0]
    out = [e0[:, None, :] * cart[None]]
    for ax in range(3):
        out.append(e0[:, None, :] * dcart[ax][None] + e1[:, None, :] * (xyz[ax] * cart)[None])
    return np.stack(out).reshape(4, nc * cart.shape[0], -1)


def eval_ao(atm, bas, env, coords, Ls, rcut, kpts=None, blksize=BLKSIZE, deriv=0):
    """AO values, real spherical GTOs, lattice summed.

    Returns ``ao[ngrids, nao]`` float64 for the Gamma point (``kpts`` None),
    else a complex array ``[nkpts, ngrids, nao]``.  With ``deriv=1``
    (PBCGTOval_sph_deriv1, pyscf/lib/pbc/grid_ao.c:548-558) the result gains an
    axis of 4 components (value, d/dx, d/dy, d/dz) in front of ``ngrids``.
    """
    if deriv == 1:
        return _eval_ao_deriv1(atm, bas, env, coords, Ls, rcut, kpts, blksize)
    atm = np.asarray(atm); bas = np.asarray(bas); env = np.asarray(env)
    coords = np.asarray(coords, dtype=float)
    Ls = np.asarray(Ls, dtype=float).reshape(-1, 3)
    ngrids = len(coords)
    nblk = (ngrids + blksize - 1) // blksize
    npad = nblk * blksize
    cpad = np.full((npad, 3), 1e9)
    cpad[:ngrids] = coords
    dims = (bas[:, 1] * 2 + 1) * bas[:, 3]
    ao_loc = np.concatenate([[0], np.cumsum(dims)])
    nao = int(ao_loc[-1])
    gamma = kpts is None
    if gamma:
        ao = np.zeros((1, npad, nao))
        phase = np.ones((len(Ls), 1))
    else:
        kpts = np.reshape(kpts, (-1, 3))
        ao = np.zeros((len(kpts), npad, nao), dtype=np.complex128)
        phase = np.exp(1j * Ls.dot(kpts.T))
    for ia in range(len(atm)):
        shls = np.where(bas[:, 0] == ia)[0]
        if len(shls) == 0:
            continue
        ratm = env[atm[ia, 1]:atm[ia, 1] + 3]
        rcmax = rcut[shls].max()
        for m, L in enumerate(Ls):
            dr = cpad - (ratm + L)
            rr = np.einsum('gx,gx->g', dr, dr)
            mind = np.sqrt(rr.reshape(nblk, blksize).min(axis=1))
            if not (mind < rcmax).any():
                continue
            for ib in shls:
                keep = mind < rcut[ib]
                if not keep.any():
                    continue
                pts = np.repeat(keep, blksize).nonzero()[0]
                l, npr, nc = bas[ib, 1], bas[ib, 2], bas[ib, 3]
                alpha = env[bas[ib, 5]:bas[ib, 5] + npr]
                coeff = env[bas[ib, 6]:bas[ib, 6] + npr * nc].reshape(nc, npr)
                val = shell_values(l, dr[pts], rr[pts], alpha, coeff)   # [dim, n]
                i0 = ao_loc[ib]
                for k in range(ao.shape[0]):
                    ao[k, pts, i0:i0 + val.shape[0]] += val.T * phase[m, k]
    ao = ao[:, :ngrids]
    return ao[0] if gamma else ao


def _eval_ao_deriv1(atm, bas, env, coords, Ls, rcut, kpts, blksize):
    """Same driver as eval_ao (same image screening), four components per AO."""
    atm = np.asarray(atm); bas = np.asarray(bas); env = np.asarray(env)
    coords = np.asarray(coords, dtype=float)
    Ls = np.asarray(Ls, dtype=float).reshape(-1, 3)
    ngrids = len(coords)
    nblk = (ngrids + blksize - 1) // blksize
    npad = nblk * blksize
    cpad = np.full((npad, 3), 1e9)
    cpad[:ngrids] = coords
    dims = (bas[:, 1] * 2 + 1) * bas[:, 3]
    ao_loc = np.concatenate([[0], np.cumsum(dims)])
    nao = int(ao_loc[-1])
    gamma = kpts is None
    if gamma:
        ao = np.zeros((1, 4, npad, nao))
        phase = np.ones((len(Ls), 1))
    else:
        kpts = np.reshape(kpts, (-1, 3))
        ao = np.zeros((len(kpts), 4, npad, nao), dtype=np.complex128)
        phase = np.exp(1j * Ls.dot(kpts.T))
    for ia in range(len(atm)):
        shls = np.where(bas[:, 0] == ia)[0]
        if len(shls) == 0:
            continue
        ratm = env[atm[ia, 1]:atm[ia, 1] + 3]
        rcmax = rcut[shls].max()
        for m, L in enumerate(Ls):
            dr = cpad - (ratm + L)
            rr = np.einsum('gx,gx->g', dr, dr)
            mind = np.sqrt(rr.reshape(nblk, blksize).min(axis=1))
            if not (mind < rcmax).any():
                continue
            for ib in shls:
                keep = mind < rcut[ib]
                if not keep.any():
                    continue
                pts = np.repeat(keep, blksize).nonzero()[0]
                l, npr, nc = bas[ib, 1], bas[ib, 2], bas[ib, 3]
                alpha = env[bas[ib, 5]:bas[ib, 5] + npr]
                coeff = env[bas[ib, 6]:bas[ib, 6] + npr * nc].reshape(nc, npr)
                val = shell_values_deriv1(l, dr[pts], rr[pts], alpha, coeff)   # [4, dim, n]
                i0 = ao_loc[ib]
                for k in range(ao.shape[0]):
                    ao[k][:, pts, i0:i0 + val.shape[1]] += val.transpose(0, 2, 1) * phase[m, k]
    ao = ao[:, :, :ngrids]
    return ao[0] if gamma else ao
