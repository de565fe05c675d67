"""Oracle: FFT conventions and Coulomb kernel (test infrastructure only).

Restates pyscf/pbc/tools/pbc.py:149-211 (fft / ifft: batched 3-D DFT over the
last three axes of ``f.reshape(-1, *mesh)``, forward unnormalised, inverse
1/N; the reference's own test pyscf/pbc/tools/test/test_pbc.py:185-217 pins
both to numpy.fft) and pyscf/pbc/tools/pbc.py:230-420 (get_coulG, 3-D branch).
"""
import numpy as np


def fp(a):
    """Fingerprint used by the reference's known-answer tests
    (pyscf/lib/misc.py:1150-1154)."""
    a = np.asarray(a)
    return np.dot(np.cos(np.arange(a.size)), a.ravel())


# Threads for the batched transforms: 1 = numpy.fft (what pyscf/pbc/tools/pbc.py:114-118 calls);
# bench.py sets it to the host's core count for the timed CPU baseline, which switches to
# scipy.fft (the same pocketfft algorithm, run on `workers` threads over the batch).
FFT_WORKERS = 1


def _fftn(x):
    if FFT_WORKERS > 1:
        import scipy.fft
        return scipy.fft.fftn(x, axes=(1, 2, 3), workers=FFT_WORKERS)
    return np.fft.fftn(x, axes=(1, 2, 3))


def _ifftn(x):
    if FFT_WORKERS > 1:
        import scipy.fft
        return scipy.fft.ifftn(x, axes=(1, 2, 3), workers=FFT_WORKERS)
    return np.fft.ifftn(x, axes=(1, 2, 3))


def fft(f, mesh):
    f = np.asarray(f)
    if f.size == 0:
        return np.zeros_like(f)
    f3d = f.reshape(-1, *mesh)
    g3d = _fftn(f3d)
    ngrids = int(np.prod(mesh))
    return g3d.ravel() if f.ndim == 1 else g3d.reshape(-1, ngrids)


def ifft(g, mesh):
    g = np.asarray(g)
    if g.size == 0:
        return np.zeros_like(g)
    g3d = g.reshape(-1, *mesh)
    f3d = _ifftn(g3d)
    ngrids = int(np.prod(mesh))
    return f3d.ravel() if g.ndim == 1 else f3d.reshape(-1, ngrids)


def get_Gv(b, mesh):
    """G vectors in FFT index order (pyscf/pbc/gto/cell.py:523-596)."""
    base = [np.fft.fftfreq(int(n), 1. / int(n)) for n in mesh]
    gx, gy, gz = np.meshgrid(*base, indexing='ij')
    return np.stack([gx.ravel(), gy.ravel(), gz.ravel()], axis=1).dot(b)


def get_coulG(a, mesh, k=np.zeros(3), exx=False, wrap_around=True,
              madelung=None, nkpts=1):
    """4 pi / |k+G|^2 with G=0 -> 0 (pyscf/pbc/tools/pbc.py:269-356);
    ``exx='ewald'`` adds Nk*vol*madelung at G=0 (:397-398); k+G wrap-around
    as in :279-302."""
    a = np.asarray(a, dtype=float)
    b = 2 * np.pi * np.linalg.inv(a.T)
    vol = abs(np.linalg.det(a))
    Gv = get_Gv(b, mesh)
    k = np.asarray(k, dtype=float)
    kG = k + Gv if abs(k).sum() > 1e-9 else Gv.copy()
    on_boundary = None
    if wrap_around and abs(k).sum() > 1e-9:
        on_boundary = np.zeros(len(Gv), dtype=bool)
        box_edge = np.einsum('i,ij->ij', np.asarray(mesh) // 2 + 0.5, b)
        red = np.linalg.solve(box_edge.T, kG.T).T.round(9)
        on_edge = red.astype(int)
        for ax in range(3):
            on_boundary |= red[:, ax] == 1
            on_boundary |= red[:, ax] == -1
            kG[on_edge[:, ax] == 1] -= 2 * box_edge[ax]
            kG[on_edge[:, ax] == -1] += 2 * box_edge[ax]
    absG2 = np.einsum('gi,gi->g', kG, kG)
    G0 = np.where(absG2 == 0)[0]
    with np.errstate(divide='ignore'):
        coulG = 4 * np.pi / absG2
    coulG[G0] = 0
    if exx == 'ewald' and len(G0) > 0:
        coulG[G0] += nkpts * vol * madelung
    elif isinstance(exx, str) and exx not in ('ewald',):
        raise NotImplementedError(exx)
    if on_boundary is not None:
        coulG[on_boundary] = 0
    return coulG
