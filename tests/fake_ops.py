"""TEST DOUBLE for `_lib.Handle`: the same method surface, computed with NumPy on
torch CPU tensors.  It exists so that the multi-rank host logic of
pyscf_isdf_b200/dist.py (slab partition, candidate blocks, all-gather /
all-to-all / all-reduce sequence, quota policy) can run under the gloo backend
with world_size 2 on a machine without a GPU.  It is test infrastructure (it
imports the oracle); the product never imports it and has no CPU path.
"""
import numpy as np
import torch

import oracle

META = 4


def _np(t):
    return t.numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def _pad2(n):
    return (n + 1) & ~1


class _Timer:
    times = {}

    def mark(self, name):
        pass


class FakePlan:
    def __init__(self, nao, ngl, g_offset, ng_total, npts):
        self.nao, self.ngl, self.g0, self.ng, self.npts = nao, ngl, g_offset, ng_total, npts
        self.naop = _pad2(nao)
        self.tol = 0.0
        self.ip, self.resid = [], []

    def close(self):
        pass

    def begin(self, phi, ldphi):
        p = _np(phi)[:, :self.ngl]
        self.d = np.einsum('ig,ig->g', p, p) ** 2
        return float(self.d.max())

    def set_tol(self, tol):
        self.tol = tol

    def rl(self, k0):
        return META + self.naop + _pad2(k0)

    def candidates(self, phi, ldphi, Lt, ldl, k0, quota, block):
        rl = self.rl(k0)
        B = _np(block)[:(quota + 1) * rl].reshape(quota + 1, rl)
        B[:] = 0.0
        B[:, :2] = -1.0
        pos = np.nonzero(self.d > 0)[0]
        order = pos[np.lexsort((pos, -self.d[pos]))]          # value descending, index ascending
        sel = np.sort(order[:quota])
        rest = order[quota:]
        theta = float(self.d[rest].max()) if len(rest) else 0.0
        n = len(sel)
        B[:n, 0] = sel + self.g0
        B[:n, 1] = self.d[sel]
        B[:n, META:META + self.nao] = _np(phi)[:, sel].T
        if k0:
            B[:n, META + self.naop:META + self.naop + k0] = _np(Lt)[:k0, sel].T
        B[quota, 2] = n
        B[quota, 3] = theta

    def panel(self, blocks, nranks, quota, k0, nbmax, U, ldu):
        rl = self.rl(k0)
        m = nranks * (quota + 1)
        B = _np(blocks)[:m * rl].reshape(m, rl)
        gidx = B[:, 0].astype(np.int64)
        dC = B[:, 1].copy()
        Phi = B[:, META:META + self.nao]
        L = B[:, META + self.naop:META + self.naop + k0]
        theta = max(B[r * (quota + 1) + quota, 3] for r in range(nranks))
        T = Phi.dot(Phi.T) ** 2 - L.dot(L.T)
        Lpp, piv, stop = [], [], False
        for kk in range(nbmax):
            best = dC.max()
            cands = np.nonzero(dC == best)[0]
            sc = cands[np.argmin(gidx[cands])]
            if gidx[sc] < 0 or not best > self.tol:
                stop = True
                break
            if kk > 0 and not best > theta:
                break
            col = T[sc].copy()
            for l in Lpp:
                col -= l * l[sc]
            col /= np.sqrt(best)
            Lpp.append(col)
            dC -= col * col
            dC[sc] = -1.0
            piv.append(sc)
            self.resid.append(best)
            self.ip.append(int(gidx[sc]))
        nb = len(piv)
        self.PivRows = B[piv].copy()
        # LppP[k][j] = L value of pivot k's point in within-panel column j (j <= k)
        self.LppP = np.array([[Lpp[j][piv[k]] for j in range(nb)] for k in range(nb)]).reshape(nb, nb)
        if U is not None and nb:
            Un = _np(U)
            Un[:, k0:k0 + nb] = 0.0
            Un[:k0, k0:k0 + nb] = self.PivRows[:, META + self.naop:META + self.naop + k0].T
            for k in range(nb):
                for j in range(k + 1):
                    Un[k0 + j, k0 + k] = self.LppP[k, j]
        return nb, stop

    def flush(self, phi, ldphi, Lt, ldl, k0, nb):
        if nb == 0:
            return
        p = _np(phi)[:, :self.ngl]
        Ln = _np(Lt)
        PivPhi = self.PivRows[:, META:META + self.nao]
        PivL = self.PivRows[:, META + self.naop:META + self.naop + k0]
        new = PivPhi.dot(p) ** 2
        if k0:
            new -= PivL.dot(Ln[:k0, :self.ngl])
        for k in range(nb):
            for j in range(k):
                new[k] -= self.LppP[k, j] * new[j]
            new[k] /= np.sqrt(self.resid[k0 + k])
            self.d -= new[k] ** 2
        Ln[k0:k0 + nb, :self.ngl] = new
        for g in self.ip[k0:k0 + nb]:
            if self.g0 <= g < self.g0 + self.ngl:
                self.d[g - self.g0] = -1.0

    def finish(self, rank):
        return np.array(self.ip[:rank], dtype=np.int64), np.array(self.resid[:rank])


class FakeOps:
    """NumPy stand-in for _lib.Handle (CPU tensors)."""

    def empty(self, *shape, dtype=None):
        return torch.zeros(*shape, dtype=dtype or torch.float64)

    zeros = empty

    def empty_complex(self, *shape):
        return torch.zeros(*shape, dtype=torch.complex128)

    @staticmethod
    def as_real(t):
        v = torch.view_as_real(t)
        return v.reshape(*t.shape[:-1], 2 * t.shape[-1])

    @staticmethod
    def as_complex(t):
        return torch.view_as_complex(t.view(t.shape[0], t.shape[1] // 2, 2))

    def from_host(self, arr):
        return torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64))

    def j_from_potential(self, phi, ldphi, nao, ngl, vR, ldv, nset, weight, vj, ws=None):
        P = _np(phi)[:nao, :ngl]
        for i in range(nset):
            vj[i] = torch.from_numpy(weight * (P * _np(vR)[i, :ngl]).dot(P.T))

    def phase_expand(self, tab_re, tab_im, ldt, nrow, Gv, ng, R, sign, out_re, out_im, ldo):
        """isdf_phase_expand: out = tab * exp(sign i G.R) as real / imaginary planes."""
        t = _np(tab_re)[:nrow, :ng] + (1j * _np(tab_im)[:nrow, :ng] if tab_im is not None else 0.0)
        ph = np.exp(1j * sign * _np(Gv)[:ng].dot(np.asarray(R, dtype=float)))
        z = t * ph[None, :]
        out_re[:nrow, :ng] = torch.from_numpy(np.ascontiguousarray(z.real))
        out_im[:nrow, :ng] = torch.from_numpy(np.ascontiguousarray(z.imag))

    def transpose(self, src, ldin, rows, cols, out, ldout):
        """isdf_transpose: out[cols, rows] = src[rows, cols]^T."""
        a = _np(src).reshape(-1)[:rows * ldin].reshape(rows, ldin)[:, :cols]
        o = _np(out).reshape(-1)
        o[:cols * ldout].reshape(cols, ldout)[:, :rows] = a.T

    def index(self, idx):
        return torch.as_tensor(np.asarray(idx, dtype=np.int64))

    def synchronize(self):
        pass

    def timer(self):
        return _Timer()

    def workspace(self, n):
        return torch.zeros(1)

    # -- stage 1 ------------------------------------------------------------------
    def eval_ao(self, atm, bas, env, a, mesh, Ls, rcut, g0, g1, phi, ldphi):
        from pyscf_isdf_b200.cell import cartesian_prod
        mesh = [int(m) for m in mesh]
        frac = cartesian_prod([np.fft.fftfreq(n) for n in mesh])
        coords = frac.dot(np.asarray(a))
        # the oracle screens images per 56-point block counted from grid point 0
        # (pyscf/lib/gto/grid_ao_drv.h:35), so evaluate from the enclosing block boundary
        b0 = g0 - g0 % 56
        b1 = min(len(coords), (g1 + 55) // 56 * 56)
        ao = oracle.eval_ao(atm, bas, env, coords[b0:b1], Ls, rcut)
        _np(phi)[:, :g1 - g0] = ao[g0 - b0:g1 - b0].T

    def eval_ao_deriv1(self, atm, bas, env, a, mesh, Ls, rcut, g0, g1, phi4, ldphi, comp_stride):
        from pyscf_isdf_b200.cell import cartesian_prod
        mesh = [int(m) for m in mesh]
        coords = cartesian_prod([np.fft.fftfreq(n) for n in mesh]).dot(np.asarray(a))
        b0 = g0 - g0 % 56
        b1 = min(len(coords), (g1 + 55) // 56 * 56)
        ao = oracle.eval_ao(atm, bas, env, coords[b0:b1], Ls, rcut, deriv=1)         # [4, n, nao]
        _np(phi4)[:, :, :g1 - g0] = ao[:, g0 - b0:g1 - b0].transpose(0, 2, 1)

    # -- exchange-correlation (NumPy restatement of what csrc/xc.cu computes) ---------
    def xc_code(self, name):
        if str(name).lower() not in ('lda,', 'blyp', 'pbe', 'lda,vwn', 'pbe0', 'b3lyp'):
            raise NotImplementedError('xc functional %r' % (name,))
        return str(name).lower()

    def xc_hybrid_fraction(self, name):
        from oracle import xc
        return xc.HYBRID.get(self.xc_code(name), 0.0)

    def xc_density(self, phi4, ldphi, comp_stride, nao, ngl, dm, rho4, ldr, ws=None):
        P = _np(phi4)[:, :nao, :ngl]
        T = _np(dm).dot(P[0])
        r = _np(rho4)
        for c in range(4):
            r[c, :ngl] = (1.0 if c == 0 else 2.0) * np.einsum('ig,ig->g', P[c], T)
        return ws

    def xc_potential(self, code, rho4, ldr, ngl, weight, wv4, sums, ws):
        from oracle import xc
        r = _np(rho4)[:, :ngl]
        exc, vrho, vsigma = xc.eval_xc(code, r[0], np.einsum('xg,xg->g', r[1:], r[1:]))
        w = _np(wv4)
        w[0, :ngl] = 0.5 * weight * vrho
        w[1:, :ngl] = 2 * weight * vsigma * r[1:]
        _np(sums)[:] = [(r[0] * weight).sum(), (r[0] * weight).dot(exc)]

    def xc_matrix(self, phi4, ldphi, comp_stride, nao, ngl, wv4, ldr, vmat, ws):
        P = _np(phi4)[:, :nao, :ngl]
        aow = np.einsum('cig,cg->ig', P, _np(wv4)[:, :ngl])
        m = P[0].dot(aow.T)
        _np(vmat)[:] = (m + m.T).reshape(-1)

    # -- stage 2 / 3 ----------------------------------------------------------------
    def select_create(self, nao, ngl, g_offset, ng_total, npts, ws=None):
        return FakePlan(nao, ngl, g_offset, ng_total, npts)

    def fit_aux_from_U(self, U, ldu, npts, Z, ldz, ngl):
        import scipy.linalg
        Un = np.triu(_np(U)[:npts, :npts])
        Zn = _np(Z)
        Zn[:npts, :ngl] = scipy.linalg.solve_triangular(Un, Zn[:npts, :ngl], lower=False)

    def gather_cols(self, src, ldin, nrows, idx, out, ldout):
        _np(out)[:nrows, :len(idx)] = _np(src)[:nrows, np.asarray(idx)]

    # -- stage 4 ------------------------------------------------------------------
    def fft_zy(self, vin, ldin, nvec, nxl, mesh, spec, lds):
        nx, ny, nz = (int(m) for m in mesh)
        nzh = nz // 2 + 1
        v = _np(vin)[:nvec, :nxl * ny * nz].reshape(nvec, nxl, ny, nz)
        s = np.fft.fft(np.fft.rfft(v, axis=3), axis=2).reshape(nvec, nxl * ny * nzh)
        S = _np(spec).reshape(-1)
        for i in range(nvec):
            S[i * lds:i * lds + nxl * ny * nzh] = s[i]

    def fft_pack(self, spec, lds, nvec, nxl, ncols, col_off, send):
        S = _np(spec).reshape(-1)
        out = _np(send).reshape(-1)
        src = np.stack([S[i * lds:i * lds + nxl * ncols] for i in range(nvec)]).reshape(nvec, nxl, ncols)
        for r in range(len(col_off) - 1):
            c0, c1 = int(col_off[r]), int(col_off[r + 1])
            blk = src[:, :, c0:c1].reshape(-1)
            out[nvec * nxl * c0:nvec * nxl * c0 + blk.size] = blk

    def fft_unpack(self, recv, nvec, x_off, ncl, Y, ldy):
        R = _np(recv).reshape(-1)
        Yn = _np(Y)
        for r in range(len(x_off) - 1):
            x0, x1 = int(x_off[r]), int(x_off[r + 1])
            blk = R[nvec * ncl * x0:nvec * ncl * x1].reshape(nvec, x1 - x0, ncl)
            for i in range(nvec):
                Yn[i, x0 * ncl:x1 * ncl] = blk[i].reshape(-1)

    def fft_x_coulomb(self, Y, ldy, nvec, mesh, a, col0, ncl):
        nx, ny, nz = (int(m) for m in mesh)
        nzh = nz // 2 + 1
        coulG = oracle.get_coulG(np.asarray(a), [nx, ny, nz]).reshape(nx, ny, nz)[:, :, :nzh]
        w = np.full(nzh, 2.0)
        w[0] = 1.0
        if nz % 2 == 0:
            w[-1] = 1.0
        f = np.sqrt(coulG * w).reshape(nx, ny * nzh)[:, col0:col0 + ncl]
        Yn = _np(Y)
        y = Yn[:nvec, :nx * ncl].reshape(nvec, nx, ncl)
        Yn[:nvec, :nx * ncl] = (np.fft.fft(y, axis=1) * f).reshape(nvec, nx * ncl)

    def dgemm(self, M, N, K, A, lda, a_kmajor, B, ldb, b_kmajor, C, ldc,
              Cin=None, ldcin=0, E=None, lde=0, alpha=1.0, neg=0, epi=0, upper_only=0):
        An, Bn = _np(A), _np(B)
        At = An[:K, :M].T if a_kmajor else An[:M, :K]
        Bt = Bn[:K, :N] if b_kmajor else Bn[:N, :K].T
        acc = At.dot(Bt)
        if neg:
            acc = -acc
        if Cin is not None:
            acc = acc + _np(Cin)[:M, :N]
        acc = acc * acc if epi == 1 else (acc * _np(E)[:M, :N] if epi == 2 else alpha * acc)
        _np(C)[:M, :N] = acc

    def syrk_i8(self, n, K, A, lda, C, ldc, alpha=1.0, nslices=8, ws=None):
        An = _np(A)[:n, :K]
        _np(C)[:n, :n] = np.triu(alpha * An.dot(An.T))

    def symmetrize_upper(self, W, ld, n):
        Wn = _np(W)
        Wn[:n, :n] = np.triu(Wn[:n, :n]) + np.triu(Wn[:n, :n], 1).T

    # -- stage 5 ------------------------------------------------------------------
    def get_jk_block(self, X, Wrows, ldw, p0, npl, dm, nset, nao, npts, with_j, with_k, vj, vk, ws=None):
        Xn, Wb = _np(X), _np(Wrows)[:npl, :npts]
        Xb = Xn[:, p0:p0 + npl]
        for i in range(nset):
            T = _np(dm)[i].dot(Xn)
            if with_j:
                rho = np.einsum('mp,mp->p', Xn, T)
                _np(vj)[i] = (Xb * Wb.dot(rho)).dot(Xb.T)
            if with_k:
                _np(vk)[i] = Xb.dot((Xb.T.dot(T) * Wb).dot(Xn.T))
        return ws
