"""ISDF density-fitting object: host-side mirror of the reference's ``with_df``
plugin protocol, driving the sm_100a kernels through the C-ABI.

Protocol (what ``pbc.scf.RHF/RKS`` touches on ``mf.with_df``), taken from
pyscf/pbc/df/fft.py:155-359 and pyscf/pbc/scf/hf.py:521-698:
``cell, kpts, mesh, stdout, verbose, max_memory, exxdiv, build(), reset(),
dump_flags(), get_jk(dm, hermi, kpts, kpts_band, with_j, with_k, omega,
exxdiv), update_mf(mf)``.  ``get_jk`` returns ``(vj, vk)`` with the shape of
``dm``, float64 at the Gamma point, ``None`` for the part not requested.

Stages (all on the GPU, no CPU fallback):
  build():  1 AO values -> 2 interpolation points -> 3 aux fit -> 4 Coulomb W
  get_jk(): 5 J/K contraction per SCF iteration.
The mounted reference has no ISDF class; the constructor follows SURVEY.md
section 8b (``ISDF(cell, kpts, c=15, n_ip=None, select='cholesky')``).

``fit`` selects how stages 3+4 are organised (all give the same J/K to 1e-8):
  'cholesky'      zeta = A^-1 B by Cholesky of A = (X^T X)^2 (the oracle's route)
  'pivot_factor'  zeta = U^-1 Lt from the selection factor (A = U^T U, B = U^T Lt exactly)
  'implicit'      zeta is never formed: M = Lt K Lt^T, W = U^-1 M U^-T   (default, fewest flops)
"""
import sys
import time
import numpy as np

from . import _lib
from .cell import madelung as _madelung


class ISDF:
    _keys = {'cell', 'kpts', 'mesh', 'c', 'n_ip', 'select', 'fit', 'tol', 'exxdiv',
             'ip_index', 'ip_resid', 'W', 'aoR', 'aoR_ip', 'stdout', 'verbose', 'max_memory',
             'keep_aoR', 'timings', 'comm', 'wgemm', 'exact_j', 'keep_buffers'}

    def __init__(self, cell, kpts=np.zeros((1, 3)), c=15, n_ip=None, select='cholesky',
                 fit='implicit', tol=-1.0, device=None, keep_aoR=True, comm=None, wgemm='i8',
                 exact_j=False):
        self.cell = cell
        self.stdout = sys.stdout
        self.verbose = getattr(cell, 'verbose', 0)
        self.max_memory = getattr(cell, 'max_memory', 4000)
        self.kpts = np.reshape(kpts, (-1, 3))
        self.c = c
        self.n_ip = n_ip
        self.select = select
        self.fit = fit
        self.tol = tol
        self.exxdiv = None
        self.keep_aoR = keep_aoR
        # wgemm: how the Coulomb-metric product Y Y^T is computed: 'i8' / 'i8:<planes>' (int8 tcgen05
        # split product, csrc/syrk_i8.cu; 8 digit planes = FP64-GEMM accuracy) or 'fp64' (FP64 tensor pipe)
        self.wgemm = wgemm
        # exact_j: J from the exact density on the FFT grid (the reference's FFTDF route,
        # pyscf/pbc/df/fft_jk.py:33-109) instead of the ISDF contraction; K stays ISDF.  Needs keep_aoR.
        self.exact_j = exact_j
        # keep_buffers: the two build-time work buffers (factor rows, spectra) stay with the object
        # for the next build() after reset(); release_buffers() drops them
        self.keep_buffers = True
        self._arena = None
        self._bufs = {}            # other build work buffers kept between builds (dist._kept)
        self._device = device
        self._handle = None
        # comm: a pyscf_isdf_b200.dist.Comm (one process per GPU, grid sharded by x-planes)
        self.comm = comm
        self._rsh_df = {}
        self._pinned = None        # pinned host staging buffers of get_jk, kept across reset()
        self._kmesh = None         # kmesh.KMesh when self.kpts is a uniform mesh (served by the supercell's ISDF)
        self._sc = None
        self.h2d_bytes = self.d2h_bytes = 0      # bytes the last get_jk() copied over PCIe (counted at the copies)
        self.timings = {}
        self.reset()

    # -- protocol attributes -------------------------------------------------
    @property
    def mesh(self):
        return self.cell.mesh

    @mesh.setter
    def mesh(self, x):
        self.cell.mesh = x
        self.reset()

    def reset(self, cell=None):
        """Drop everything derived from the geometry (pyscf/pbc/df/fft.py:198-203)."""
        if cell is not None:
            self.cell = cell
        self.ip_index = None
        self.ip_resid = None
        self.W = None
        self.aoR = None
        self.aoR_ip = None
        self._ld = None
        self._ovlp = None
        self._madelung = None
        self._jk_ws = None
        self._part = None
        self._slab = None
        self._aoR4 = None          # AO values + gradients of nr_rks
        # (the pinned host staging buffers of get_jk, self._pinned, depend on nao only and are re-checked there:
        # re-pinning 17 MB per geometry step costs more than the copies they serve)
        if getattr(self, '_sc', None) is not None:
            self._sc.reset()
        if cell is not None:
            self._kmesh = self._sc = None
        return self

    def release_buffers(self):
        """Give the build-time work buffers back to the allocator (they are only needed by build())."""
        self._arena = None
        self._bufs = {}
        return self

    def dump_flags(self, verbose=None):
        v = self.verbose if verbose is None else verbose
        if v >= 4:
            self.stdout.write('\n******** %s ********\n' % self.__class__)
            self.stdout.write('mesh = %s (%d PWs)\n' % (list(self.mesh), int(np.prod(self.mesh))))
            self.stdout.write('c = %s  n_ip = %s  select = %s  fit = %s\n' %
                              (self.c, self.n_ip, self.select, self.fit))
        return self

    def check_sanity(self):
        if self.select != 'cholesky':
            raise NotImplementedError("select=%r (only 'cholesky' pivoting is implemented)" % self.select)
        if self.fit not in ('implicit', 'cholesky', 'pivot_factor'):
            raise ValueError("fit must be 'implicit', 'cholesky' or 'pivot_factor'")
        if self.wgemm != 'fp64' and not str(self.wgemm).startswith('i8'):
            raise ValueError("wgemm must be 'fp64', 'i8' or 'i8:<planes>'")
        if self.comm is not None and self.fit == 'cholesky':
            raise NotImplementedError("fit='cholesky' is single-GPU only; use 'implicit' or 'pivot_factor'")
        if getattr(self.cell, 'dimension', 3) != 3:
            raise RuntimeError('ISDF supports 3-D periodic cells only')   # cf. fft.py:216-220
        if abs(self.kpts).sum() > 1e-9 and self._kmesh is None:
            from .kmesh import KMesh
            self._kmesh = KMesh(self.cell, self.kpts)     # NotImplementedError unless a complete uniform mesh
        return self

    def update_mf(self, mf):
        """Attach to an SCF object and switch off the in-core ERI shortcut that
        would bypass get_jk (pyscf/pbc/scf/hf.py:670-679, :796-802)."""
        mf = mf.copy() if hasattr(mf, 'copy') else mf
        mf.with_df = self
        mf._eri = None
        mf._is_mem_enough = lambda: False
        return mf

    @property
    def handle(self):
        if self._handle is None:
            self._handle = _lib.Handle(self._device)
        return self._handle

    # -- stages 1-4 ----------------------------------------------------------------
    def build(self):
        import torch
        self.check_sanity()
        if self._kmesh is not None:
            # k-point mesh: the Gamma-point ISDF of the supercell does all the work (kmesh.py)
            self._supercell_df()
            self._sc.build()
            self.ip_index, self.ip_resid, self.W = self._sc.ip_index, self._sc.ip_resid, self._sc.W
            self.timings = self._sc.timings
            return self
        if self.comm is not None or self.fit == 'implicit':
            from . import dist
            return dist.build(self, self.comm if self.comm is not None else dist.SerialComm())
        h = self.handle
        cell = self.cell
        mesh = np.asarray(cell.mesh, dtype=np.int32)
        ng = int(np.prod(mesh))
        nao = cell.nao_nr()
        npts = int(self.n_ip) if self.n_ip else int(self.c * nao)
        npts = min(npts, ng)
        a = np.ascontiguousarray(cell.lattice_vectors(), dtype=np.float64)
        ld = _lib.pad8(ng)
        t0 = time.perf_counter()

        # stage 1: AO values Phi[nao, ng]
        rcut, Ls = cell.ao_eval_params()
        phi = h.empty((nao, ld))
        if ld > ng:
            phi[:, ng:].zero_()
        h.eval_ao(cell._atm, cell._bas, cell._env, a, mesh, Ls, rcut, 0, ng, phi, ld)

        # stage 2: interpolation points (factor rows Lt[npts, ng] double as the zeta buffer)
        big = h.empty((npts, ld))
        ip, resid = h.select_ip(phi, nao, ng, ld, npts, self.tol, big, ld)
        rank = len(ip)
        if rank < npts and self.verbose >= 3:
            self.stdout.write('ISDF: residual reached tol after %d of %d points\n' % (rank, npts))
        npts = rank

        # stage 3: aux functions zeta[npts, ng]
        if self.fit == 'pivot_factor':
            h.fit_aux_from_factor(big, ld, ng, ip)
        else:
            h.fit_aux(phi, nao, ng, ld, ip, big, ld)

        # stage 4: Coulomb metric W[npts, npts]
        W = h.empty((npts, npts))
        h.build_W(big, npts, ld, mesh, a, W)

        X = h.empty((nao, npts))
        h.gather_cols(phi, ld, nao, ip, X, npts)
        torch.cuda.synchronize(h.torch_device)
        del big
        self.timings = {
            'ao_ms': h.stage_ms(_lib.STAGE_AO), 'select_ms': h.stage_ms(_lib.STAGE_SELECT),
            'fit_ms': h.stage_ms(_lib.STAGE_FIT), 'coulomb_ms': h.stage_ms(_lib.STAGE_COULOMB),
            'fft_ms': h.stage_ms(_lib.STAGE_FFT), 'build_wall_s': time.perf_counter() - t0}
        self.n_ip = npts
        self.ip_index = ip
        self.ip_resid = resid
        self.W = W
        self.aoR_ip = X
        self._ld = ld
        self.aoR = phi if self.keep_aoR else None
        self._jk_ws = None
        return self

    def aoR_host(self):
        """[nao, ng] AO matrix copied to the host (tests / diagnostics)."""
        if self.aoR is None:
            raise RuntimeError('aoR was not kept (keep_aoR=False)')
        ng = int(np.prod(self.mesh))
        if self._slab is not None:
            ng = self._slab[1] - self._slab[0]        # this rank's slab only
        return self.aoR[:, :ng].cpu().numpy()

    # -- overlap / Madelung for exxdiv='ewald' ---------------------------------------
    def get_ovlp(self):
        """S = (vol/ng) Phi Phi^T on the FFT grid.  (The reference takes S from
        libcint's int1e_ovlp, pyscf/pbc/df/df_jk.py:1447; libcint is not
        available here, and for GTH bases the grid quadrature is exact to the
        mesh's precision.)"""
        if self._ovlp is None:
            h = self.handle
            # the build's AO matrix when it was kept; otherwise (keep_aoR=False, or after load())
            # stage 1 is evaluated again for this one product
            phi, ld = self._ao_on_grid()
            nao = phi.shape[0]
            ng = int(np.prod(self.mesh))
            S = h.empty((nao, nao))
            # with a sharded grid every rank sums its slab, then the partial overlaps are added
            g0, g1 = self._slab if self._slab is not None else (0, ng)
            h.dgemm(nao, nao, g1 - g0, phi, ld, 0, phi, ld, 0, S, nao,
                    alpha=self.cell.vol / ng)
            if self.comm is not None:
                self.comm.all_reduce_sum(S)
            self._ovlp = S
        return self._ovlp

    def get_madelung(self):
        if self._madelung is None:
            self._madelung = _madelung(self.cell)
        return self._madelung

    # -- stage 5 -------------------------------------------------------------------------
    def get_jk(self, dm, hermi=1, kpts=None, kpts_band=None, with_j=True, with_k=True,
               omega=None, exxdiv=None):
        """Signature and return convention of FFTDF.get_jk (pyscf/pbc/df/fft.py:298-315)."""
        import torch
        if omega is not None:
            raise NotImplementedError('range-separated Coulomb (omega) is not supported by ISDF')
        if kpts_band is not None:
            raise NotImplementedError('kpts_band is not supported by ISDF')
        if exxdiv not in (None, 'ewald'):
            raise NotImplementedError('exxdiv=%s' % exxdiv)
        if self._kmesh is None and abs(self.kpts).sum() > 1e-9:
            self.check_sanity()
        if self._kmesh is not None:
            # a uniform k-point mesh (KSCF.get_jk, pyscf/pbc/scf/khf.py:519-532) through its supercell
            kpts = self.kpts if kpts is None else np.reshape(kpts, (-1, 3))
            if kpts.shape != self.kpts.shape or abs(kpts - self.kpts).max() > 1e-9:
                raise NotImplementedError('ISDF at k-points: get_jk serves exactly the mesh it was built for')
            if self.W is None:
                self.build()
            from .kmesh import get_jk_kmesh
            out = get_jk_kmesh(self, dm, hermi, with_j, with_k, exxdiv)
            self.h2d_bytes, self.d2h_bytes = self._sc.h2d_bytes, self._sc.d2h_bytes
            return out
        if kpts is not None and abs(np.asarray(kpts)).sum() > 1e-9:
            raise NotImplementedError('ISDF is Gamma-point only unless built for a uniform k-point mesh')
        if self.W is None:
            self.build()                       # SCF never calls build() itself
        h = self.handle
        nao, npts = self.aoR_ip.shape
        dm0 = np.asarray(dm)
        if np.iscomplexobj(dm0):
            if abs(dm0.imag).max() > 1e-12:
                raise NotImplementedError('complex density matrices (Gamma point only)')
            dm0 = dm0.real
        dms = dm0.reshape(-1, nao, nao)                                           # df_jk.py:1411-1416
        nset = dms.shape[0]
        t0 = (time.process_time(), time.perf_counter())

        # pinned staging buffers live with the object: one host->device copy in, one device->host copy out
        stage = self._pinned
        if stage is None or stage[0].shape[0] < nset or stage[0].shape[1] != nao:
            stage = self._pinned = (torch.empty((nset, nao, nao), dtype=torch.float64).pin_memory(),
                                    torch.empty((2, nset, nao, nao), dtype=torch.float64).pin_memory())
        h_in, h_out = stage[0][:nset], stage[1][:, :nset]
        h_in.numpy()[...] = dms                                                   # casts integer DMs to float64
        d_dm = h_in.to(h.torch_device, non_blocking=True)
        self.h2d_bytes = h_in.numel() * 8
        vj, vk = self.get_jk_device(d_dm, with_j, with_k, exxdiv, mo=self._mo_tags(dm, dm0, nset, nao))
        self.d2h_bytes = 0
        if with_j:
            h_out[0].copy_(vj, non_blocking=True)
            self.d2h_bytes += vj.numel() * 8
        if with_k:
            h_out[1].copy_(vk, non_blocking=True)
            self.d2h_bytes += vk.numel() * 8
        torch.cuda.current_stream(h.torch_device).synchronize()
        out_j = h_out[0].numpy().reshape(dm0.shape).copy() if with_j else None
        out_k = h_out[1].numpy().reshape(dm0.shape).copy() if with_k else None
        self._timer('vj and vk', t0)
        return out_j, out_k

    @staticmethod
    def _mo_tags(dm, dm0, nset, nao):
        """(mo_coeff, mo_occ) carried by a tagged density matrix (pyscf/lib/numpy_helper.py tag_array; read by
        the reference's K build at pyscf/pbc/df/fft_jk.py:206-210): real orbitals with non-negative occupations
        of a single Gamma-point density matrix, else None."""
        mo_coeff, mo_occ = getattr(dm, 'mo_coeff', None), getattr(dm, 'mo_occ', None)
        if mo_coeff is None or mo_occ is None or nset != 1:
            return None
        C, occ = np.asarray(mo_coeff), np.asarray(mo_occ)
        if C.ndim == 3 and C.shape[0] == 1:
            C, occ = C[0], occ.reshape(1, -1)[0]
        if C.ndim != 2 or C.shape[0] != nao or occ.shape != (C.shape[1],) or np.iscomplexobj(C):
            return None
        if (occ < 0).any():
            return None
        return C, occ

    def get_jk_device(self, d_dm, with_j=True, with_k=True, exxdiv=None, mo=None):
        """Stage 5 on a device-resident density matrix [nset, nao, nao]; returns
        device tensors (no host copies; asynchronous on the handle's stream)."""
        if self.W is None:
            self.build()
        h = self.handle
        nao, npts = self.aoR_ip.shape
        nset = d_dm.shape[0]
        isdf_j = with_j and not self.exact_j
        d_cs = None
        if mo is not None and nset == 1 and (isdf_j or with_k):
            # occupied orbitals of a tagged density matrix: D = C diag(n) C^T, Cs = sqrt(n) C^T [nocc, nao]
            C, occ = mo
            keep = occ > 0
            if keep.any() and keep.sum() <= nao:
                cs = np.ascontiguousarray((C[:, keep] * np.sqrt(occ[keep])).T, dtype=np.float64)
                d_cs = h.from_host(cs)
        self.last_jk_form = 'occupied-orbital' if d_cs is not None else 'density-matrix'
        if self.comm is not None and (isdf_j or with_k):
            from . import dist                  # self.W is this rank's row block of W
            vj, vk = dist.get_jk_device(self, self.comm, d_dm, isdf_j, with_k, occ=d_cs)
        elif isdf_j or with_k:
            vj = h.empty((nset, nao, nao)) if isdf_j else None
            vk = h.empty((nset, nao, nao)) if with_k else None
            if self._jk_ws is None:
                self._jk_ws = h.workspace(h.lib.isdf_get_jk_ws_bytes(nao, npts))
            if d_cs is not None:
                h.get_jk_block_occ(self.aoR_ip, self.W, npts, 0, npts, d_cs, d_cs.shape[0], nao, npts,
                                   isdf_j, with_k, vj, vk, ws=self._jk_ws)
            else:
                h.get_jk(self.aoR_ip, self.W, d_dm, nset, nao, npts, isdf_j, with_k, vj, vk, ws=self._jk_ws)
        else:
            vj = vk = None
        if with_j and self.exact_j:
            vj = self.get_j_exact_device(d_dm)
        if with_k and exxdiv == 'ewald':
            # vk += madelung * S D S   (pyscf/pbc/df/df_jk.py:1446-1455)
            S = self.get_ovlp()
            mad = self.get_madelung()
            tmp = h.empty((nao, nao))
            sds = h.empty((nao, nao))
            for i in range(nset):
                h.dgemm(nao, nao, nao, S, nao, 0, d_dm[i], nao, 1, tmp, nao)            # S D
                h.dgemm(nao, nao, nao, tmp, nao, 0, S, nao, 1, sds, nao, alpha=mad)    # mad (S D) S
                vk[i] += sds
        return vj, vk

    def get_j_exact_device(self, d_dm):
        """J from the exact density on the FFT grid: rho = diag(Phi^T D Phi), vR = ifft(coulG fft(rho)).real,
        J = (vol/ng) Phi diag(vR) Phi^T (pyscf/pbc/df/fft_jk.py:33-109).  With a sharded grid every rank
        forms rho on its slab, the slabs are gathered (8 ng bytes), every rank solves the one-vector
        Poisson problem on the whole grid and contracts its slab; J is summed over ranks."""
        if self.W is None:
            self.build()
        h = self.handle
        aoR, ld = self._ao_on_grid()
        nao = aoR.shape[0]
        nset = d_dm.shape[0]
        ng = int(np.prod(self.mesh))
        g0, g1 = self._slab if self._slab is not None else (0, ng)
        ngl = g1 - g0
        ws = h.workspace(h.lib.isdf_grid_j_ws_bytes(nao, ld))
        ngp = _lib.pad8(ng)
        rho = h.zeros((nset, ngp))
        h.density_on_grid(aoR, ld, nao, ngl, d_dm, nset, rho[:, g0:], ngp, ws=ws)
        if self.comm is not None and self.comm.size > 1:
            self.comm.all_reduce_sum(rho)           # slabs are disjoint: the sum is the gather
        vR = h.empty((nset, ngp))
        h.coulomb_apply(rho, ngp, nset, self.mesh, self.cell.lattice_vectors(), vR, ngp)
        vj = h.empty((nset, nao, nao))
        h.j_from_potential(aoR, ld, nao, ngl, vR[:, g0:], ngp, nset, self.cell.vol / ng, vj, ws=ws)
        if self.comm is not None and self.comm.size > 1:
            self.comm.all_reduce_sum(vj)
        return vj

    # -- hcore pieces on the same device-resident AO values ("next" row 2 of SURVEY 8f) ----------
    def _check_gamma(self, kpts):
        if self._kmesh is not None or abs(self.kpts).sum() > 1e-9:
            raise NotImplementedError('ISDF at k-points: not served through the supercell')
        if kpts is not None and abs(np.asarray(kpts)).sum() > 1e-9:
            raise NotImplementedError('ISDF is Gamma-point only')

    def _supercell_df(self):
        """The Gamma-point ISDF object of the k-mesh's supercell (kmesh.py), created on first use, not built."""
        if self._sc is None:
            self._sc = ISDF(self._kmesh.supercell, c=self.c, n_ip=None, select=self.select, fit=self.fit,
                            tol=self.tol, device=self._device, keep_aoR=self.keep_aoR, comm=self.comm,
                            wgemm=self.wgemm, exact_j=self.exact_j)
            if self.n_ip:
                self._sc.n_ip = int(self.n_ip) * self._kmesh.nk
            self._sc.verbose, self._sc.stdout = self.verbose, self.stdout
        return self._sc

    def _on_kmesh(self, kpts):
        """True when the call is for the uniform k-point mesh this object was made for (then the supercell
        serves it); raises for any other k-point list."""
        if self._kmesh is None and abs(self.kpts).sum() > 1e-9:
            self.check_sanity()
        if self._kmesh is None:
            return False
        kpts = self.kpts if kpts is None else np.reshape(kpts, (-1, 3))
        if kpts.shape != self.kpts.shape or abs(kpts - self.kpts).max() > 1e-9:
            raise NotImplementedError('ISDF at k-points serves exactly the uniform mesh it was made for')
        return True

    def _folded(self, name, *args):
        """A lattice-periodic one-electron operator at the mesh's k-points from the supercell's Gamma-point
        matrix: V_k = 1/Nk sum_RR' exp(ik.(R'-R)) V_sc[(R mu),(R' nu)]  (kmesh.KMesh.mat_to_kpts)."""
        return self._kmesh.mat_to_kpts(getattr(self._supercell_df(), name)(*args))

    def _ao_on_grid(self):
        """AO values of this rank's grid slab: the build's copy when it was kept, else stage 1 alone
        (keep_aoR=False, or an object restored by load())."""
        if self.aoR is not None:
            return self.aoR, self._ld
        h = self.handle
        cell = self.cell
        ng = int(np.prod(self.mesh))
        if self.comm is not None and self.comm.size > 1 and self._slab is None:
            raise RuntimeError('the grid slabs are assigned by build() or load() on a sharded grid')
        g0, g1 = self._slab if self._slab is not None else (0, ng)
        ld = _lib.pad8(g1 - g0)
        rcut, Ls = cell.ao_eval_params()
        phi = h.empty((cell.nao_nr(), ld))
        if ld > g1 - g0:
            phi[:, g1 - g0:].zero_()
        h.eval_ao(cell._atm, cell._bas, cell._env,
                  np.ascontiguousarray(cell.lattice_vectors(), dtype=np.float64),
                  np.asarray(cell.mesh, dtype=np.int32), Ls, rcut, g0, g1, phi, ld)
        if self.keep_aoR:                       # e.g. after load(): evaluated once, then kept like build() does
            self.aoR, self._ld = phi, ld
        return phi, ld

    def potential_matrix_device(self, vR):
        """V = Phi diag(vR) Phi^T for a host vector vR[ng] (the contraction of
        pyscf/pbc/df/fft.py:55-56 and :79-81) as a device tensor [nao, nao]; with a sharded grid every
        rank contracts its slab and the partial matrices are summed."""
        h = self.handle
        phi, ld = self._ao_on_grid()
        nao = phi.shape[0]
        ng = int(np.prod(self.mesh))
        g0, g1 = self._slab if self._slab is not None else (0, ng)
        ngl = g1 - g0
        ngp = _lib.pad8(ngl)
        v = np.zeros((1, ngp))
        v[0, :ngl] = np.asarray(vR, dtype=np.float64).reshape(-1)[g0:g1]
        d_v = h.from_host(v)
        out = h.empty((1, nao, nao))
        h.j_from_potential(phi, ld, nao, ngl, d_v, ngp, 1, 1.0, out)
        if self.comm is not None and self.comm.size > 1:
            self.comm.all_reduce_sum(out)
        return out[0]

    def get_nuc(self, kpts=None):
        """Nuclear attraction of point charges -Z, G=0 removed (FFTDF.get_nuc, pyscf/pbc/df/fft.py:39-62)."""
        from . import pseudo
        if self._on_kmesh(kpts):
            return self._folded('get_nuc')
        self._check_gamma(kpts)
        return self.potential_matrix_device(pseudo.nuc_on_grid(self.cell, self.mesh)).cpu().numpy()

    def get_pp_loc(self, kpts=None):
        """Local part of the GTH pseudopotential matrix (pyscf/pbc/df/fft.py:70-84)."""
        from . import pseudo
        if self._on_kmesh(kpts):
            return self._folded('get_pp_loc')
        self._check_gamma(kpts)
        return self.potential_matrix_device(pseudo.pp_loc_on_grid(self.cell, self.mesh)).cpu().numpy()

    def get_pp_nl(self, kpts=None, on_device=True):
        """Non-local projector part of FFTDF.get_pp (pyscf/pbc/df/fft.py:86-141), in reciprocal space from
        analytic Fourier transforms of the AOs as the reference has it.  The per-species radial x angular
        tables are closed forms evaluated on the host per batch of G vectors; the structure-factor phases
        of every atom and the [projectors x G] . [G x nao] contraction run on the device
        (`isdf_phase_expand` + `isdf_dgemm`; `on_device=False` keeps everything on the host like the
        reference).  With several ranks each takes a contiguous range of reciprocal vectors and the
        projector-AO overlaps are summed over ranks before h_ij is applied."""
        from . import pseudo
        if self._on_kmesh(kpts):
            return self._folded('get_pp_nl', None, on_device)
        self._check_gamma(kpts)
        cell = self.cell
        nao = cell.nao_nr()
        if not pseudo.has_nonlocal(cell):
            return np.zeros((nao, nao))
        ng = int(np.prod(self.mesh))
        comm = self.comm
        rank, size = (comm.rank, comm.size) if comm is not None else (0, 1)
        g0, g1 = ng * rank // size, ng * (rank + 1) // size
        if on_device:
            import torch
            zre, zim = pseudo.pp_nl_overlaps_device(cell, self.handle, g0, g1, self.mesh)
            buf = torch.stack([zre, zim])
            if size > 1:
                comm.all_reduce_sum(buf)
            zz = buf.cpu().numpy()
            return pseudo.pp_nl_from_overlaps(cell, zz[0] + 1j * zz[1])
        Z = pseudo.pp_nl_overlaps(cell, g0, g1, self.mesh)
        if size > 1:
            buf = self.handle.from_host(np.stack([Z.real, Z.imag]))
            comm.all_reduce_sum(buf)
            zz = buf.cpu().numpy()
            Z = zz[0] + 1j * zz[1]
        return pseudo.pp_nl_from_overlaps(cell, Z)

    def get_pp(self, kpts=None):
        """FFTDF.get_pp (pyscf/pbc/df/fft.py:64-152): local part on the device-resident AO values,
        non-local projector part on the host as the reference has it."""
        from . import pseudo
        vpp = self.get_pp_loc(kpts)
        if pseudo.has_nonlocal(self.cell):
            vpp = vpp + self.get_pp_nl(kpts)
        return vpp

    # -- exchange-correlation on the same grid ("next" row 1 of SURVEY 8f) -----------------------
    XC_CACHE_BYTES = 48 << 30       # AO values + gradients are kept between SCF iterations up to this size

    def _ao_deriv1(self, c0, c1):
        """AO values and gradients of the grid points [c0, c1): a new device tensor [4, nao, ld]."""
        h = self.handle
        cell = self.cell
        ld = _lib.pad8(c1 - c0)
        nao = cell.nao_nr()
        rcut, Ls = cell.ao_eval_params()
        phi4 = h.zeros((4, nao, ld))
        h.eval_ao_deriv1(cell._atm, cell._bas, cell._env,
                         np.ascontiguousarray(cell.lattice_vectors(), dtype=np.float64),
                         np.asarray(cell.mesh, dtype=np.int32), Ls, rcut, c0, c1, phi4, ld, nao * ld)
        return phi4

    def _ao_deriv1_on_grid(self):
        """AO values and gradients of this rank's whole grid slab, evaluated on first use and kept
        (4 x the AO matrix; drop with ``self._aoR4 = None``)."""
        if self._aoR4 is None:
            ng = int(np.prod(self.mesh))
            g0, g1 = self._slab if self._slab is not None else (0, ng)
            self._aoR4 = self._ao_deriv1(g0, g1)
        return self._aoR4

    def nr_rks(self, xc_code, dm, kpts=None, max_bytes=None):
        """(nelec, exc, vmat) of a closed-shell density matrix at the Gamma point: NumInt.nr_rks of the
        reference (pyscf/pbc/dft/numint.py:283-381) for the functionals 'lda,', 'lda,vwn', 'blyp', 'pbe' and the semilocal parts of 'pbe0' / 'b3lyp', on the
        device: density and gradient from the AO values, the functional, and the potential matrix.  With a
        sharded grid every rank works on its slab and the three results are summed over ranks.

        The AO values and gradients of the slab (32 nao ngl bytes) are kept between calls when they fit in
        ``max_bytes`` (default ``XC_CACHE_BYTES``); a larger slab is walked in chunks of that size whose AO
        gradients are re-evaluated on every call, the way the reference's ``block_loop`` does (numint.py:354-356)."""
        if self._on_kmesh(kpts):
            # KNumInt.nr_rks (pyscf/pbc/dft/numint.py:1090-1204) for a uniform mesh: the density of the k-point
            # density matrices is the supercell density of D_sc (only its real symmetric part reaches rho with
            # real AOs); electrons and energy come per unit cell, the potential matrix is folded back
            km = self._kmesh
            dsc = km.dm_to_supercell(np.asarray(dm).reshape(km.nk, km.nao, km.nao))
            ne, exc, vsc = self._supercell_df().nr_rks(xc_code, np.ascontiguousarray(dsc.real), None, max_bytes)
            return ne / km.nk, exc / km.nk, km.mat_to_kpts(vsc)
        self._check_gamma(kpts)
        h = self.handle
        dm = np.asarray(dm, dtype=np.float64)
        if dm.ndim != 2 or abs(dm - dm.T).max() > 1e-10 * max(1.0, abs(dm).max()):
            raise ValueError('nr_rks: one symmetric density matrix expected')
        if self.comm is not None and self.comm.size > 1 and self._slab is None:
            raise RuntimeError('nr_rks on a sharded grid needs build() first (it assigns the grid slabs)')
        code = h.xc_code(xc_code)
        nao = self.cell.nao_nr()
        ng = int(np.prod(self.mesh))
        g0, g1 = self._slab if self._slab is not None else (0, ng)
        weight = self.cell.vol / ng
        max_bytes = self.XC_CACHE_BYTES if max_bytes is None else max_bytes
        chunk = max(56, int(max_bytes // (32 * nao)) // 56 * 56)      # whole 56-point blocks of the AO kernel
        d_dm = h.from_host(np.ascontiguousarray(dm))
        total = h.zeros((nao * nao + 2,))            # vmat and the two sums in one buffer: one all-reduce
        ws = None
        for c0 in range(g0, g1, chunk):
            c1 = min(c0 + chunk, g1)
            phi4 = self._ao_deriv1_on_grid() if chunk >= g1 - g0 else self._ao_deriv1(c0, c1)
            ld = phi4.shape[2]
            rho4 = h.zeros((4, ld))
            wv4 = h.zeros((4, ld))
            out = h.zeros((nao * nao + 2,))
            ws = h.xc_density(phi4, ld, nao * ld, nao, c1 - c0, d_dm, rho4, ld, ws if c1 - c0 == chunk else None)
            h.xc_potential(code, rho4, ld, c1 - c0, weight, wv4, out[nao * nao:], ws)
            h.xc_matrix(phi4, ld, nao * ld, nao, c1 - c0, wv4, ld, out[:nao * nao], ws)
            total += out
        if self.comm is not None and self.comm.size > 1:
            self.comm.all_reduce_sum(total)
        res = total.cpu().numpy()
        return float(res[-2]), float(res[-1]), res[:nao * nao].reshape(nao, nao).copy()

    def xc_hybrid_fraction(self, xc_code):
        """Exact-exchange fraction of a functional nr_rks knows ('pbe0' 0.25, 'b3lyp' 0.20, else 0): for the
        hybrids nr_rks returns the semilocal part and the caller adds ``-fraction/2 K`` from get_jk, as
        pyscf/pbc/dft/rks.py:103-118 does."""
        return self.handle.xc_hybrid_fraction(xc_code)

    def numint(self):
        """Object with the ``nr_rks`` call signature the reference's RKS uses on ``mf._numint``
        (pyscf/pbc/dft/rks.py:88-89), backed by :meth:`nr_rks`."""
        return ISDFNumInt(self)

    def get_j(self, dm, hermi=1, kpts=None, kpts_band=None, omega=None):
        return self.get_jk(dm, hermi, kpts, kpts_band, True, False, omega)[0]

    def get_k(self, dm, hermi=1, kpts=None, kpts_band=None, omega=None, exxdiv=None):
        return self.get_jk(dm, hermi, kpts, kpts_band, False, True, omega, exxdiv)[1]

    # -- the rest of FFTDF's surface (pyscf/pbc/df/fft.py:246-272, :317-359): not part of the J/K hot path.
    # Named here so that a caller reaching them through `mf.with_df` gets the reason, not an AttributeError
    # (pyscf/pbc/scf/hf.py:676-679 calls get_ao_eri whenever `_is_mem_enough()` was not patched by update_mf).
    def _unsupported(self, name, why):
        raise NotImplementedError('ISDF.%s: %s' % (name, why))

    def get_ao_eri(self, *args, **kwargs):
        self._unsupported('get_ao_eri', 'the 4-index integrals are never formed; J/K come from get_jk '
                          '(attach with ISDF.update_mf(mf) so that SCF does not take its in-core ERI shortcut)')

    get_eri = get_ao_eri

    def get_mo_eri(self, *args, **kwargs):
        self._unsupported('get_mo_eri', 'integral transformation is outside the ISDF J/K path')

    ao2mo = get_mo_eri

    def ao2mo_7d(self, *args, **kwargs):
        self._unsupported('ao2mo_7d', 'integral transformation is outside the ISDF J/K path')

    def get_ao_pairs_G(self, *args, **kwargs):
        self._unsupported('get_ao_pairs_G', 'pair densities are replaced by the interpolation (aoR_ip, W)')

    get_ao_pairs = get_ao_pairs_G

    def get_mo_pairs_G(self, *args, **kwargs):
        self._unsupported('get_mo_pairs_G', 'pair densities are replaced by the interpolation (aoR_ip, W)')

    get_mo_pairs = get_mo_pairs_G

    def aoR_loop(self, grids=None, kpts=None, deriv=0):
        self._unsupported('aoR_loop', 'the AO values stay on the device (ISDF.aoR, [nao, grid]); use aoR_host()')

    def loop(self, blksize=None):
        self._unsupported('loop', 'no 3-index tensor exists in ISDF')

    def range_coulomb(self, omega):
        self._unsupported('range_coulomb', 'range-separated Coulomb (omega) is not supported')

    def get_jk_e1(self, *args, **kwargs):
        self._unsupported('get_jk_e1', 'nuclear gradients are outside the ISDF J/K path')

    get_j_e1 = get_k_e1 = get_jk_e1

    def get_naoaux(self):
        """Number of auxiliary functions = interpolation points (FFTDF returns 2 ngrids, fft.py:350-353)."""
        if self.n_ip:
            return int(self.n_ip)
        return min(int(self.c * self.cell.nao_nr()), int(np.prod(self.mesh)))

    @property
    def grids(self):
        self._unsupported('grids', 'the uniform grid is generated on the device from the mesh; coordinates: cell.get_uniform_grids()')

    @property
    def _numint(self):
        return self.numint()

    def _timer(self, msg, t0):
        """The reference's `logger.timer` line (pyscf/lib/logger.py:167-180; call sites
        pyscf/pbc/scf/hf.py:666,697, pyscf/pbc/df/fft_jk.py:249,292), printed at verbose >= 5 (DEBUG)."""
        t1 = (time.process_time(), time.perf_counter())
        if self.verbose >= 5:
            self.stdout.write('    CPU time for %s %9.2f sec, wall time %9.2f sec\n' % (msg, t1[0] - t0[0], t1[1] - t0[1]))
        return t1

    # -- persistence ("next" row 3 of SURVEY 8f) ---------------------------------------------
    @staticmethod
    def _shard_path(path, rank, size):
        path = str(path)
        base = path[:-4] if path.endswith('.npz') else path
        return path if size == 1 else '%s.r%dof%d.npz' % (base, rank, size)

    def save(self, path):
        """Interpolation points, X = Phi[:, ip], the overlap (when it has been formed) and W.  With several
        ranks every rank writes its own file `<path>.r<rank>of<size>.npz` holding its row block of W
        (the role of the reference's `_cderi` cache, examples/pbc/11-gamma_point_all_electron_scf.py:66-75)."""
        if self.W is None:
            raise RuntimeError('nothing to save: build() first')
        rank, size = (self.comm.rank, self.comm.size) if self.comm is not None else (0, 1)
        p0 = int(self._part.ip_off[rank]) if (size > 1 and self._part is not None) else 0
        extra = {}
        if self._ovlp is not None:
            extra['ovlp'] = self._ovlp.cpu().numpy()
        np.savez(self._shard_path(path, rank, size), ip_index=self.ip_index, ip_resid=self.ip_resid,
                 W=self.W.cpu().numpy(), aoR_ip=self.aoR_ip.cpu().numpy(),
                 mesh=np.asarray(self.mesh), n_ip=self.n_ip, nao=self.cell.nao_nr(),
                 rank=rank, size=size, p0=p0, **extra)

    def load(self, path):
        """Restore what save() wrote.  The AO matrix on the grid is not stored: the pieces that need it
        (overlap for exxdiv='ewald' when it was not saved, exact_j, get_pp, nr_rks) evaluate stage 1 again."""
        import torch
        from .dist import SlabPartition
        rank, size = (self.comm.rank, self.comm.size) if self.comm is not None else (0, 1)
        z = np.load(self._shard_path(path, rank, size))
        if not np.array_equal(z['mesh'], np.asarray(self.mesh)):
            raise ValueError('mesh of the saved ISDF object does not match the cell')
        nao = self.cell.nao_nr()
        npts = int(z['n_ip'])
        if z['aoR_ip'].shape != (nao, npts):
            raise ValueError('saved ISDF object has %d AOs x %d points; the cell has %d AOs'
                             % (z['aoR_ip'].shape + (nao,)))
        fsize = int(z['size']) if 'size' in z.files else 1
        frank = int(z['rank']) if 'rank' in z.files else 0
        if (frank, fsize) != (rank, size):
            raise ValueError('file was written by rank %d of %d, this is rank %d of %d' % (frank, fsize, rank, size))
        part = SlabPartition(self.mesh, size, npts)
        p0, p1 = int(part.ip_off[rank]), int(part.ip_off[rank + 1])
        if z['W'].shape != (p1 - p0, npts) or len(z['ip_index']) != npts:
            raise ValueError('saved W has shape %s, expected %s' % (z['W'].shape, (p1 - p0, npts)))
        self.reset()
        h = self.handle
        self.ip_index, self.ip_resid, self.n_ip = z['ip_index'], z['ip_resid'], npts
        self.W = torch.from_numpy(z['W']).to(h.torch_device)
        self.aoR_ip = torch.from_numpy(z['aoR_ip']).to(h.torch_device)
        if 'ovlp' in z.files:
            self._ovlp = torch.from_numpy(z['ovlp']).to(h.torch_device)
        if size > 1:
            self._part = part
            self._slab = part.grid_range(rank)
        return self


class ISDFNumInt:
    """Stand-in for ``pyscf.pbc.dft.numint.NumInt`` on the exchange-correlation call of the SCF loop:
    ``ni.nr_rks(cell, grids, xc_code, dms, relativity, hermi, kpt, kpts_band, max_memory=...)``
    (pyscf/pbc/dft/rks.py:88-89; signature pyscf/pbc/dft/numint.py:283-285) -> ``(nelec, excsum, vmat)``,
    scalars and one matrix for a single density matrix, arrays over the set otherwise (numint.py:370-378).
    With PySCF present, subclass its NumInt and override ``nr_rks`` with this method so that the
    functional parsing helpers (``libxc.is_hybrid_xc`` ...) stay the reference's own."""

    def __init__(self, mydf):
        self.mydf = mydf

    def rsh_and_hybrid_coeff(self, xc_code, spin=0):
        """(omega, alpha, hyb) in the reference's convention (pyscf/dft/numint.py:rsh_and_hybrid_coeff):
        no range separation here, so (0, 0, exact-exchange fraction)."""
        return 0.0, 0.0, self.mydf.xc_hybrid_fraction(xc_code)

    def nr_rks(self, cell, grids, xc_code, dms, relativity=0, hermi=1, kpt=None, kpts_band=None,
               max_memory=2000, verbose=None):
        if kpts_band is not None:
            raise NotImplementedError('kpts_band')
        if hermi != 1:
            raise NotImplementedError('nr_rks: hermitian density matrices only (hermi=1)')
        dms = np.asarray(dms, dtype=np.float64)
        single = dms.ndim == 2
        out = [self.mydf.nr_rks(xc_code, dm, kpts=kpt) for dm in (dms[None] if single else dms)]
        if single:
            return out[0]
        return (np.array([o[0] for o in out]), np.array([o[1] for o in out]), np.stack([o[2] for o in out]))
