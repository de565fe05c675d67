"""Multi-GPU ISDF: one process per GPU, grid sharded by x-planes.

Host-side orchestration only -- every number is produced by the sm_100a
kernels behind the C-ABI (``_lib.Handle``); ``torch.distributed`` (NCCL over
NVLink on the GPU box, gloo in the CPU unit tests) carries the exchanges.
The reference has no distributed form of this path;
the partitioning below is this repository's design:

  stage 1  AO values         rank r evaluates its slab of x-planes       no exchange
  stage 2  IP selection      per panel of <=128 pivots: every rank packs its
                             `quota` largest residuals (+ their phi / factor
                             columns), ONE all-gather, then every rank factors
                             the same candidate set redundantly (pivots are
                             exact: certificate max_C d > max_notC d) and
                             updates its own slab                          all-gather
  stage 3  aux fit           fit='pivot_factor': zeta = U^-1 Lt, columns
                             independent.  fit='implicit' (default): zeta is
                             never formed, see stage 4                     no exchange
  stage 4  Coulomb metric W  slab FFT: local z,y passes, ONE all-to-all to
                             (y,z)-column slabs, x pass fused with the
                             Coulomb weight, partial metric over the rank's
                             spectral slab, summed.  'implicit': the metric
                             M = Lt K Lt^T of the factor rows is formed and
                             W = U^-1 M U^-T follows from two triangular
                             solves sharded by columns (one small all-to-all
                             transposes between them)                      all-to-all + all-reduce
  stage 5  J/K               rank r contracts its row block of W           all-reduce of J,K

Results are bit-identical for any number of ranks in stages 1-3 (same FMA
chains, exact pivots); W, J, K differ only by the summation order of the
final reductions.
"""
import os

import numpy as np

from . import _lib

# upper bound of one all-to-all batch of half spectra (per rank, bytes)
A2A_BATCH_BYTES = int(os.environ.get('ISDF_A2A_BATCH_BYTES', 1 << 30))


class Comm:
    """Thin wrapper over a torch.distributed process group (float64 payloads only)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.size = dist.get_world_size(group)

    def all_gather(self, out, inp):
        self.dist.all_gather_into_tensor(out, inp, group=self.group)

    def all_to_all(self, out, inp, out_splits, in_splits):
        self.dist.all_to_all_single(out, inp, output_split_sizes=list(out_splits),
                                    input_split_sizes=list(in_splits), group=self.group)

    def all_reduce_sum(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)

    def max_scalar(self, x, like):
        t = like.new_tensor([float(x)])
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return float(t.item())


class SerialComm:
    """Single-rank stand-in (no process group needed)."""
    rank, size = 0, 1

    def all_gather(self, out, inp):
        out.copy_(inp)

    def all_to_all(self, out, inp, out_splits, in_splits):
        out.copy_(inp)

    def all_reduce_sum(self, t):
        pass

    def max_scalar(self, x, like):
        return float(x)


def _kept(mydf, ops, name, numel, dtype=None, zero=False):
    """A work buffer of the build that stays with the ISDF object between builds (ISDF.keep_buffers): handing
    gigabyte blocks back to the caching allocator after every build lets it split and re-map them, which showed
    up as sporadic 0.3 - 1 s stalls of a stage.  Returns a flat view of `numel` elements."""
    keep = getattr(mydf, 'keep_buffers', True)
    total = ops.device_mem_bytes() if hasattr(ops, 'device_mem_bytes') else 1 << 62
    nbytes = int(numel) * (1 if dtype is not None else 8)          # dtype is only ever uint8 here
    keep = keep and nbytes <= 0.15 * total                          # e.g. Si216's 50 GB product workspace stays transient
    cache = mydf.__dict__.setdefault('_bufs', {}) if keep else {}
    t = cache.get(name)
    if t is None or t.numel() < numel or (dtype is not None and t.dtype != dtype):
        cache.pop(name, None)
        t = None
        t = ops.empty((int(numel),), dtype=dtype) if dtype is not None else ops.empty((int(numel),))
        cache[name] = t
    v = t[:int(numel)]
    if zero:
        v.zero_()
    return v


class SlabPartition:
    """Who owns what: x-planes (stages 1-3 and the z,y FFT passes), (y,z)-columns of
    the half spectrum (x pass, W), rows of W (stage 5)."""

    def __init__(self, mesh, nranks, npts):
        nx, ny, nz = (int(m) for m in mesh)
        self.mesh = (nx, ny, nz)
        self.nranks = nranks
        self.nzh = nz // 2 + 1
        self.ncols = ny * self.nzh
        if nranks > nx:
            raise ValueError('more ranks (%d) than x-planes (%d)' % (nranks, nx))
        self.x_off = np.array([(r * nx) // nranks for r in range(nranks + 1)], dtype=np.int32)
        self.col_off = np.array([(r * self.ncols) // nranks for r in range(nranks + 1)], dtype=np.int32)
        self.set_npts(npts)

    def set_npts(self, npts):
        # row blocks of W: multiples of 8 so that every block start is 16-byte aligned
        n8 = (npts + 7) // 8
        self.ip_off = np.array([min(npts, 8 * ((r * n8) // self.nranks)) for r in range(self.nranks + 1)],
                               dtype=np.int64)
        self.ip_off[-1] = npts

    def grid_range(self, r):
        nyz = self.mesh[1] * self.mesh[2]
        return int(self.x_off[r]) * nyz, int(self.x_off[r + 1]) * nyz

    def nx_local(self, r):
        return int(self.x_off[r + 1] - self.x_off[r])

    def ncol_local(self, r):
        return int(self.col_off[r + 1] - self.col_off[r])


def build(mydf, comm, ops=None):
    """Stages 1-4 on `comm.size` ranks.  Fills the same attributes as ISDF.build()."""
    import time
    ops = ops or mydf.handle
    cell = mydf.cell
    mesh = np.asarray(cell.mesh, dtype=np.int32)
    nx, ny, nz = (int(m) for m in mesh)
    ng = nx * ny * nz
    nao = cell.nao_nr()
    npts = int(mydf.n_ip) if mydf.n_ip else int(mydf.c * nao)
    npts = min(npts, ng)
    a = np.ascontiguousarray(cell.lattice_vectors(), dtype=np.float64)
    N, r = comm.size, comm.rank
    part = SlabPartition(mesh, N, npts)
    g0, g1 = part.grid_range(r)
    ngl = g1 - g0
    nxl = part.nx_local(r)
    ldl = _lib.pad8(ngl)
    lib = _lib.load_library()
    t0 = time.perf_counter()
    ev = ops.timer()

    # stage 1: AO values on the local slab
    rcut, Ls = cell.ao_eval_params()
    phi = ops.empty((nao, ldl))
    if ldl > ngl:
        phi[:, ngl:].zero_()
    ops.eval_ao(cell._atm, cell._bas, cell._env, a, mesh, Ls, rcut, g0, g1, phi, ldl)
    ev.mark('ao_ms')

    # stage 2: interpolation points
    NB = lib.isdf_select_panel_width()
    # The two big buffers (factor rows, weighted spectra) are two flat allocations of one common
    # size owned by the ISDF object: handing tens of GB back to the caching allocator between
    # rebuilds lets it split the blocks for small requests and map new ones (seen as 0.2-0.5 s
    # stalls in the AO / selection / FFT stages of later builds).
    ncl_r = part.ncol_local(r)
    n_lt, n_y = npts * ldl, 2 * npts * nx * ncl_r            # doubles: factor rows, weighted spectra of this rank
    nbig = max(n_lt, n_y)
    # One arena or two?  Two (factor rows and spectra side by side) is the fast layout.  When that does not fit the
    # device next to the other buffers of the build, the spectra overwrite the factor rows as they are consumed
    # (`alias`): the rows are placed so that spectrum row v ends before factor row v+1 starts (right-aligned when
    # a spectrum row is the longer one), and every batch goes through a bounce buffer.
    total = ops.device_mem_bytes() if hasattr(ops, 'device_mem_bytes') else 1 << 62
    wmode = str(getattr(mydf, 'wgemm', 'i8'))
    planes = int(wmode.split(':')[1]) if ':' in wmode else 8
    ws_w = lib.isdf_syrk_i8_ws_bytes(npts, 2 * nx * ncl_r, planes) if (wmode.startswith('i8') and hasattr(lib, 'isdf_syrk_i8_ws_bytes')) else 0
    other = 8 * (nao * ldl + 3 * npts * _lib.pad8(npts)) + ws_w + (3 << 30)
    alias = os.environ.get('ISDF_ALIAS_ARENAS')
    alias = (16 * nbig + other > 0.92 * total) if alias is None else (alias == '1')
    arena = getattr(mydf, '_arena', None)
    if arena is None or arena[0].numel() < nbig + 32 or (arena[1] is None) != alias:
        mydf._arena = arena = None                # release the old buffers before mapping larger ones
        arena = (ops.empty((nbig + 32,)), None if alias else ops.empty((nbig + 32,)))
        # kept by the object across reset()/build() cycles unless that would pin most of the device
        if getattr(mydf, 'keep_buffers', True) and (8 if alias else 16) * nbig <= 0.45 * total:
            mydf._arena = arena
    big_lt, big_y = arena
    del arena
    if alias:
        big_y = big_lt
        # right-aligned (to 256 bytes: the GEMMs want their rows on sector boundaries) when a spectrum row is the longer one
        lt_off = (nbig - n_lt + 31) // 32 * 32 if 2 * nx * ncl_r >= ldl else 0
        if 2 * nx * ncl_r < ldl and n_y > n_lt:
            raise RuntimeError('aliased build buffers: inconsistent row sizes')
    else:
        lt_off = 0
    Lt = big_lt[lt_off:lt_off + n_lt].view(npts, ldl)
    ldu = _lib.pad8(npts)
    import torch
    U = _kept(mydf, ops, 'U', npts * ldu, zero=True).view(npts, ldu)
    sel_ws = _kept(mydf, ops, 'select_ws', ops.select_ws_bytes(nao, ngl, npts), dtype=torch.uint8) \
        if hasattr(ops, 'select_ws_bytes') else None
    plan = ops.select_create(nao, ngl, g0, ng, npts, ws=sel_ws)
    dmax = comm.max_scalar(plan.begin(phi, ldl), phi)
    tol = mydf.tol if mydf.tol >= 0 else float(ng) * 2.220446049250313e-16 * dmax
    plan.set_tol(tol)
    # the quota must be the same on every rank (equal all-gather contributions)
    ngl_min = min(part.grid_range(q)[1] - part.grid_range(q)[0] for q in range(N))
    cap = min(lib.isdf_select_max_rows() // N - 1, ngl_min)
    quota = min(lib.isdf_select_first_quota(N), cap)
    rl_max = lib.isdf_select_row_len(nao, npts)
    blockbuf = _kept(mydf, ops, 'blockbuf', (cap + 1) * rl_max)
    allbuf = _kept(mydf, ops, 'allbuf', N * (cap + 1) * rl_max) if N > 1 else blockbuf
    k0 = 0
    npanels = 0
    prof = ops.timer() if os.environ.get('ISDF_PROFILE_SELECT') else None      # per-step breakdown
    acc = {'sel_candidates_ms': 0.0, 'sel_allgather_ms': 0.0, 'sel_panel_ms': 0.0, 'sel_flush_ms': 0.0}
    while k0 < npts:
        nbmax = min(NB, npts - k0)
        n_el = (quota + 1) * lib.isdf_select_row_len(nao, k0)
        block = blockbuf[:n_el]
        if prof:
            prof = ops.timer()
        plan.candidates(phi, ldl, Lt, ldl, k0, quota, block)
        if prof:
            prof.mark('sel_candidates_ms')
        blocks = allbuf[:N * n_el]
        if N > 1:
            comm.all_gather(blocks, block)
        if prof:
            prof.mark('sel_allgather_ms')
        nb, stop = plan.panel(blocks, N, quota, k0, nbmax, U, ldu)
        if prof:
            prof.mark('sel_panel_ms')
        plan.flush(phi, ldl, Lt, ldl, k0, nb)
        if prof:
            prof.mark('sel_flush_ms')
            for kk, vv in prof.times.items():
                acc[kk] += vv
        k0 += nb
        npanels += 1
        if stop or nb == 0:
            break
        quota = min(lib.isdf_select_next_quota(quota, nb, nbmax, N), cap)
    ip, resid = plan.finish(k0)
    plan.close()
    del blockbuf, allbuf
    npts = k0
    part.set_npts(npts)
    ev.mark('select_ms')

    # stage 3: zeta = U^-1 Lt on the local slab (in place) -- skipped by fit='implicit',
    # which transforms the metric of the factor rows afterwards: W = U^-1 (Lt K Lt^T) U^-T
    implicit = (mydf.fit == 'implicit')
    if not implicit:
        ops.fit_aux_from_U(U, ldu, npts, Lt, ldl, ngl)
    ev.mark('fit_ms')

    # X = Phi[:, ip]: every rank contributes the pivots that live in its slab
    mine = np.nonzero((ip >= g0) & (ip < g1))[0]
    if len(mine) == npts:                       # one rank owns every pivot: gather straight into X
        X = ops.empty((nao, npts))
        ops.gather_cols(phi, ldl, nao, ip - g0, X, npts)
    else:
        X = ops.zeros((nao, npts))
        if len(mine):
            tmp = ops.empty((nao, len(mine)))
            ops.gather_cols(phi, ldl, nao, ip[mine] - g0, tmp, len(mine))
            X[:, ops.index(mine)] = tmp
    comm.all_reduce_sum(X)

    # stage 4: Coulomb metric W through the slab-decomposed FFT
    ncols, ncl, col0 = part.ncols, part.ncol_local(r), int(part.col_off[r])
    ldy = nx * ncl
    Yr = big_y[:2 * npts * ldy].view(npts, 2 * ldy)
    Y = ops.as_complex(Yr)
    mydf.timings_flags = {'aliased_arenas': bool(alias)}
    # vectors per all-to-all batch: computed from rank-independent maxima so that every rank
    # runs the same number of exchanges
    nxl_max = max(part.nx_local(q) for q in range(N))
    ncl_max = max(part.ncol_local(q) for q in range(N))
    per_vec = max(nxl_max * ncols, nx * ncl_max) * 16
    vb = int(max(1, min(npts, A2A_BATCH_BYTES // per_vec)))
    if N > 1:
        spec = ops.empty_complex((vb * nxl * ncols,))
        send = ops.empty_complex((vb * nxl * ncols,))
        recv = ops.empty_complex((vb * nx * ncl,))
    elif alias:
        spec = ops.empty_complex((vb * ldy,))
    for v0 in range(0, npts, vb):
        nv = min(vb, npts - v0)
        if N == 1:
            if alias:      # bounce: the batch's spectra land on its own (already transformed) factor rows
                ops.fft_zy(Lt[v0:], ldl, nv, nxl, mesh, spec, ldy)
                Y[v0:v0 + nv].copy_(spec[:nv * ldy].view(nv, ldy))
            else:
                ops.fft_zy(Lt[v0:], ldl, nv, nxl, mesh, Y[v0:], ldy)
            continue
        if prof:
            prof = ops.timer()
        ops.fft_zy(Lt[v0:], ldl, nv, nxl, mesh, spec, nxl * ncols)
        ops.fft_pack(spec, nxl * ncols, nv, nxl, ncols, part.col_off, send)
        if prof:
            prof.mark('fft_zy_pack_ms')
        in_splits = [2 * nv * nxl * part.ncol_local(s) for s in range(N)]
        out_splits = [2 * nv * part.nx_local(s) * ncl for s in range(N)]
        comm.all_to_all(ops.as_real(recv)[:sum(out_splits)], ops.as_real(send)[:sum(in_splits)],
                        out_splits, in_splits)
        if prof:
            prof.mark('fft_alltoall_ms')
        ops.fft_unpack(recv, nv, part.x_off, ncl, Y[v0:], ldy)
        if prof:
            prof.mark('fft_unpack_ms')
            for kk, vv in prof.times.items():
                acc[kk] = acc.get(kk, 0.0) + vv
    del Lt, big_lt                              # the factor rows are not needed past this point
    if prof:
        prof = ops.timer()
    ops.fft_x_coulomb(Y, ldy, npts, mesh, a, col0, ncl)
    if prof:
        prof.mark('fft_x_ms')
        acc.update(prof.times)
    ev.mark('fft_ms')
    W = ops.empty((npts, npts))
    if wmode.startswith('i8'):
        w_ws = _kept(mydf, ops, 'syrk_ws', ops.syrk_i8_ws_bytes(npts, 2 * nx * ncl, planes), dtype=torch.uint8) \
            if hasattr(ops, 'syrk_i8_ws_bytes') else None
        ops.syrk_i8(npts, 2 * nx * ncl, Yr, 2 * ldy, W, npts,
                    alpha=cell.vol / (float(ng) * float(ng)), nslices=planes, ws=w_ws)
    else:
        ops.dgemm(npts, npts, 2 * nx * ncl, Yr, 2 * ldy, 0, Yr, 2 * ldy, 0, W, npts,
                  alpha=cell.vol / (float(ng) * float(ng)), upper_only=1)
    ops.symmetrize_upper(W, npts, npts)
    comm.all_reduce_sum(W)
    ev.mark('wgemm_ms')
    del Y, Yr, big_y
    p0, p1 = int(part.ip_off[r]), int(part.ip_off[r + 1])
    if implicit:
        W = _congruence_solve(ops, comm, part, U, ldu, W, npts, mydf)
        ev.mark('wsolve_ms')
    else:
        W = W[p0:p1]
    ops.synchronize()
    del U

    mydf.timings = dict(ev.times)
    mydf.timings['coulomb_ms'] = (ev.times.get('fft_ms', 0.0) + ev.times.get('wgemm_ms', 0.0) +
                                  ev.times.get('wsolve_ms', 0.0))
    mydf.timings['build_wall_s'] = time.perf_counter() - t0
    mydf.timings['select_panels'] = npanels
    if prof:
        mydf.timings.update(acc)
    mydf.n_ip = npts
    mydf.ip_index = ip
    mydf.ip_resid = resid
    mydf.W = W
    mydf.aoR_ip = X
    mydf._ld = ldl
    mydf._part = part
    mydf._slab = (g0, g1)
    mydf.aoR = phi if mydf.keep_aoR else None
    mydf._jk_ws = None
    return mydf


def _congruence_solve(ops, comm, part, U, ldu, M, npts, mydf=None):
    """Rows [p0, p1) of W = U^-1 M U^-T for this rank (M symmetric, replicated; U upper
    triangular, replicated).  Both solves act on independent columns, so rank r takes the
    column block [p0, p1); the transposition between them is one all-to-all of Np*Np/N doubles."""
    N, r = comm.size, comm.rank
    if N == 1 and hasattr(ops, 'congruence'):           # one GPU: both solves in place, the second on the upper trapezoid only
        ops.congruence(U, ldu, npts, M, npts)
        return M
    off = [int(x) for x in part.ip_off]
    p0, p1 = off[r], off[r + 1]
    ncr = p1 - p0                                       # may be 0 on a rank when Np < 8 N: still take part
    Z1 = _kept(mydf, ops, 'cong_Z1', npts * ncr).view(npts, ncr) if mydf is not None else ops.empty((npts, ncr))
    if ncr:
        Z1.copy_(M[:, p0:p1])                           # column block of M: a pitched device-to-device copy
    del M
    if ncr:
        ops.fit_aux_from_U(U, ldu, npts, Z1, ncr, ncr)  # Z1 = U^-1 M[:, cols]
    R = _kept(mydf, ops, 'cong_R', npts * ncr).view(npts, ncr) if mydf is not None else ops.empty((npts, ncr))
    if N == 1:
        ops.transpose(Z1, ncr, npts, ncr, R, ncr)       # R = Z1^T (square here: ncr == Np)
    else:
        recv = ops.empty((npts * ncr,))
        splits = [(off[s + 1] - off[s]) * ncr for s in range(N)]
        comm.all_to_all(recv, Z1.reshape(-1), splits, splits)
        o = 0
        for s in range(N):
            ncs = off[s + 1] - off[s]
            if ncs and ncr:                             # rank s sent its Z1[p0:p1, :] = [ncr, ncs]; R[off_s:off_s+ncs, :] is its transpose
                ops.transpose(recv[o:o + ncr * ncs], ncs, ncr, ncs, R[off[s]:off[s + 1]], ncr)
            o += ncr * ncs
        del recv
    del Z1
    if ncr:
        ops.fit_aux_from_U(U, ldu, npts, R, ncr, ncr)   # R = U^-1 (U^-1 M)^T[:, cols] = W[:, cols]
    Wrows = ops.empty((ncr, npts))
    if ncr:
        ops.transpose(R, ncr, npts, ncr, Wrows, npts)   # W[cols, :] by symmetry
    return Wrows


def get_jk_device(mydf, comm, d_dm, with_j=True, with_k=True, ops=None, occ=None):
    """Stage 5: every rank contracts its row block of W, then J and K are all-reduced.
    occ: device tensor Cs[nocc, nao] = sqrt(n) C^T of a single tagged density matrix (occupied-orbital form)."""
    ops = ops or mydf.handle
    nao, npts = mydf.aoR_ip.shape
    part = mydf._part
    r = comm.rank
    p0, p1 = int(part.ip_off[r]), int(part.ip_off[r + 1])
    nset = d_dm.shape[0]
    out = ops.zeros((2, nset, nao, nao))
    vj = out[0] if with_j else None
    vk = out[1] if with_k else None
    if p1 > p0:
        # mydf.W holds this rank's row block W[p0:p1, :] (the whole matrix on one rank)
        if occ is not None and nset == 1 and hasattr(ops, 'get_jk_block_occ'):
            mydf._jk_ws = ops.get_jk_block_occ(mydf.aoR_ip, mydf.W, npts, p0, p1 - p0, occ, occ.shape[0], nao, npts,
                                               with_j, with_k, vj, vk, ws=mydf._jk_ws)
        else:
            mydf._jk_ws = ops.get_jk_block(mydf.aoR_ip, mydf.W, npts, p0, p1 - p0, d_dm, nset, nao, npts,
                                           with_j, with_k, vj, vk, ws=mydf._jk_ws)
    comm.all_reduce_sum(out)
    return vj, vk
