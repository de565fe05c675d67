#!/usr/bin/env python
"""Benchmark of the ISDF hot path (BASELINE.json metric: ISDF build s + J/K s
per SCF iteration).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c64] [--c 15]
  python bench.py --impl reference ...     # CPU arm: the NumPy/SciPy oracle port

One "step" = one full ISDF build (AO values -> interpolation points -> aux fit
-> Coulomb metric W) plus one get_jk(with_j, with_k) on a synthetic cell.
`value` is measured with everything device-resident; `e2e` goes through the
plugin's public host API (ISDF.build() + ISDF.get_jk(numpy dm) -> numpy J/K).
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'ISDF build s + J/K s per SCF iter'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default=None, help='c8 | c64 | h2o32 | si216 | c8-small')
    ap.add_argument('--c', type=float, default=15.0, help='ISDF rank parameter: N_IP = c * nao')
    ap.add_argument('--fit', default='implicit', choices=['implicit', 'pivot_factor', 'cholesky'])
    ap.add_argument('--wgemm', default='i8',
                    help="Coulomb-metric product Y Y^T: 'fp64' (FP64 tensor pipe) or 'i8[:planes]' "
                         "(int8 tcgen05 split product, 8 digit planes = FP64-GEMM accuracy)")
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--jk-reps', type=int, default=5)
    ap.add_argument('--no-e2e', action='store_true', help='exploratory runs only: skip the host-API leg')
    ap.add_argument('--no-parity', action='store_true', help='skip the C8-vs-oracle parity block (N = 1 only)')
    ap.add_argument('--no-extra', action='store_true',
                    help='skip the one-off runs of the larger configs (h2o32 at every N, si216 at 8 GPUs)')
    ap.add_argument('--exxdiv', default='ewald', choices=['ewald', 'none'],
                    help="exchange divergence treatment of the timed J/K: 'ewald' is the pbc.scf.RHF default")
    ap.add_argument('--xc', default=None,
                    help="also time ISDF.nr_rks(XC, dm) per SCF iteration ('pbe' for BASELINE configs[1]); reported "
                         "beside the headline under 'xc', never part of 'value'")
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clock / throttle-reason sampler running beside the timed region."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None
        # sampling period (ms); ISDF_BENCH_SMI_MS=0 switches the sampler off (diagnostics only: the contract wants it on)
        self.period = int(os.environ.get('ISDF_BENCH_SMI_MS', '200'))

    def start(self):
        if self.period <= 0:
            return self
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                 '--format=csv,noheader,nounits', '-lms', str(self.period)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')] + [time.time()])      # arrival time = sample time

    def mean_clock_between(self, marks):
        """Mean SM clock of the samples that arrived inside each interval marks[i] .. marks[i+1] (host wall times)."""
        out = []
        for t0, t1 in zip(marks[:-1], marks[1:]):
            v = []
            for r in self.rows:
                if len(r) >= 8 and t0 <= r[-1] < t1:
                    try:
                        v.append(float(r[0]))
                    except ValueError:
                        pass
            out.append(float(np.mean(v)) if v else None)
        return out

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(nm)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples']}
        # median over samples taken under load (power above the idle floor)
        load = [s for s, p in zip(sm, pw) if p > 0.5 * max(pw)] or sm
        return {'sm_mhz': float(np.median(load)), 'sm_mhz_mean': float(np.mean(load)), 'sm_max_mhz': float(max(mx)),
                'sm_mhz_min_under_load': float(min(load)), 'sm_mhz_p10_under_load': float(np.percentile(load, 10)),
                'power_w_max': float(max(pw)), 'power_w_median_under_load': float(np.median([p for p in pw if p > 0.5 * max(pw)])),
                'samples': len(sm), 'period_ms': self.period, 'reasons': sorted(reasons)}


# --------------------------------------------------------------------------- CPU arm
_POOL_LIMIT = None


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 when it starts more than one process per node, and NumPy's BLAS
    reads that at import: lift the BLAS / OpenMP pools back to every host core for the CPU legs."""
    global _POOL_LIMIT
    ncores = os.cpu_count() or 1
    try:
        import threadpoolctl
        _POOL_LIMIT = threadpoolctl.threadpool_limits(limits=ncores)      # kept alive for the whole run
    except Exception:                                                     # noqa: BLE001
        pass
    os.environ['OMP_NUM_THREADS'] = str(ncores)                           # for libgomp (the C oracle) if not yet loaded
    try:
        import ctypes
        ctypes.CDLL('libgomp.so.1').omp_set_num_threads(ncores)
    except OSError:
        pass
    return ncores


def cpu_blocked_select_seconds(phi_slab, npts, ng, nb=128):
    """Seconds of the PANEL-BLOCKED selection on the host BLAS (the algorithm of csrc/select.cu and
    oracle/c/isdf_oracle_blocked.c: per panel of 128 pivots one pair-density GEMM and one flush GEMM that streams
    the finished factor columns once), measured on a slab of grid points and scaled to the grid: the two
    GEMMs of a panel at k0 = Np/8, 3Np/8, 5Np/8, 7Np/8 (cost affine in k0: the midpoint rule is exact),
    best of 3 each.  The sequential pivot search on the candidates is not included (<2 % of the flops)."""
    import scipy.linalg.blas as blas
    nao, ngs = phi_slab.shape
    rng = np.random.default_rng(3)
    Lt = np.asfortranarray(rng.standard_normal((ngs, npts)) * 1e-3)      # [grid, column]: dgemm reads it as is
    per_panel = []
    for frac in (1, 3, 5, 7):
        k0 = max(nb, npts * frac // 8)
        P = np.asfortranarray(rng.standard_normal((k0, nb)))
        X = np.asfortranarray(rng.standard_normal((nao, nb)))
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            G = blas.dgemm(1.0, phi_slab, X, trans_a=True)                # pair densities [ngs, nb]
            G *= G
            blas.dgemm(-1.0, Lt[:, :k0], P, beta=1.0, c=G, overwrite_c=True)   # flush
            best = min(best, time.perf_counter() - t0)
        per_panel.append(best)
    return float(np.mean(per_panel)) * (npts / nb) * (ng / ngs)


def cpu_sample_target(cell, npts, seed=5):
    """One bounded sample of the CPU port on the TARGET workload itself (its cell, nao, mesh, Np):
    every stage of the NumPy/SciPy oracle is run on a slab of that workload's own arrays and the
    slab time is multiplied by the slab count (the oracle's stages are loops over grid points,
    pivots or aux functions, so a slab is a fixed fraction of the stage).  Returns the per-stage
    seconds scaled to the whole workload and a description of what was run.

      ao      oracle.eval_ao on ng/64 grid points of the cell                          x 64
      select  oracle.pivot_step on ng/16 and ng/32 points at k = Np/8, 3Np/8, 5Np/8, 7Np/8 (best of
              7 pivots each; the step cost is affine in k, so the midpoint rule is exact); the two
              slab sizes give the per-grid-point slope, extended to the whole grid      x Np
      fit     A = (X^T X)^2 and its Cholesky factorisation in full; B and the two triangular
              solves on ng/256 grid points                                               x 256
      fft     oracle.build_V (fft, coulG, ifft) of 16 aux functions on the full mesh     x Np/16
      wgemm   zeta V^T over ng/256 grid points                                           x 256
      jk      oracle.isdf_get_jk in full (one call)
    """
    import oracle
    import scipy.linalg
    from oracle import isdf_ref
    from pyscf_isdf_b200.workloads import synthetic_dm
    rng = np.random.default_rng(seed)
    nao = cell.nao
    mesh = list(cell.mesh)
    ng = int(np.prod(mesh))
    a = cell.lattice_vectors()
    rcut, Ls = cell.ao_eval_params()
    coords = cell.get_uniform_grids()
    t = {}
    # stage 1 on a slab of ng/64 points spread over the cell
    sl = np.arange(0, ng, 64)
    ngs = len(sl)
    t0 = time.perf_counter()
    phi_s = np.ascontiguousarray(oracle.eval_ao(cell._atm, cell._bas, cell._env, coords[sl], Ls, rcut).T)
    t['ao'] = (time.perf_counter() - t0) * (ng / ngs)
    # stage 2: per-pivot cost on that slab with a factor of the real shape
    # (the AO slab is repeated: the step's time does not depend on the values.  Two slab sizes,
    # ng/32 and ng/16, separate the per-call overhead from the part that grows with the grid:
    # a full-size step costs  T(ng/16) + slope * (ng - ng/16), the overhead counted once.)
    phi_sel = np.ascontiguousarray(np.tile(phi_s, (1, 4)))
    ngsel = phi_sel.shape[1]
    half = ngsel // 2
    Lt = np.full((npts, ngsel), 1e-3)
    Lt_half = np.full((npts, half), 1e-3)
    phi_half = np.ascontiguousarray(phi_sel[:, :half])
    d = np.einsum('ig,ig->g', phi_sel, phi_sel) ** 2 + 1.0
    d_half = d[:half].copy()

    def one_step(ph, L, dd, k):
        best = 1e30
        for r in range(7):                  # best of 7: thread wake-ups of the BLAS pool are noisy on shared hosts
            pcol = int(rng.integers(ph.shape[1]))
            dd[pcol] = abs(dd[pcol]) + 1.0
            t0 = time.perf_counter()
            int(np.argmax(dd))
            isdf_ref.pivot_step(ph, L, dd, k, pcol)
            best = min(best, time.perf_counter() - t0)
        return best

    per_pivot = []
    for frac in (1, 3, 5, 7):
        k = max(1, npts * frac // 8)
        t_full = one_step(phi_sel, Lt, d, k)
        t_half = one_step(phi_half, Lt_half, d_half, k)
        slope = max(t_full - t_half, 0.0) / (ngsel - half)
        per_pivot.append(t_full + slope * (ng - ngsel))
    t['select_unblocked'] = float(np.mean(per_pivot)) * npts
    del Lt, Lt_half, phi_sel, phi_half
    # the same selection panel-blocked on the host BLAS: this is the figure the CPU total uses
    t['select'] = cpu_blocked_select_seconds(np.ascontiguousarray(np.tile(phi_s, (1, 4))), npts, ng)
    # stage 3 (the oracle's route: Cholesky of A, two triangular solves)
    ipx = rng.integers(ngs, size=npts)
    X = np.ascontiguousarray(phi_s[:, ipx])
    t0 = time.perf_counter()
    A = X.T.dot(X) ** 2
    t_A = time.perf_counter() - t0
    A[np.diag_indices_from(A)] += 2.0 * np.abs(A).sum(axis=1).max()      # slab columns repeat: make it factorisable (not timed)
    t0 = time.perf_counter()
    cfac = scipy.linalg.cholesky(A, lower=True)
    t_chol = time.perf_counter() - t0
    del A
    sl3 = slice(0, max(1, ngs // 4))
    n3 = len(range(*sl3.indices(ngs)))
    t0 = time.perf_counter()
    B = X.T.dot(phi_s[:, sl3]) ** 2
    y = scipy.linalg.solve_triangular(cfac, B, lower=True)
    zeta_s = scipy.linalg.solve_triangular(cfac.T, y, lower=False)
    t['fit'] = t_A + t_chol + (time.perf_counter() - t0) * (ng / n3)
    del cfac, B, y
    # stage 4: FFT + coulG + inverse FFT of 16 aux functions on the whole mesh; metric product on a slab
    nvec = min(16, npts)
    zfull = rng.standard_normal((nvec, ng))
    t0 = time.perf_counter()
    isdf_ref.build_V(zfull, a, mesh)
    t['fft'] = (time.perf_counter() - t0) * (npts / nvec)
    del zfull
    V_s = rng.standard_normal(zeta_s.shape)
    t0 = time.perf_counter()
    W = (cell.vol / ng) * zeta_s.dot(V_s.T)
    t['wgemm'] = (time.perf_counter() - t0) * (ng / n3)
    del zeta_s, V_s
    # stage 5 in full
    W = W + W.T
    dm = synthetic_dm(nao)
    t0 = time.perf_counter()
    oracle.isdf_get_jk(X, W, dm)
    t['jk'] = time.perf_counter() - t0
    desc = ('NumPy/SciPy oracle on slabs of the workload itself: AO on ng/64 (%d points), selection = panel-blocked form on the host BLAS '
            '(pair-density + flush dgemm of a 128-pivot panel at k0=Np/8..7Np/8 on ng/16 = %d points, scaled); the unblocked GEMV loop is '
            'timed beside it (select_unblocked, not part of the total), full A=(X^T X)^2 + Cholesky (Np=%d), triangular solves '
            'and metric product on ng/256 (%d points), FFT+coulG+iFFT of %d aux functions, one full J/K; '
            'other slab times multiplied by the slab count' % (ngs, ngsel, npts, n3, nvec))
    return t, desc


def cpu_full_run(cell, npts):
    """Small workloads (C8 and below): every stage of the NumPy/SciPy oracle in full, timed as it runs."""
    import oracle
    from oracle import isdf_ref
    from pyscf_isdf_b200.workloads import synthetic_dm
    t = {}
    rcut, Ls = cell.ao_eval_params()
    a = cell.lattice_vectors()
    t0 = time.perf_counter()
    phi = np.ascontiguousarray(oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut).T)
    t['ao'] = time.perf_counter() - t0
    from oracle import c_oracle
    t0 = time.perf_counter()
    ip, resid = c_oracle.select_ip_blocked(phi, npts)       # panel-blocked, OpenMP (bit-equal to the loop below)
    t['select'] = time.perf_counter() - t0
    if float(phi.shape[1]) * npts * npts <= 2e10:           # the unblocked NumPy loop beside it while that is cheap
        t0 = time.perf_counter()
        isdf_ref.select_ip(phi, npts)
        t['select_unblocked'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    zeta = isdf_ref.fit_aux(phi, ip)
    t['fit'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    V = isdf_ref.build_V(zeta, a, cell.mesh)
    t['fft'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    W = isdf_ref.build_W(zeta, a, cell.mesh, V=V)
    t['wgemm'] = time.perf_counter() - t0
    X = np.ascontiguousarray(phi[:, ip])
    t0 = time.perf_counter()
    oracle.isdf_get_jk(X, W, synthetic_dm(cell.nao))
    t['jk'] = time.perf_counter() - t0
    return t, 'the whole workload, every stage of the oracle in full (no sampling, no extrapolation)'


def cpu_baseline_run(cell, npts):
    """(total s, build s, J/K s, per-stage s, description, cores) of one bounded CPU sample."""
    from oracle import pbc_tools
    ncores = use_all_host_threads()
    pbc_tools.FFT_WORKERS = ncores          # BLAS uses every core; give the FFTs the same
    # the sample holds [Np, ng/16] and [Np, Np] arrays: above the C64 size it is taken at Np = 12480
    # and the Np-dependence of each stage's work model is applied on top (said in the description)
    nps = npts if npts <= 16384 else 12480
    try:
        ng = int(np.prod(cell.mesh))
        if float(ng) * npts * npts <= 2e11:
            t, desc = cpu_full_run(cell, npts)       # small workloads: the whole oracle, no sampling
        else:
            t, desc = cpu_sample_target(cell, nps)
    finally:
        pbc_tools.FFT_WORKERS = 1
    if nps != npts:
        r = npts / nps
        power = {'ao': 0, 'fft': 1}
        t = {k: v * r ** power.get(k, 2) for k, v in t.items()}
        desc += '; sampled at Np=%d and scaled to Np=%d by each stage\'s leading power of Np (EXTRAPOLATED in Np)' % (nps, npts)
    build = t['ao'] + t['select'] + t['fit'] + t['fft'] + t['wgemm']
    return build + t['jk'], build, t['jk'], t, desc, ncores


def target_dims(args):
    from pyscf_isdf_b200 import workloads
    name = args.workload or 'c64'
    cell = workloads.make_cell(name)
    nao = cell.nao
    ng = int(np.prod(cell.mesh))
    npts = min(int(args.c * nao), ng)
    return name, cell, nao, ng, npts


def run_reference(args):
    """CPU arm: the oracle port on the host cores.  Under torchrun only rank 0 works."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    from pyscf_isdf_b200 import workloads
    name, cell, nao, ng, npts = target_dims(args)
    for _ in range(args.warmup):
        cpu_baseline_run(cell, npts)
    tot = []
    last = None
    for _ in range(args.steps):
        v, b, j, per, desc, ncores = cpu_baseline_run(cell, npts)
        tot.append(v)
        last = (desc, b, j, per)
    val = float(np.mean(tot))
    sample = last[0] + ' (%s, nao=%d Ng=%d Np=%d)' % (name, nao, ng, npts)
    sampled = float(ng) * npts * npts > 2e11
    check = None
    if sampled and args.gpus == 1:
        # error bar of the slab extrapolation: BASELINE configs[0] (C8) timed in full AND through the same
        # sampling code; ratio = sampled estimate / measured, per stage
        c8 = workloads.make_cell('c8')
        n8 = int(args.c * c8.nao)
        full, _ = cpu_full_run(c8, n8)
        est, _ = cpu_sample_target(c8, n8)
        check = {'workload': workloads.describe('c8'), 'measured_s': full, 'sampled_estimate_s': est,
                 'ratio': {k: (est[k] / full[k] if full.get(k) else None) for k in est if k in full}}
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 's', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': val * 1e3,
        'higher_is_better': False, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': {'workload': workloads.describe(name), 'c': args.c, 'nao': nao, 'ngrids': ng, 'n_ip': npts},
        'build_s': last[1], 'jk_s': last[2], 'stages_s': last[3],
        # a step of this arm is a BOUNDED SAMPLE: for workloads above C8 `value` is the slab times scaled to the
        # whole workload (it does not fit in this run's wall time by construction); selection enters as the
        # panel-blocked BLAS form, the unblocked loop is listed beside it in stages_s
        'extrapolated': sampled, 'extrapolation_check': check,
        'cpu_baseline': {'value': val, 'unit': 's', 'cores': ncores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': val, 'unit': 's', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------- GPU arm
def measure_fp64_peak(torch, dev, n=8192, reps=10):
    """cuBLAS DGEMM n^3 (torch.matmul, best of reps) -- the FP64 roofline
    denominator; MEASURED_PEAKS.json carries no FP64 entry."""
    a = torch.randn((n, n), dtype=torch.float64, device=dev)
    b = torch.randn((n, n), dtype=torch.float64, device=dev)
    c = torch.empty_like(a)
    for _ in range(2):
        torch.matmul(a, b, out=c)
    best = 1e30
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b, out=c); e1.record()
        torch.cuda.synchronize(dev)
        best = min(best, e0.elapsed_time(e1))
    del a, b, c
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def parity_block(device):
    """BASELINE configs[0] (C8, mesh 35^3, c = 15) through the product path against the CPU oracle, outside
    every timed region: interpolation points bit-equal to the C oracle fed the device's AO values; J/K of a
    synthetic density matrix against the oracle's own build (Cholesky fit, NumPy FFT)."""
    import oracle
    from oracle import c_oracle, isdf_ref
    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    cell = workloads.make_cell('c8')
    nao = cell.nao
    mydf = ISDF(cell, c=15, device=device).build()
    dm = workloads.synthetic_dm(nao)
    vj, vk = mydf.get_jk(dm, exxdiv=None)
    aoR = mydf.aoR_host()
    ip, resid = c_oracle.select_ip_blocked(aoR, int(15 * nao))
    rcut, Ls = cell.ao_eval_params()
    phi = np.ascontiguousarray(oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut).T)
    W = isdf_ref.build_W(isdf_ref.fit_aux(phi, ip), cell.lattice_vectors(), cell.mesh)
    vj0, vk0 = oracle.isdf_get_jk(np.ascontiguousarray(phi[:, ip]), W, dm)
    return {'config': workloads.describe('c8') + ', c=15', 'n_ip': int(mydf.n_ip),
            'ip_equal': bool(np.array_equal(ip, mydf.ip_index)),
            'resid_equal': bool(np.array_equal(resid, mydf.ip_resid)),
            'max_dAO': float(abs(aoR - phi).max()), 'max_dJ': float(abs(vj - vj0).max()),
            'max_dK': float(abs(vk - vk0).max()), 'tolerance': 1e-8,
            'oracle': 'oracle/c/isdf_oracle_blocked.c (points), oracle/isdf_ref.py (fit, W, J/K); parity unpinned: no ISDF in the reference'}


def extra_workload(name, args, local, comm, world, torch, dist):
    """One build + J/K of a larger BASELINE config on the same ranks (not part of `value`): 1 warm-up build,
    1 timed build, 5 J/K; stage times and the FP64 GEMM rate of the timed build."""
    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    dev = torch.device('cuda', local)
    try:
        cell = workloads.make_cell(name)
        nao = cell.nao
        mydf = ISDF(cell, c=args.c, fit=args.fit, device=local, keep_aoR=False, comm=comm, wgemm=args.wgemm)
        h = mydf.handle
        # does it fit these ranks?  Decided by ALL ranks before anything is allocated: a rank that ran out of
        # memory alone would leave the others waiting in a collective.
        from pyscf_isdf_b200.dist import SlabPartition
        from pyscf_isdf_b200 import _lib
        ng = int(np.prod(cell.mesh))
        npts = min(int(args.c * nao), ng)
        rank = comm.rank if comm is not None else 0
        part = SlabPartition(cell.mesh, world, npts)
        g0, g1 = part.grid_range(rank)
        ldl = _lib.pad8(g1 - g0)
        nbig = max(npts * ldl, 2 * npts * int(cell.mesh[0]) * part.ncol_local(rank))
        ws_w = int(h.lib.isdf_syrk_i8_ws_bytes(npts, 2 * int(cell.mesh[0]) * part.ncol_local(rank), 8)) if args.wgemm.startswith('i8') else 0
        need = 8 * nbig + 8 * (nao * ldl + 3 * npts * _lib.pad8(npts)) + ws_w + (3 << 30)
        free = torch.cuda.mem_get_info(dev)[0]
        fits = torch.tensor([1.0 if need <= 0.95 * free else 0.0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(fits, op=dist.ReduceOp.MIN)
        if float(fits.item()) < 0.5:
            return {'workload': workloads.describe(name), 'n_gpus': world,
                    'skipped': 'needs %.0f GB per rank on %d GPU(s), %.0f GB free' % (need / 1e9, world, free / 1e9)}
        d_dm = torch.from_numpy(workloads.synthetic_dm(nao)[None]).to(dev)
        mydf.build()
        mydf.reset()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        h.gemm_profile(True)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); mydf.build(); e1.record()
        torch.cuda.synchronize(dev)
        gflops, gms, gcount = h.gemm_profile_read()
        h.gemm_profile(False)
        build_ms = e0.elapsed_time(e1)
        jk = []
        for _ in range(6):
            a0 = torch.cuda.Event(enable_timing=True); a1 = torch.cuda.Event(enable_timing=True)
            a0.record(); mydf.get_jk_device(d_dm, True, True, None); a1.record()
            torch.cuda.synchronize(dev)
            jk.append(a0.elapsed_time(a1))
        tt = torch.tensor([build_ms, float(np.median(jk[1:]))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        out = {'workload': workloads.describe(name), 'nao': nao, 'ngrids': int(np.prod(cell.mesh)), 'n_ip': int(mydf.n_ip),
               'n_gpus': world, 'build_s': float(tt[0]) / 1e3, 'jk_s': float(tt[1]) / 1e3,
               'value': (float(tt[0]) + float(tt[1])) / 1e3, 'unit': 's', 'timed': '1 warm-up build, 1 timed build, median of 5 J/K; max over ranks',
               'stages_ms': {k: float(v) for k, v in mydf.timings.items() if k.endswith('_ms')},
               'fp64_gemm_tflops_rank0': gflops / (gms * 1e-3) / 1e12 if gms > 0 else None,
               'aliased_build_buffers': bool(getattr(mydf, 'timings_flags', {}).get('aliased_arenas', False))}
        mydf.release_buffers()
        del mydf
        torch.cuda.empty_cache()
        return out
    except Exception as e:                    # noqa: BLE001 -- an extra must never take the headline line down
        torch.cuda.empty_cache()
        return {'workload': name, 'error': '%s: %s' % (type(e).__name__, str(e)[:300])}


def host_stats():
    """Host-side facts that explain GPU idle time inside a step: CFS throttling of the container and involuntary
    context switches of this process."""
    import resource
    out = {}
    try:
        for line in open('/sys/fs/cgroup/cpu.stat'):
            k, v = line.split()
            if k in ('nr_throttled', 'throttled_usec', 'nr_periods'):
                out[k] = int(v)
        out['cpu_max'] = open('/sys/fs/cgroup/cpu.max').read().strip()
    except (OSError, ValueError):
        pass
    ru = resource.getrusage(resource.RUSAGE_SELF)
    out['nvcsw'], out['nivcsw'] = ru.ru_nvcsw, ru.ru_nivcsw
    out['utime_s'], out['stime_s'] = ru.ru_utime, ru.ru_stime
    return out


def kernel_traffic(kernel):
    """DRAM bytes per launch of the roofline kernel, read from the committed ncu summary of this round."""
    path = os.path.join(ROOT, 'profiles', 'kernel_traffic.json')
    try:
        with open(path) as f:
            return json.load(f).get(kernel)
    except (OSError, ValueError):
        return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a B200: the ISDF hot path has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        import datetime
        # short watchdog: a mismatched collective must fail fast, not burn GPU time
        dist.init_process_group('nccl', device_id=dev, timeout=datetime.timedelta(seconds=180))

    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    name, cell, nao, ng, npts = target_dims(args)
    comm = None
    if world > 1:
        from pyscf_isdf_b200.dist import Comm
        comm = Comm()                 # grid sharded by x-planes, one rank per GPU (pyscf_isdf_b200/dist.py)
    exxdiv = None if args.exxdiv == 'none' else args.exxdiv
    mydf = ISDF(cell, c=args.c, fit=args.fit, device=local, keep_aoR=(exxdiv == 'ewald'), comm=comm, wgemm=args.wgemm)
    h = mydf.handle
    dm_host = workloads.synthetic_dm(nao)[None]
    dm_pinned = torch.from_numpy(dm_host).pin_memory()
    d_dm = dm_pinned.to(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_device():
        mydf.reset()
        mydf.build()
        return mydf.get_jk_device(d_dm, True, True, exxdiv)

    def step_e2e():
        mydf.reset()
        mydf.build()
        return mydf.get_jk(dm_host[0], exxdiv=exxdiv)

    fp64_peak = measure_fp64_peak(torch, dev) if rank == 0 else None
    if os.environ.get('ISDF_BENCH_GC') == '0':          # diagnostics: is the interpreter's garbage collector behind outlier steps?
        import gc
        gc.collect()
        gc.disable()

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local).start() if rank == 0 else None
    l0 = h.launches()
    h.gemm_profile(True)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    stage_acc = {}
    marks = []
    host0 = host_stats()
    wall = [time.time()]
    e0.record()
    for _ in range(args.steps):
        step_device()
        wall.append(time.time())       # the build ends with a device synchronisation: host time tracks device time to ~20 ms
        for k, v in mydf.timings.items():
            stage_acc.setdefault(k, []).append(v)
        m = torch.cuda.Event(enable_timing=True)
        m.record()
        marks.append(m)
    e1.record()
    barrier()
    total_ms = e0.elapsed_time(e1)
    host1 = host_stats()
    host_delta = {k: (host1[k] - host0[k] if isinstance(host1.get(k), (int, float)) and k in host0 else host1.get(k)) for k in host1}
    host_delta['wall_s'] = wall[-1] - wall[0]
    each_ms = [a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks)]
    gflops, gms, gcount = h.gemm_profile_read()
    h.gemm_profile(False)
    l1 = h.launches()
    clocks = sampler.stop() if sampler else None
    if clocks is not None and sampler.rows:
        clocks['sm_mhz_mean_each_step'] = sampler.mean_clock_between(wall)

    # J/K alone (per SCF iteration), device-resident
    jk_ms = []
    for _ in range(args.jk_reps + 2):
        a0 = torch.cuda.Event(enable_timing=True); a1 = torch.cuda.Event(enable_timing=True)
        a0.record(); mydf.get_jk_device(d_dm, True, True, exxdiv); a1.record()
        torch.cuda.synchronize(dev)
        jk_ms.append(a0.elapsed_time(a1))
    jk_ms = float(np.median(jk_ms[2:]))

    # exchange-correlation per SCF iteration (optional): first call evaluates the AO gradients
    xc_info = None
    if args.xc:
        t0 = time.perf_counter()
        mydf.nr_rks(args.xc, dm_host[0])
        torch.cuda.synchronize(dev)
        first_s = time.perf_counter() - t0
        xs = []
        for _ in range(args.jk_reps):
            t0 = time.perf_counter()
            ne, exc, vxc = mydf.nr_rks(args.xc, dm_host[0])
            xs.append(time.perf_counter() - t0)
        xc_info = {'functional': args.xc, 'ms': float(np.median(xs)) * 1e3, 'first_call_ms': first_s * 1e3,
                   'nelec': ne, 'exc': exc, 'timed': 'host API, wall clock, includes H2D of dm and D2H of vmat'}
        mydf._aoR4 = None

    # end to end through the plugin's host API
    if args.no_e2e:
        vj, vk = mydf.get_jk(dm_host[0], exxdiv=exxdiv)
        e2e_s = float('nan')
        clocks_e2e = None
        e2e_each = None
    else:
        step_e2e()
        barrier()
        sampler2 = ClockSampler(local).start() if rank == 0 else None
        t0 = time.perf_counter()
        e2e_marks = [t0]
        for _ in range(args.steps):                  # all K steps, every one a full build + J/K through host buffers
            vj, vk = step_e2e()
            e2e_marks.append(time.perf_counter())
        barrier()
        e2e_s = (time.perf_counter() - t0) / args.steps
        e2e_each = [b - a for a, b in zip(e2e_marks[:-1], e2e_marks[1:])]     # rank 0's view; `value` is the mean over all
        clocks_e2e = sampler2.stop() if sampler2 else None
    # bytes that cross the PCIe link per step: the cell tables and lattice images isdf_eval_ao takes as host
    # arrays, the density matrix through the object's pinned staging buffer (counted where it is copied,
    # ISDF.get_jk) on the way in; J, K (counted likewise) and the selected points + residuals on the way out
    rcut, Ls = cell.ao_eval_params()
    h2d = int(cell._atm.nbytes + cell._bas.nbytes + cell._env.nbytes + np.asarray(Ls).nbytes +
              np.asarray(rcut).nbytes + mydf.h2d_bytes)
    d2h = int(mydf.d2h_bytes + 16 * mydf.n_ip)

    tt = torch.tensor([total_ms, jk_ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    total_ms, jk_ms, e2e_s = [float(x) for x in tt.tolist()]

    # one-off runs of the larger BASELINE configs on the same ranks (never part of `value`): the north_star
    # target Si216 needs the 8-GPU sharding (its factor alone is 424 GB), the water box fits from 4 GPUs up
    extra = {}
    if not args.no_extra and name == 'c64':
        headline = (mydf.n_ip, mydf.timings)
        mydf.reset(); mydf.release_buffers()
        torch.cuda.empty_cache()
        # config 3 fits one GPU since round 2 (the spectra overwrite the factor rows: one 80 GB arena instead of two)
        extra['h2o32'] = extra_workload('h2o32', args, local, comm, world, torch, dist)
        if world >= 8:
            extra['si216'] = extra_workload('si216', args, local, comm, world, torch, dist)
        mydf.n_ip = headline[0]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    step_ms = total_ms / args.steps
    stages_ms = {k: float(np.median(v)) for k, v in stage_acc.items() if k.endswith('_ms')}
    build_s = step_ms / 1e3 - jk_ms / 1e3
    fm = workloads.flop_model(nao, ng, mydf.n_ip, cell.mesh, args.fit)
    achieved = gflops / (gms * 1e-3) / 1e12 if gms > 0 else 0.0
    roofline = {
        'kernel': 'dgemm_tma_kernel (FP64 DMMA GEMM fed by 2-D tensor-map TMA, all layouts/epilogues; '
                  'dgemm_kernel for operands TMA cannot address)', 'bound': 'tensor',
        'achieved': achieved, 'peak': fp64_peak, 'unit': 'TFLOP/s',
        'frac': achieved / fp64_peak if fp64_peak else None,
        'peak_source': 'cuBLAS DGEMM 8192^3 best-of-10 measured in this run (MEASURED_PEAKS.json has no FP64 entry)',
        'launches_per_step': gcount / args.steps, 'flops_per_step': gflops / args.steps,
        'gemm_ms_per_step': gms / args.steps, 'gemm_share_of_step': gms / total_ms,
        # DRAM bytes of one representative launch of this kernel from an `ncu --set full` capture
        # (flush GEMM M=128 N=357912 K=6240: 19.1 GB against 18.6 GB algorithmic, profiles/r02_gemm_tma.md)
    }
    tr = kernel_traffic('dgemm_tma_kernel') if name == 'c64' else None
    roofline['traffic'] = tr['dram_bytes'] if tr else None
    roofline['traffic_launch'] = ('%s: algorithmic %.3g B (%s)' % (tr['launch'], tr['algorithmic_bytes'], tr['source'])) if tr else None
    line = {
        'metric': METRIC, 'value': step_ms / 1e3, 'unit': 's', 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': step_ms,
        'higher_is_better': False, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': {'workload': workloads.describe(name), 'c': args.c, 'nao': nao, 'ngrids': ng,
                   'n_ip': int(mydf.n_ip), 'fit': args.fit, 'wgemm': args.wgemm, 'l2': 'inputs_exceed_l2' if 8.0 * ng * nao > 126e6 else 'inputs_fit_l2',
                   'exxdiv': args.exxdiv,
                   'parallelism': 'single' if world == 1 else 'grid-slab x%d (all-gather/all-to-all/all-reduce over NCCL)' % world},
        'build_s': build_s, 'jk_s': jk_ms / 1e3, 'stages_ms': stages_ms,
        'each_step_ms': each_ms,
        'stages_ms_each': {k: [float(x) for x in v] for k, v in stage_acc.items() if k.endswith('_ms')},
        'flop_model': fm,
        'e2e': {'value': e2e_s, 'unit': 's', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'sm_mhz_mean': (clocks_e2e or {}).get('sm_mhz_mean'), 'sm_mhz_median': (clocks_e2e or {}).get('sm_mhz'),
                'each_step_s': e2e_each},
        'gpu_launches': int((l1[0] - l0[0]) + (l1[1] - l0[1])),
        'roofline': roofline, 'clocks': clocks, 'host': host_delta,
    }
    if args.wgemm.startswith('i8') and stages_ms.get('wgemm_ms'):
        planes = int(args.wgemm.split(':')[1]) if ':' in args.wgemm else 8
        ops = float(getattr(h, 'syrk_i8_ops', 0.0))         # int8 ops issued by this rank's last W product
        crt = os.environ.get('ISDF_SYRK_CRT')
        try:
            crt = None if crt is None else str(int(crt))
        except ValueError:
            crt = '0'                                        # the library reads it with atoi()
        form = ('digit planes: S(S+1)/2 int8 products per K chunk' if crt is not None and int(crt) <= 0 else
                'residues modulo %d coprime moduli <= 256 (one int8 product per modulus, Garner reconstruction)'
                % (16 if crt is None else max(8, min(16, int(crt)))))
        line['wgemm_i8'] = {'kernel': 'syrk_i8_kernel2 (tcgen05 kind::i8, TMEM accumulators, bulk-copy ring, cluster multicast)',
                            'form': form, 'max_planes': planes, 'int8_ops_per_rank': ops, 'ms': stages_ms['wgemm_ms'],
                            'achieved_pops': ops / (stages_ms['wgemm_ms'] * 1e-3) / 1e15,
                            'fp64_equivalent_tflops': fm['coulomb_gemm'] / world / (stages_ms['wgemm_ms'] * 1e-3) / 1e12}
    if xc_info is not None:
        line['xc'] = xc_info
    if extra:
        line['extra'] = extra
    if world == 1 and not args.no_parity:
        try:
            line['parity'] = parity_block(local)
        except Exception as e:                # noqa: BLE001
            line['parity'] = {'error': '%s: %s' % (type(e).__name__, str(e)[:300])}
    if world == 1 and not args.no_cpu_baseline:
        v, b, j, per, desc, ncores = cpu_baseline_run(cell, int(mydf.n_ip))
        line['cpu_baseline'] = {'value': v, 'unit': 's', 'cores': ncores, 'kind': 'port', 'sample': desc,
                                'build_s': b, 'jk_s': j, 'stages_s': per}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == 'reference':
        return run_reference(args)
    return run_b200(args)


if __name__ == '__main__':
    sys.exit(main())
