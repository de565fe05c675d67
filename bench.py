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

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                 '--format=csv,noheader,nounits', '-lms', '200'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

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
        return {'sm_mhz': float(np.median(load)), 'sm_max_mhz': float(max(mx)),
                'power_w_max': float(max(pw)), 'samples': len(sm), 'reasons': sorted(reasons)}


# --------------------------------------------------------------------------- CPU arm
def cpu_sample_step(nthreads):
    """One bounded sample of the CPU port (the NumPy/SciPy oracle): the C8 cell
    of BASELINE configs[0] at mesh 35^3, c = 8, AO values on 1/4 of the grid.
    Returns measured per-stage seconds and the sample's algorithmic flop counts."""
    import oracle
    from pyscf_isdf_b200 import workloads
    cell = workloads.make_cell('c8')
    nao = cell.nao
    mesh = list(cell.mesh)
    ng = int(np.prod(mesh))
    npts = 8 * nao
    rcut, Ls = cell.ao_eval_params()
    coords = cell.get_uniform_grids()
    t = {}
    t0 = time.perf_counter()
    part = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords[:ng // 4], Ls, rcut)
    t['ao'] = (time.perf_counter() - t0) * 4.0          # scaled to the whole grid
    rest = oracle.eval_ao(cell._atm, cell._bas, cell._env, coords[ng // 4:], Ls, rcut)
    phi = np.ascontiguousarray(np.vstack([part, rest]).T)
    t0 = time.perf_counter()
    ip, _ = oracle.select_ip(phi, npts)
    t['select'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    zeta = oracle.fit_aux(phi, ip)
    t['fit'] = time.perf_counter() - t0
    a = cell.lattice_vectors()
    t0 = time.perf_counter()
    V = oracle.build_V(zeta, a, mesh)
    t['fft'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    W = oracle.build_W(zeta, a, mesh, V=V)
    t['wgemm'] = time.perf_counter() - t0
    from pyscf_isdf_b200.workloads import synthetic_dm
    dm = synthetic_dm(nao)
    X = np.ascontiguousarray(phi[:, ip])
    t0 = time.perf_counter()
    for _ in range(5):
        oracle.isdf_get_jk(X, W, dm)
    t['jk'] = (time.perf_counter() - t0) / 5
    dims = dict(nao=nao, ng=ng, npts=len(ip), mesh=mesh)
    return t, dims


def cpu_work(nao, ng, npts):
    """Work units of the CPU port per stage (what its time scales with)."""
    return {
        'ao': 1.0 * ng * nao,
        'select': 2.0 * ng * nao * npts + 1.0 * ng * npts ** 2,
        'fit': 2.0 * npts * nao * ng + 2.0 * npts ** 2 * nao + npts ** 3 / 3.0 + 2.0 * npts ** 2 * ng,
        'fft': 2.0 * 5.0 * npts * ng * np.log2(ng),
        'wgemm': 2.0 * npts ** 2 * ng,
        'jk': 4.0 * npts ** 2 * nao + 6.0 * npts * nao ** 2 + 2.0 * npts ** 2,
    }


def cpu_extrapolate(t, dims, nao, ng, npts):
    """Scale the sample's per-stage times to the target workload by the stage's work model."""
    ws = cpu_work(dims['nao'], dims['ng'], dims['npts'])
    wt = cpu_work(nao, ng, npts)
    per = {k: t[k] * wt[k] / ws[k] for k in t}
    build = per['ao'] + per['select'] + per['fit'] + per['fft'] + per['wgemm']
    return build + per['jk'], build, per['jk'], per


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
    ncores = os.cpu_count() or 1
    from pyscf_isdf_b200 import workloads
    name, cell, nao, ng, npts = target_dims(args)
    for _ in range(args.warmup):
        cpu_sample_step(ncores)
    tot = []
    last = None
    for _ in range(args.steps):
        t, dims = cpu_sample_step(ncores)
        v, b, j, per = cpu_extrapolate(t, dims, nao, ng, npts)
        tot.append(v)
        last = (t, b, j, per)
    val = float(np.mean(tot))
    sample = ('C8 diamond mesh 35^3 c=8 (AO on 1/4 grid) through the NumPy/SciPy oracle: measured %.1f s; '
              'per-stage times scaled by the stage work model to %s nao=%d Ng=%d Np=%d (EXTRAPOLATED)'
              % (sum(last[0].values()), name, nao, ng, npts))
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 's', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': val * 1e3,
        'higher_is_better': False, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': {'workload': workloads.describe(name), 'c': args.c, 'nao': nao, 'ngrids': ng, 'n_ip': npts},
        'build_s': last[1], 'jk_s': last[2], 'stages_s': last[3],
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
    mydf = ISDF(cell, c=args.c, fit=args.fit, device=local, keep_aoR=False, comm=comm, wgemm=args.wgemm)
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
        return mydf.get_jk_device(d_dm, True, True, None)

    def step_e2e():
        mydf.reset()
        mydf.build()
        return mydf.get_jk(dm_host[0], exxdiv=None)

    fp64_peak = measure_fp64_peak(torch, dev) if rank == 0 else None

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local).start() if rank == 0 else None
    l0 = h.launches()
    h.gemm_profile(True)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    stage_acc = {}
    marks = []
    e0.record()
    for _ in range(args.steps):
        step_device()
        for k, v in mydf.timings.items():
            stage_acc.setdefault(k, []).append(v)
        m = torch.cuda.Event(enable_timing=True)
        m.record()
        marks.append(m)
    e1.record()
    barrier()
    total_ms = e0.elapsed_time(e1)
    each_ms = [a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks)]
    gflops, gms, gcount = h.gemm_profile_read()
    h.gemm_profile(False)
    l1 = h.launches()
    clocks = sampler.stop() if sampler else None

    # J/K alone (per SCF iteration), device-resident
    jk_ms = []
    for _ in range(args.jk_reps + 2):
        a0 = torch.cuda.Event(enable_timing=True); a1 = torch.cuda.Event(enable_timing=True)
        a0.record(); mydf.get_jk_device(d_dm, True, True, None); a1.record()
        torch.cuda.synchronize(dev)
        jk_ms.append(a0.elapsed_time(a1))
    jk_ms = float(np.median(jk_ms[2:]))

    # end to end through the plugin's host API
    if args.no_e2e:
        vj, vk = mydf.get_jk(dm_host[0], exxdiv=None)
        e2e_s = float('nan')
    else:
        step_e2e()
        barrier()
        t0 = time.perf_counter()
        e2e_steps = max(1, min(args.steps, 2))
        for _ in range(e2e_steps):
            vj, vk = step_e2e()
        barrier()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    rcut, Ls = cell.ao_eval_params()
    h2d = int(cell._atm.nbytes + cell._bas.nbytes + cell._env.nbytes + np.asarray(Ls).nbytes +
              np.asarray(rcut).nbytes + dm_host.nbytes + 8 * mydf.n_ip)
    d2h = int(vj.nbytes + vk.nbytes + 16 * mydf.n_ip)

    tt = torch.tensor([total_ms, jk_ms, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    total_ms, jk_ms, e2e_s = [float(x) for x in tt.tolist()]
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
        'kernel': 'dgemm_kernel (FP64 FMA GEMM, all layouts/epilogues)', 'bound': 'tensor',
        'achieved': achieved, 'peak': fp64_peak, 'unit': 'TFLOP/s',
        'frac': achieved / fp64_peak if fp64_peak else None,
        'peak_source': 'cuBLAS DGEMM 8192^3 best-of-10 measured in this run (MEASURED_PEAKS.json has no FP64 entry)',
        'launches_per_step': gcount / args.steps, 'flops_per_step': gflops / args.steps,
        'gemm_ms_per_step': gms / args.steps, 'gemm_share_of_step': gms / total_ms,
        # DRAM bytes of one representative launch of this kernel from an `ncu --set full` capture
        # (flush GEMM M=128 N=357912 K=6240: 19.3 GB against 18.6 GB algorithmic, profiles/r01_ncu_final.md)
        'traffic': 1.93e10 if name == 'c64' else None,
        'traffic_launch': 'C[128,357912] -= A^T B, K=6240 (selection flush): algorithmic 1.86e10 B' if name == 'c64' else None,
    }
    line = {
        'metric': METRIC, 'value': step_ms / 1e3, 'unit': 's', 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': step_ms,
        'higher_is_better': False, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64',
        'data': 'synthetic',
        'config': {'workload': workloads.describe(name), 'c': args.c, 'nao': nao, 'ngrids': ng,
                   'n_ip': int(mydf.n_ip), 'fit': args.fit, 'wgemm': args.wgemm, 'l2': 'inputs_exceed_l2' if 8.0 * ng * nao > 126e6 else 'inputs_fit_l2',
                   'parallelism': 'single' if world == 1 else 'grid-slab x%d (all-gather/all-to-all/all-reduce over NCCL)' % world},
        'build_s': build_s, 'jk_s': jk_ms / 1e3, 'stages_ms': stages_ms,
        'each_step_ms': each_ms,
        'stages_ms_each': {k: [float(x) for x in v] for k, v in stage_acc.items() if k.endswith('_ms')},
        'flop_model': fm,
        'e2e': {'value': e2e_s, 'unit': 's', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h},
        'gpu_launches': int((l1[0] - l0[0]) + (l1[1] - l0[1])),
        'roofline': roofline, 'clocks': clocks,
    }
    if args.wgemm.startswith('i8') and stages_ms.get('wgemm_ms'):
        planes = int(args.wgemm.split(':')[1]) if ':' in args.wgemm else 8
        ops = float(getattr(h, 'syrk_i8_ops', 0.0))         # int8 ops issued by this rank's last W product
        line['wgemm_i8'] = {'kernel': 'syrk_i8_kernel (tcgen05 kind::i8, TMEM accumulators, bulk-copy ring, cluster multicast)',
                            'max_planes': planes, 'int8_ops_per_rank': ops, 'ms': stages_ms['wgemm_ms'],
                            'achieved_pops': ops / (stages_ms['wgemm_ms'] * 1e-3) / 1e15,
                            'fp64_equivalent_tflops': fm['coulomb_gemm'] / world / (stages_ms['wgemm_ms'] * 1e-3) / 1e12}
    if world == 1 and not args.no_cpu_baseline:
        ncores = os.cpu_count() or 1
        t, dims = cpu_sample_step(ncores)
        v, b, j, per = cpu_extrapolate(t, dims, nao, ng, int(mydf.n_ip))
        line['cpu_baseline'] = {
            'value': v, 'unit': 's', 'cores': ncores, 'kind': 'port',
            'sample': 'C8 diamond mesh 35^3 c=8 (AO on 1/4 grid) through the NumPy/SciPy oracle: measured %.1f s; '
                      'per-stage times scaled by the stage work model to this workload (EXTRAPOLATED)' % sum(t.values()),
            'build_s': b, 'jk_s': j}
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
