#!/usr/bin/env python
"""Parity of the benchmarked configuration itself (BASELINE configs[1]: C64, mesh 71^3, c = 15, Np = 12480)
against independent arithmetic.  Run once per round on the GPU box; writes one JSON line
(commit it under profiles/).  Three checks:

  1. int8 residue product (the default, `wgemm='i8'`) against the FP64 tensor-pipe product (`wgemm='fp64'`) and
     the digit-plane form (ISDF_SYRK_CRT=0) in one process: same interpolation points, max|dW|, max|dJ|, max|dK|.
  2. interpolation points against the panel-blocked OpenMP C oracle (oracle/c/isdf_oracle_blocked.c, bit-equal to
     the textbook loop) fed the device's AO matrix: the first `--oracle-points` pivots (all 12480 by default)
     and their residuals, bit for bit.  63 TFLOP of FMA chains on the host cores.
  3. a 16 x 16 block of W against the oracle's own route (Cholesky of A = (X^T X)^2, zeta = A^-1 B, FFT, coulG,
     quadrature), computed for 16 rows only so that it stays cheap.

Test infrastructure: this is the only tool that imports oracle/."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='c64')
    ap.add_argument('--c', type=float, default=15.0)
    ap.add_argument('--oracle-points', type=int, default=-1, help='pivots the CPU oracle selects (-1: all, 0: skip)')
    ap.add_argument('--rows', type=int, default=16, help='rows of W recomputed by the oracle route')
    ap.add_argument('--skip-planes', action='store_true')
    args = ap.parse_args()
    import torch
    import scipy.linalg
    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    from oracle import c_oracle, pbc_tools
    from oracle.isdf_ref import build_V

    cell = workloads.make_cell(args.workload)
    nao = cell.nao
    ng = int(np.prod(cell.mesh))
    dm = workloads.synthetic_dm(nao)
    out = {'workload': workloads.describe(args.workload), 'c': args.c, 'nao': nao, 'ngrids': ng,
           'host_cores': os.cpu_count()}

    # ---- 1. the three forms of the metric product
    res = {}
    forms = [('i8_crt', 'i8', None), ('fp64', 'fp64', None)] + ([] if args.skip_planes else [('i8_planes', 'i8', '0')])
    for name, wgemm, crt in forms:
        if crt is None:
            os.environ.pop('ISDF_SYRK_CRT', None)
        else:
            os.environ['ISDF_SYRK_CRT'] = crt
        mydf = ISDF(cell, c=args.c, wgemm=wgemm, keep_aoR=(name == 'i8_crt'))
        mydf.build()
        vj, vk = mydf.get_jk(dm, exxdiv=None)
        res[name] = dict(W=mydf.W.cpu().numpy(), vj=vj, vk=vk, ip=np.array(mydf.ip_index), resid=np.array(mydf.ip_resid))
        out[name + '_wgemm_ms'] = mydf.timings.get('wgemm_ms')
        if name == 'i8_crt':
            aoR = mydf.aoR_host()
            npts = int(mydf.n_ip)
        mydf.release_buffers()
        del mydf
        torch.cuda.empty_cache()
    os.environ.pop('ISDF_SYRK_CRT', None)
    ref = res['fp64']
    out['n_ip'] = npts
    out['W_max_abs'] = float(abs(ref['W']).max())
    out['vj_max_abs'], out['vk_max_abs'] = float(abs(ref['vj']).max()), float(abs(ref['vk']).max())
    for name in res:
        if name == 'fp64':
            continue
        out[name + '_vs_fp64'] = {
            'same_points': bool(np.array_equal(res[name]['ip'], ref['ip'])),
            'max_dW': float(abs(res[name]['W'] - ref['W']).max()),
            'max_dJ': float(abs(res[name]['vj'] - ref['vj']).max()),
            'max_dK': float(abs(res[name]['vk'] - ref['vk']).max())}
    print(json.dumps(out), flush=True)

    # ---- 2. the interpolation points against the blocked C oracle on the device's AO matrix
    gpu = res['i8_crt']
    nsel = npts if args.oracle_points < 0 else min(args.oracle_points, npts)
    if nsel > 0:
        t0 = time.perf_counter()
        ip, resid = c_oracle.select_ip_blocked(aoR, nsel)
        dt = time.perf_counter() - t0
        out['oracle_select'] = {
            'points': int(nsel), 'seconds': dt, 'tflop': (ng * float(nsel) ** 2 + 2.0 * ng * nao * nsel) / 1e12,
            'indices_bit_equal': bool(np.array_equal(ip, gpu['ip'][:nsel])),
            'residuals_bit_equal': bool(np.array_equal(resid, gpu['resid'][:nsel])),
            'first_mismatch': int(np.argmax(ip != gpu['ip'][:len(ip)])) if not np.array_equal(ip, gpu['ip'][:nsel]) else None}
        print(json.dumps(out), flush=True)

    # ---- 3. rows of W by the oracle's route (Cholesky of A, zeta = A^-1 B), on the device's points
    if args.rows > 0:
        t0 = time.perf_counter()
        ipg = gpu['ip']
        X = np.ascontiguousarray(aoR[:, ipg])
        A = X.T.dot(X) ** 2
        cf = scipy.linalg.cho_factor(A, lower=True)
        del A
        rows = np.linspace(0, npts - 1, args.rows).astype(int)
        E = np.zeros((npts, len(rows)))
        E[rows, np.arange(len(rows))] = 1.0
        Ainv_rows = scipy.linalg.cho_solve(cf, E).T                      # [rows, Np] = rows of A^-1 (A symmetric)
        del cf
        zeta = np.empty((len(rows), ng))
        step = 16384
        for g0 in range(0, ng, step):
            B = X.T.dot(aoR[:, g0:g0 + step]) ** 2
            zeta[:, g0:g0 + step] = Ainv_rows.dot(B)
        pbc_tools.FFT_WORKERS = os.cpu_count() or 1
        V = build_V(zeta, cell.lattice_vectors(), cell.mesh)
        Wblk = (cell.vol / ng) * zeta.dot(V.T)
        Wg = ref['W'][np.ix_(rows, rows)]
        Wi = gpu['W'][np.ix_(rows, rows)]
        out['oracle_W_block'] = {'rows': [int(r) for r in rows], 'seconds': time.perf_counter() - t0,
                                 'max_abs': float(abs(Wblk).max()),
                                 'max_d_fp64': float(abs(Wg - Wblk).max()), 'max_d_i8': float(abs(Wi - Wblk).max())}
    print(json.dumps(out), flush=True)
    return 0


if __name__ == '__main__':
    sys.exit(main())
