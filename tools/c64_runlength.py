#!/usr/bin/env python
"""How many pivots does ONE candidate set certify when it is kept across panels with the stale bound
?  C64, the oracle's factor on the host: for quota q the candidate set is the union of
each of N slabs' q/N largest residuals; it is kept until the next pivot of the exact sequence is outside it or
not above theta (the largest residual outside the set WHEN IT WAS CHOSEN).  Prints run-length statistics.
Test infrastructure (imports oracle/)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    from oracle import c_oracle
    name = sys.argv[1] if len(sys.argv) > 1 else 'c64'
    cell = workloads.make_cell(name)
    mydf = ISDF(cell, c=15, keep_aoR=True).build()
    aoR = mydf.aoR_host()
    ip_gpu = np.array(mydf.ip_index)
    npts = int(mydf.n_ip)
    del mydf
    torch.cuda.empty_cache()
    t0 = time.perf_counter()
    ip, resid, Lt = c_oracle.select_ip_blocked(aoR, npts, return_factor=True)
    out = {'workload': name, 'oracle_seconds': time.perf_counter() - t0, 'same_points': bool(np.array_equal(ip, ip_gpu))}
    d0 = np.einsum('ig,ig->g', aoR, aoR) ** 2
    ng = d0.shape[0]
    nx = int(cell.mesh[0])
    nyz = ng // nx
    for nranks in (1, 8):
        bounds = [((r * nx) // nranks) * nyz for r in range(nranks + 1)]
        for quota in (1023, 2047):
            q = quota // nranks
            dd = d0.copy()
            k = 0
            runs = []
            while k < len(ip):
                C = []
                theta = -1.0
                for r in range(nranks):
                    seg = dd[bounds[r]:bounds[r + 1]]
                    part = np.argpartition(-seg, q)[:q + 1]
                    vals = seg[part]
                    srt = np.argsort(-vals)
                    C.append(part[srt[:q]] + bounds[r])
                    theta = max(theta, vals[srt[q]])
                inC = np.zeros(ng, dtype=bool)
                inC[np.concatenate(C)] = True
                run = 0
                while k < len(ip):
                    p = ip[k]
                    if not inC[p] or not (dd[p] > theta):
                        break
                    dd -= Lt[k] ** 2
                    dd[p] = -1.0
                    k += 1
                    run += 1
                if run == 0:
                    out['stuck_at'] = k
                    break
                runs.append(run)
            runs = np.array(runs)
            out['ranks%d_quota%d' % (nranks, quota)] = {
                'candidate_sets': int(len(runs)), 'mean_run': float(runs.mean()), 'median_run': float(np.median(runs)),
                'min_run': int(runs.min()), 'max_run': int(runs.max()),
                'panels_of_128': int(np.sum((runs + 127) // 128))}
            print(json.dumps(out), flush=True)
    return 0


if __name__ == '__main__':
    sys.exit(main())
