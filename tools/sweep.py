#!/usr/bin/env python
"""BASELINE configs[4]: interpolation-point selection + FFT Coulomb microbench sweep over
N_IP in {2k, 5k, 10k, 20k, 40k} x mesh in {48, 64, 96, 128}^3 on ONE B200 (SURVEY 8d "S2 and S4 standalone").

Per point: stage 2 (`isdf_select_ip`: exact blocked pivoted Cholesky, N_IP points out of Ng) and the stage-4
transform (`isdf_fft_zy` + `isdf_fft_x_coulomb`: batched 3-D R2C fused with sqrt(w coulG), N_IP vectors in
batches through one spectrum buffer), each timed with CUDA events after one warm-up, on synthetic AO-like
data (nao = N_IP / 15 periodic Gaussians at random centres; same flop / byte model as the real cells).
Points whose factor rows [N_IP, Ng] do not fit the device next to the AO matrix are listed with the number
of GPUs the grid-slab sharding would need, not run.  Writes one JSON document (commit under profiles/)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import _lib


def synthetic_phi(h, nao, mesh, seed):
    """[nao, ld] periodic Gaussians exp(-alpha |r - R|^2_min-image) on the mesh of a cubic cell (test data: torch)."""
    n = mesh[0]
    ng = n ** 3
    ld = _lib.pad8(ng)
    g = torch.Generator(device='cpu').manual_seed(seed)
    cen = torch.rand((nao, 3), generator=g, dtype=torch.float64).to(h.torch_device)
    # widths fixed in units of the cell (2.5 .. 4.5 spacings of a 48^3 mesh): the pair space keeps its rank as the mesh grows
    width = (n / 48.0) * (2.5 + 2.0 * torch.rand((nao,), generator=g, dtype=torch.float64)).to(h.torch_device)
    ax = torch.arange(n, dtype=torch.float64, device=h.torch_device) / n
    phi = h.empty((nao, ld))
    phi[:, ng:] = 0
    for m0 in range(0, nao, 64):
        c = cen[m0:m0 + 64]
        a = 1.0 / (2.0 * (width[m0:m0 + 64] / n) ** 2)
        parts = []
        for d in range(3):
            dx = ax[None, :] - c[:, d:d + 1]
            dx = dx - torch.round(dx)
            parts.append(torch.exp(-a[:, None] * dx * dx))
        blk = parts[0][:, :, None, None] * parts[1][:, None, :, None] * parts[2][:, None, None, :]
        phi[m0:m0 + 64, :ng] = blk.reshape(blk.shape[0], ng)
    return phi, ld


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--nip', default='2000,5000,10000,20000,40000')
    ap.add_argument('--mesh', default='48,64,96,128')
    ap.add_argument('--budget-gb', type=float, default=150.0, help='device memory a point may use')
    ap.add_argument('--out', default='gpurun_out/sweep.json')
    args = ap.parse_args()
    h = _lib.Handle(0)
    dev = h.torch_device
    a = np.eye(3) * 10.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))
    except (OSError, ValueError):
        pass
    hbm = float(peaks.get('hbm_gbs', 6537.0))
    rows = []
    for n in [int(x) for x in args.mesh.split(',')]:
        mesh = [n, n, n]
        ng = n ** 3
        ngh = n * n * (n // 2 + 1)
        for nip in [int(x) for x in args.nip.split(',')]:
            nao = max(8, nip // 15)
            ld = _lib.pad8(ng)
            need = 8.0 * (nao + nip) * ld + h.lib.isdf_select_ip_ws_bytes(nao, ng, nip) + 16.0 * ngh * 256
            row = {'mesh': n, 'ngrids': ng, 'n_ip': nip, 'nao': nao, 'device_bytes': need}
            if need > args.budget_gb * 1e9 or nip > ng:
                row['skipped'] = 'factor rows [N_IP, Ng] = %.0f GB: needs the grid-slab sharding over >= %d GPUs' % (
                    8.0 * nip * ld / 1e9, int(np.ceil(need / (args.budget_gb * 1e9))))
                rows.append(row)
                print(json.dumps(row), flush=True)
                continue
            phi, ld = synthetic_phi(h, nao, mesh, seed=n * 100003 + nip)
            Lt = h.empty((nip, ld))
            sel = []
            for rep in range(2):                       # warm-up + timed
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                ip, resid = h.select_ip(phi, nao, ng, ld, nip, -1.0, Lt, ld)
                e1.record()
                torch.cuda.synchronize(dev)
                sel.append(e0.elapsed_time(e1))
            rank = len(ip)
            flops = 2.0 * ng * nao * rank + 1.0 * ng * rank * rank
            row.update(select_ms=sel[-1], select_rank=rank, select_tflops=flops / (sel[-1] * 1e-3) / 1e12,
                       resid_ratio=float(resid[-1] / resid[0]))
            # stage-4 transform of the factor rows, in batches through one spectrum buffer
            vb = max(1, min(rank, int((1 << 30) // (16 * ngh))))
            spec = h.empty_complex((vb * ngh,))
            fft = []
            for rep in range(2):
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
                for v0 in range(0, rank, vb):
                    nv = min(vb, rank - v0)
                    h.fft_zy(Lt[v0:], ld, nv, n, mesh, spec, ngh)
                    h.fft_x_coulomb(spec, ngh, nv, mesh, a, 0, n * (n // 2 + 1))
                e1.record()
                torch.cuda.synchronize(dev)
                fft.append(e0.elapsed_time(e1))
            byt = rank * (8.0 * ng + 16.0 * ngh)
            row.update(fft_ms=fft[-1], fft_algorithmic_gb=byt / 1e9, fft_gbs=byt / (fft[-1] * 1e-3) / 1e9,
                       fft_frac_of_hbm=byt / (fft[-1] * 1e-3) / 1e9 / hbm)
            rows.append(row)
            print(json.dumps(row), flush=True)
            del phi, Lt, spec
            torch.cuda.empty_cache()
    doc = {'what': 'BASELINE configs[4] sweep, 1 x B200, FP64', 'hbm_peak_gbs': hbm,
           'flop_model': 'select: 2 Ng nao Np + Ng Np^2; fft bytes: 8 Np Ng + 16 Np Ngh (one read of the real vectors, one write of the half spectra)',
           'when': time.strftime('%Y-%m-%dT%H:%M:%SZ', time.gmtime()), 'rows': rows}
    os.makedirs(os.path.dirname(args.out) or '.', exist_ok=True)
    with open(args.out, 'w') as f:
        json.dump(doc, f, indent=1)
    return 0


if __name__ == '__main__':
    sys.exit(main())
