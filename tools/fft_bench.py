#!/usr/bin/env python
"""FFT microbench: forward 3-D R2C + fused Coulomb weight (the stage-4 FFT part of
isdf_build_W) on the meshes of the BASELINE configs; reports ms and the fraction of HBM
peak for the algorithmic bytes (8*Ng read + 16*Ngh written per vector).
Tuning knob: ISDF_FFT_SMEM_KB (shared memory per CTA)."""
import os
import sys
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import _lib


def hbm_peak():
    import json
    try:
        return float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))['hbm_gbs'])
    except (OSError, ValueError, KeyError):
        return 6537.0


def main():
    h = _lib.Handle(0)
    peak = hbm_peak()
    a = np.eye(3) * 13.0
    cases = (([35] * 3, 1560), ([71] * 3, 1024), ([96] * 3, 512), ([108] * 3, 384))
    if len(sys.argv) > 1:                       # e.g. `fft_bench.py 71 64` -> only those meshes (any size)
        known = {c[0][0]: c for c in cases}
        cases = [known.get(int(a), ([int(a)] * 3, max(64, int(4.0e8 / int(a) ** 3)))) for a in sys.argv[1:]]
    for mesh, nvec in cases:
        ng = int(np.prod(mesh))
        ngh = mesh[0] * mesh[1] * (mesh[2] // 2 + 1)
        ld = _lib.pad8(ng)
        z = torch.rand((nvec, ld), dtype=torch.float64, device=h.torch_device) - 0.5
        W = h.empty((nvec, nvec))
        best = 1e30
        for _ in range(3):
            h.build_W(z, nvec, ld, mesh, a, W)
            torch.cuda.synchronize()
            best = min(best, h.stage_ms(_lib.STAGE_FFT))
        byt = nvec * (8.0 * ng + 16.0 * ngh)
        print('mesh %3d^3 nvec %5d: fft %8.3f ms  algorithmic %.1f GB -> %.0f GB/s (%.1f%% of %.0f)'
              % (mesh[0], nvec, best, byt / 1e9, byt / best / 1e6, 100 * byt / best / 1e6 / peak, peak), flush=True)
        del z, W


if __name__ == '__main__':
    main()
