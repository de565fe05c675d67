#!/usr/bin/env python
"""Correctness + rate of isdf_syrk_i8 (int8 tcgen05 split product) against torch FP64."""
import os
import sys
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import _lib


def main():
    h = _lib.Handle(0)
    sizes = [(128, 128), (200, 300), (700, 5000), (1000, 70001)]
    if len(sys.argv) > 1:
        sizes = [tuple(int(x) for x in a.split('x')) for a in sys.argv[1:]]
    for n, K in sizes:
        lda = K + (K % 2)
        A = torch.zeros((n, lda), dtype=torch.float64, device='cuda')
        A[:, :K] = torch.randn((n, K), dtype=torch.float64, device='cuda')
        C = h.empty((n, n))
        for S in (8,):
            h.syrk_i8(n, K, A, lda, C, n, nslices=S)
            h.symmetrize_upper(C, n, n)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            h.syrk_i8(n, K, A, lda, C, n, nslices=S)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            h.symmetrize_upper(C, n, n)
            if n * K <= 4e9:
                ref = A[:, :K] @ A[:, :K].t()
                err = float((C - ref).abs().max() / ref.abs().max())
            else:
                err = float('nan')
            ops = float(getattr(h, 'syrk_i8_ops', 0.0)) or S * (S + 1) / 2 * n * n * K     # int8 ops the library issued (either form)
            print('n=%6d K=%7d S=%d  rel err %.2e   %.2f ms  %.2f POP/s (int8)  %.1f TFLOP/s FP64-equivalent'
                  % (n, K, S, err, ms, ops / ms / 1e12, n * n * K / ms / 1e9), flush=True)
        del A, C


if __name__ == '__main__':
    main()
