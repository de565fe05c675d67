#!/usr/bin/env python
"""GEMM microbench through the C-ABI (isdf_dgemm): the shapes/layouts of the ISDF
stages, CUDA-event timed, against torch.matmul (cuBLAS) on the same shape.
usage: gemm_bench.py [tile ...]   (tile: 0 auto, 2 64x128, 3 64x64)"""
import os
import sys
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import _lib
if os.environ.get('LIB'):
    _lib.LIB_PATH = os.path.abspath(os.environ['LIB'])
from pyscf_isdf_b200._lib import Handle


def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    tiles = [int(x) for x in sys.argv[1:]] or [0, 2, 3]
    only = os.environ.get('GEMM_CASE')
    h = Handle(0)
    dev = h.torch_device
    g = torch.Generator(device=dev); g.manual_seed(0)
    rnd = lambda *s: torch.rand(*s, dtype=torch.float64, device=dev, generator=g) - 0.5
    NG = 357912
    cases = [
        # name, M, N, K, a_kmajor, b_kmajor, neg, cin, upper
        ('flush  C-=A^T B  (km,km)', 128, NG, 6240, 1, 1, 1, True, 0),
        ('pairs  (A^T B)^2 (km,km)', 128, NG, 832, 1, 1, 0, False, 0),
        ('solveU C-=A B    (kc,km)', 128, NG, 6240, 0, 1, 1, True, 0),
        ('W      A A^T     (kc,kc) upper', 4096, 4096, 65536, 0, 0, 0, False, 1),
        ('T      A^T A     (km,km)', 1024, 1024, 6240, 1, 1, 0, False, 0),
        ('Tkc    A A^T     (kc,kc)', 1024, 1024, 6240, 0, 0, 0, False, 0),
        ('wsolvL C-=A B    (kc,km) narrow', 128, 1560, 6240, 0, 1, 1, True, 0),
        ('wsolvR C-=A B    (kc,km) narrow', 6240, 1560, 128, 0, 1, 1, True, 0),
        ('JK     X^T T     (km,km)', 12480, 12480, 832, 1, 1, 0, False, 0),
        ('JK     U X^T     (kc,kc)', 832, 832, 12480, 0, 0, 0, False, 0),
    ]
    for name, M, N, K, ak, bk, neg, cin, upper in cases:
        if only and only not in name:
            continue
        A = rnd((K, M) if ak else (M, K))
        B = rnd((K, N) if bk else (N, K))
        C = torch.zeros((M, N), dtype=torch.float64, device=dev)
        flops = 2.0 * M * N * K * (0.5 if upper else 1.0)
        row = '%-34s M=%6d N=%6d K=%6d ' % (name, M, N, K)
        for t in tiles:
            h.set_gemm_tile(t)
            ms = timeit(lambda: h.dgemm(M, N, K, A, A.shape[1], ak, B, B.shape[1], bk, C, N,
                                        Cin=C if cin else None, ldcin=N, neg=neg, upper_only=upper))
            row += '| tile%d %7.2f TF/s ' % (t, flops / ms / 1e9)
        At = A.t() if ak else A
        Bt = B if bk else B.t()
        ms = timeit(lambda: torch.matmul(At, Bt, out=C))
        row += '| cuBLAS %7.2f TF/s' % (2.0 * M * N * K / ms / 1e9)
        print(row, flush=True)
        del A, B, C


if __name__ == '__main__':
    main()
