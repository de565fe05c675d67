#!/usr/bin/env python
"""End-to-end sensitivity of J/K to the digit count S of the int8 split product used for
the W GEMM (prototype through the library int8 GEMM): builds the ISDF object with the FP64
GEMM and with S = 6..9, prints max |dJ|, |dK| against the S = 10 result."""
import os
import sys
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ozaki_proto import ozaki_syrk
from pyscf_isdf_b200 import workloads
from pyscf_isdf_b200.isdf import ISDF


def run(name, c, S):
    cell = workloads.make_cell(name)
    mydf = ISDF(cell, c=c, fit='implicit')
    h = mydf.handle
    orig = h.dgemm

    def spy(M, N, K, A, lda, akm, B, ldb, bkm, C, ldc, **kw):
        if kw.get('upper_only') and S:
            C[:M, :N] = kw.get('alpha', 1.0) * ozaki_syrk(A[:M, :K], S, kc=32768)
            return None
        return orig(M, N, K, A, lda, akm, B, ldb, bkm, C, ldc, **kw)
    h.dgemm = spy
    mydf.build()
    dm = workloads.synthetic_dm(cell.nao)
    vj, vk = mydf.get_jk(dm, exxdiv=None)
    return np.asarray(vj), np.asarray(vk), mydf.W.cpu().numpy()


def main():
    for name, c in (('c8-small', 10), ('c8', 15)):
        ref = run(name, c, 10)
        print('%s c=%g  |J|max %.3e |K|max %.3e |W|max %.3e' % (name, c, abs(ref[0]).max(), abs(ref[1]).max(), abs(ref[2]).max()), flush=True)
        for S in (0, 6, 7, 8, 9):
            vj, vk, W = run(name, c, S)
            print('   %-8s dJ %.2e  dK %.2e  dW %.2e' % ('fp64' if S == 0 else 'S=%d' % S, abs(vj - ref[0]).max(),
                                                        abs(vk - ref[1]).max(), abs(W - ref[2]).max()), flush=True)


if __name__ == '__main__':
    main()
