#!/usr/bin/env python
"""Prototype (torch, library int8 GEMM) of the error-free-split FP64 product used to
decide on the slice count of the tcgen05 int8 W kernel: C = A A^T with A split into S
signed 7-bit digit planes per row-scaled entry.  Prints the error against the FP64
product (and a double-double style reference on a sub-block) for several S, on
(1) synthetic rows with a large dynamic range, (2) the real weighted spectra Y of the
C8 build.  Also times the library int8 GEMM (upper bound on what the digit products cost).
"""
import os
import sys
import time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def split(A, S):
    mx = A.abs().amax(dim=1).clamp_min(1e-300)
    _, e = torch.frexp(mx)                       # mx = m * 2^e, m in [0.5, 1)
    scale = torch.ldexp(torch.ones_like(mx), e)  # |A / scale| < 1
    x = A / scale[:, None]
    out = []
    for s in range(S):
        x = x * (64.0 if s == 0 else 128.0)
        q = torch.round(x)
        x = x - q
        out.append(q.to(torch.int8))
    return out, scale


def ozaki_syrk(A, S, kc=65536):
    n, K = A.shape
    C = torch.zeros((n, n), dtype=torch.float64, device=A.device)
    for k0 in range(0, K, kc):
        Ab = A[:, k0:k0 + kc]
        if Ab.shape[1] % 8:
            Ab = torch.nn.functional.pad(Ab, (0, 8 - Ab.shape[1] % 8))
        q, sc = split(Ab.contiguous(), S)
        acc = torch.zeros((n, n), dtype=torch.float64, device=A.device)
        for g in range(S - 1, -1, -1):
            P = torch.zeros((n, n), dtype=torch.int64, device=A.device)
            for s in range(g + 1):
                P += torch._int_mm(q[s], q[g - s].t())
            acc += P.double() * 2.0 ** (-12 - 7 * g)
        C += acc * sc[:, None] * sc[None, :]
    return C


def report(name, A):
    ref = A @ A.t()
    # higher-precision reference on a 64x64 corner: exact products summed in extended precision on the CPU
    Ah = A[:64].cpu().numpy().astype(np.longdouble)
    ref_ld = (Ah @ Ah.T)
    e64 = np.abs(ref[:64, :64].cpu().numpy().astype(np.longdouble) - ref_ld).max()
    nrm = float(np.abs(ref_ld).max())
    print('%s: n=%d K=%d  |C|max=%.3e   fp64 GEMM vs long double: %.2e (rel %.1e)'
          % (name, A.shape[0], A.shape[1], nrm, e64, e64 / nrm), flush=True)
    for S in (5, 6, 7, 8, 9):
        C = ozaki_syrk(A, S)
        d = float((C - ref).abs().max())
        dl = np.abs(C[:64, :64].cpu().numpy().astype(np.longdouble) - ref_ld).max()
        print('   S=%d  vs fp64 GEMM %.2e (rel %.1e)   vs long double %.2e (rel %.1e)'
              % (S, d, d / nrm, dl, dl / nrm), flush=True)


def real_Y():
    """Capture Yr (the real view of the weighted half spectra) of the C8 build."""
    from pyscf_isdf_b200 import workloads
    from pyscf_isdf_b200.isdf import ISDF
    cell = workloads.make_cell('c8-small')
    mydf = ISDF(cell, c=10, fit='implicit')
    h = mydf.handle
    grabbed = {}
    orig = h.dgemm

    def spy(M, N, K, A, lda, akm, B, ldb, bkm, C, ldc, **kw):
        if kw.get('upper_only'):
            grabbed['Y'] = A[:M, :K].clone()
        return orig(M, N, K, A, lda, akm, B, ldb, bkm, C, ldc, **kw)
    h.dgemm = spy
    mydf.build()
    return grabbed['Y']


def main():
    dev = torch.device('cuda:0')
    torch.manual_seed(0)
    n, K = 1024, 40000
    A = torch.randn((n, K), dtype=torch.float64, device=dev)
    report('gaussian', A)
    decay = torch.logspace(0, -9, K, dtype=torch.float64, device=dev)
    report('decaying columns 1..1e-9', A * decay[None, :])
    try:
        Y = real_Y()
        report('C8 weighted spectra', Y)
    except Exception as e:                       # noqa: BLE001  (prototype)
        print('real Y capture failed:', repr(e))
    # library int8 GEMM rate
    n, kc = 12480, 65536
    a = torch.randint(-64, 65, (n, kc), dtype=torch.int8, device=dev)
    for _ in range(2):
        torch._int_mm(a, a.t())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        torch._int_mm(a, a.t())
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print('library int8 GEMM %dx%dx%d: %.2f ms  %.2f POPS' % (n, n, kc, dt * 1e3, 2.0 * n * n * kc / dt / 1e15))


if __name__ == '__main__':
    main()
