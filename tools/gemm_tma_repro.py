#!/usr/bin/env python
"""Flush-shaped product of a C8 build (M 52, N 42875, K 194, A k-contiguous, in place), TMA forms against the
LDGSTS kernel, repeated; prints where the results differ."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyscf_isdf_b200 import _lib
if os.environ.get('LIB'):
    _lib.LIB_PATH = os.path.abspath(os.environ['LIB'])
from pyscf_isdf_b200._lib import Handle

h = Handle(0)
dev = h.torch_device
g = torch.Generator(device=dev); g.manual_seed(1)
rnd = lambda *s: torch.rand(*s, dtype=torch.float64, device=dev, generator=g) - 0.5
M, N, K, ld = int(os.environ.get('M', 52)), 42875, int(os.environ.get('K', 194)), 42880
A = rnd(M, 302 if K <= 302 else K + 2)
Lt = rnd(K + M, ld)
C0 = Lt[K:].clone()
def run(form):
    Lt[K:].copy_(C0)
    h.set_gemm_tile(form)
    C = Lt[K:]
    h.dgemm(M, N, K, A, A.shape[1], 0, Lt, ld, 1, C, ld, Cin=C, ldcin=ld, neg=1)
    torch.cuda.synchronize()
    return C[:, :N].clone()
ref = run(4)
nbad = 0
for form in [int(x) for x in os.environ.get('FORMS', '10,7,8,9').split(',')]:
    for rep in range(int(os.environ.get('REPS', 6))):
        got = run(form)
        bad = (got != ref).nonzero()
        if len(bad):
            nbad += 1
            n = bad[:, 1].cpu().numpy(); m = bad[:, 0].cpu().numpy()
            ctas = np.unique(n // 64)
            if nbad <= 3: print('form', form, 'rep', rep, 'bad', len(bad), 'ctas', ctas[:12], '... n%64', np.unique(n % 64)[:20], 'm', np.unique(m)[:20],
                  'maxdiff', float((got - ref).abs().max()), flush=True)
        else:
            pass
print('done: runs with mismatches', nbad, flush=True)
