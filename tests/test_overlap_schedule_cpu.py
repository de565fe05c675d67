"""CPU: the two-stream schedule of the residue int8 product (ISDF_SYRK_OVERLAP=1, csrc/syrk_i8.cu) as a
happens-before graph.  The enqueue order of the C++ driver is restated here operation by operation (stream,
kernel, buffers read / written, events waited on / recorded); the test derives the ordering CUDA guarantees
(program order within a stream, record -> wait across streams) and checks that every pair of conflicting
accesses (write-write, read-write on the same buffer) is ordered, that the updates of C happen in chunk
order, and that the products of chunk i really can run beside the element-wise kernels of its neighbours."""
import itertools
import pytest


def schedule(m):
    """Operations in enqueue order: (name, stream, reads, writes, waits, record)."""
    ops = []
    ops.append(('rowstats', 's0', {'A'}, {'sc'}, [], 'start'))

    def split(i):
        waits = ['start'] if not any(o[1] == 's1' for o in ops) else []
        if i >= 2:
            waits.append('syrk%d' % (i - 2))
        ops.append(('split%d' % i, 's1', {'A', 'sc'}, {'planes%d' % (i & 1)}, waits, 'split%d' % i))

    if m > 0:
        split(0)
    if m > 1:
        split(1)
    for i in range(m):
        waits = ['split%d' % i] + (['recon%d' % (i - 2)] if i >= 2 else [])
        ops.append(('syrk%d' % i, 's0', {'planes%d' % (i & 1), 'tiles'}, {'resid%d' % (i & 1)}, waits, 'syrk%d' % i))
        ops.append(('recon%d' % i, 's1', {'resid%d' % (i & 1), 'sc'}, {'C'}, ['syrk%d' % i], 'recon%d' % i))
        if i + 2 < m:
            split(i + 2)
    if m > 0:
        ops.append(('after', 's0', {'C'}, {'A', 'ws'}, ['recon%d' % (m - 1)], None))     # whatever the caller queues next
    return ops


def happens_before(ops):
    n = len(ops)
    hb = [[False] * n for _ in range(n)]
    recorded = {}
    last = {}
    for j, (name, stream, rd, wr, waits, rec) in enumerate(ops):
        if stream in last:
            hb[last[stream]][j] = True
        for e in waits:
            hb[recorded[e]][j] = True          # the event must have been recorded earlier in enqueue order
        last[stream] = j
        if rec:
            recorded[rec] = j
    for k, i, j in itertools.product(range(n), repeat=3):
        if hb[i][k] and hb[k][j]:
            hb[i][j] = True
    return hb


@pytest.mark.parametrize('m', [0, 1, 2, 3, 4, 7, 11])
def test_every_conflict_is_ordered(m):
    ops = schedule(m)
    hb = happens_before(ops)
    for i, j in itertools.combinations(range(len(ops)), 2):
        _, _, ri, wi, _, _ = ops[i]
        _, _, rj, wj, _, _ = ops[j]
        if (wi & wj) or (wi & rj) or (ri & wj):
            assert hb[i][j] or hb[j][i], (ops[i][0], ops[j][0])
    idx = {o[0]: k for k, o in enumerate(ops)}
    for i in range(m - 1):
        assert hb[idx['recon%d' % i]][idx['recon%d' % (i + 1)]]           # C is updated in chunk order
    if m >= 3:
        # the overlap the form exists for: products of chunk 1 unordered with recon(0) and split(2)
        assert not hb[idx['syrk1']][idx['recon0']] and not hb[idx['recon0']][idx['syrk1']]
        assert not hb[idx['syrk1']][idx['split2']] and not hb[idx['split2']][idx['syrk1']]
