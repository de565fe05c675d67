"""GPU: the residue (CRT) form of the int8 metric product (ISDF_SYRK_CRT=<moduli>, csrc/syrk_i8.cu +
crt_core.cuh) against the FP64 product: 16 moduli carry 53-bit inputs, so the result agrees with an
FP64 GEMM to a few 1e-15 relative to max|C| (2e-13 here, as for the digit-plane form).

The residue arithmetic itself is checked on the CPU against exact integers
(tests/test_crt_emul_cpu.py); this file checks the kernels.  Written after the last GPU session of
round 1: skipped unless ISDF_TEST_EXPERIMENTAL=1 until it has been run on a B200."""
import os
import pytest
import test_syrk_i8_gpu as base

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get('ISDF_TEST_EXPERIMENTAL') != '1',
                                 reason='opt-in kernel forms not yet validated on a B200 (set ISDF_TEST_EXPERIMENTAL=1)')]


@pytest.fixture
def crt16():
    old = os.environ.get('ISDF_SYRK_CRT')
    os.environ['ISDF_SYRK_CRT'] = '16'
    yield
    if old is None:
        os.environ.pop('ISDF_SYRK_CRT', None)
    else:
        os.environ['ISDF_SYRK_CRT'] = old


@pytest.mark.parametrize('n,K', [(200, 300), (128, 128), (257, 1000), (700, 5000), (1000, 70001)])
def test_crt_matches_fp64(crt16, n, K):
    assert base._run(n, K) < 2e-13


def test_crt_alpha_ld_and_dynamic_range(crt16):
    assert base._run(520, 40000, alpha=0.37, pad=3, decay=True) < 2e-13


def test_crt_fewer_bits_for_fewer_planes(crt16):
    """nslices sets the accuracy target per chunk; the CRT form picks the fewest moduli that carry it."""
    e6, e8 = base._run(300, 4000, S=6), base._run(300, 4000, S=8)
    assert e8 < 2e-13 and e6 < 1e-9
