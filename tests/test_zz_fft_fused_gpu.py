"""GPU: the opt-in fused Rader path of the stage-4 FFT kernels (ISDF_FFT_FUSED=1, csrc/fft_fused.cu)
against numpy.fft + coulG on prime meshes, 1e-10 relative like the default path.

The line-FFT routines are already checked on the CPU by tests/test_fft_emul_cpu.py; this file checks
the kernels that call them.  First run on a B200 at the end of round 1: all passed
(profiles/r01_optin_forms.md); the path stays opt-in (it gains 5 % of the FFT stage on C64)."""
import os
import numpy as np
import pytest
import test_g1_stages_gpu as stages

pytestmark = pytest.mark.gpu


@pytest.fixture
def fused_fft():
    old = os.environ.get('ISDF_FFT_FUSED')
    os.environ['ISDF_FFT_FUSED'] = '1'
    yield
    if old is None:
        os.environ.pop('ISDF_FFT_FUSED', None)
    else:
        os.environ['ISDF_FFT_FUSED'] = old


@pytest.mark.parametrize('mesh', [[7, 11, 13], [71, 5, 4], [6, 23, 47], [4, 22, 71], [31, 41, 12], [13, 13, 13]])
def test_fused_coulomb_apply(handle, fused_fft, mesh):
    stages.test_coulomb_apply_matches_numpy_fft(handle, mesh)


def test_fused_build_W(handle, fused_fft):
    """W through the forward passes with the sqrt(w coulG) multiply (x-pass MODE 1), prime mesh."""
    import torch
    import oracle
    from pyscf_isdf_b200._lib import pad8
    from util import to_dev
    rng = np.random.default_rng(12)
    mesh = [11, 13, 31]
    a = np.array([[5.0, 0.3, 0.1], [0.2, 4.5, -0.4], [0.0, 0.6, 6.0]])
    ng = int(np.prod(mesh))
    npts = 21
    z = rng.random((npts, ng)) - .5
    Wref = oracle.build_W(z, a, mesh)
    ld = pad8(ng)
    dz = to_dev(handle, z, ld)
    W = torch.zeros((npts, npts), dtype=torch.float64, device=handle.torch_device)
    handle.build_W(dz, npts, ld, mesh, a, W)
    torch.cuda.synchronize()
    assert abs(W.cpu().numpy() - Wref).max() < 1e-10 * abs(Wref).max()
