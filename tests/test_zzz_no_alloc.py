"""GPU: isdf_get_jk_static -- the J/K contraction with its scratch in a fixed device array inside the
library (no allocation on the call path) -- against isdf_get_jk with a caller workspace, through the C-ABI.
Collected last on purpose.  Passes alone in a fresh process and after
the other GPU tests: the only state it shares is the scratch array, which every call rewrites."""
import ctypes

import numpy as np
import pytest

from util import rand_dm

pytestmark = pytest.mark.gpu


def _case(handle, nao, npts, seed):
    import torch
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((nao, npts))
    W = rng.standard_normal((npts, npts)); W = W + W.T
    dms = np.stack([rand_dm(nao, seed), rng.standard_normal((nao, nao))])      # symmetric and general
    dev = handle.torch_device
    return (torch.from_numpy(X).to(dev), torch.from_numpy(W).to(dev), torch.from_numpy(dms).to(dev))


def test_static_scratch_jk_equals_workspace_jk_twice_in_a_row(handle):
    """Two calls in a row with different sizes and data: the second must not see the first one's scratch.
    Same kernels in the same order, so the tolerance (1e-10 relative, that of test_get_jk_matches_oracle)
    is met with room to spare."""
    import torch
    assert handle.lib.isdf_get_jk_static_max_bytes() == 32 << 20
    for nao, npts, seed in [(26, 130, 31), (40, 412, 32)]:
        X, W, dm = _case(handle, nao, npts, seed)
        assert handle.lib.isdf_get_jk_ws_bytes(nao, npts) <= handle.lib.isdf_get_jk_static_max_bytes()
        vj0 = torch.zeros_like(dm); vk0 = torch.zeros_like(dm)
        handle.get_jk(X, W, dm, 2, nao, npts, True, True, vj0, vk0)
        vj1 = torch.full_like(dm, float('nan')); vk1 = torch.full_like(dm, float('nan'))
        handle.get_jk_static(X, W, dm, 2, nao, npts, True, True, vj1, vk1)
        torch.cuda.synchronize()
        # a repeat of the same call (kernels loaded, scratch address resolved) leaves the device's free memory as it was
        vj3 = torch.full_like(dm, float('nan')); vk3 = torch.full_like(dm, float('nan'))
        free_before = torch.cuda.mem_get_info(handle.torch_device)[0]
        handle.get_jk_static(X, W, dm, 2, nao, npts, True, True, vj3, vk3)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(handle.torch_device)[0] == free_before
        assert torch.equal(vj3, vj1) and torch.equal(vk3, vk1)
        for a, b in ((vj1, vj0), (vk1, vk0)):
            a, b = a.cpu().numpy(), b.cpu().numpy()
            assert np.isfinite(a).all()
            assert abs(a - b).max() <= 1e-10 * abs(b).max()
        # K only: J output untouched
        vk2 = torch.zeros_like(dm)
        handle.get_jk_static(X, W, dm, 2, nao, npts, False, True, None, vk2)
        torch.cuda.synchronize()
        assert abs(vk2.cpu().numpy() - vk0.cpu().numpy()).max() <= 1e-10 * abs(vk0.cpu().numpy()).max()


def test_static_scratch_rejects_sizes_beyond_its_maximum(handle):
    import torch
    nao, npts = 16, 2100                       # np^2 doubles alone are 35 MB > 32 MiB
    assert handle.lib.isdf_get_jk_ws_bytes(nao, npts) > handle.lib.isdf_get_jk_static_max_bytes()
    z = torch.zeros(8, dtype=torch.float64, device=handle.torch_device)
    with pytest.raises(RuntimeError, match='needs .* bytes of scratch'):
        handle.get_jk_static(z, z, z, 1, nao, npts, True, True, z, z)
