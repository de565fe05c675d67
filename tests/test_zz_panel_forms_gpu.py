"""GPU: the three forms of the stage-2 pivot kernel (cooperative grid / cluster with DSMEM reads /
cluster with st.async pushes, ISDF_PANEL_CLUSTER = 0 / 1 / 2) give pivots, residuals and factor
bit-identical to the unblocked FMA-chain oracle, on generic data and on a symmetric cell with exact ties.

Forms 1 and 2 are opt-in in the library (the default is form 0).  First run on a B200 at the end of
round 1: both tests passed and form 2 saves 34 ms of the C64 selection (profiles/r01_optin_forms.md).  The
file stays behind ISDF_TEST_EXPERIMENTAL=1 until the cluster protocols have more than one run behind them
(multi-rank, compute-sanitizer): tools/gpu_r2_first.sh."""
import os
import numpy as np
import pytest
from oracle import c_oracle
from util import diamond_c8, oracle_phi
from test_g1_stages_gpu import _select

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get('ISDF_TEST_EXPERIMENTAL') != '1',
                                 reason='opt-in kernel forms not yet validated on a B200 (set ISDF_TEST_EXPERIMENTAL=1)')]


@pytest.fixture
def panel_mode():
    old = os.environ.get('ISDF_PANEL_CLUSTER')

    def set_mode(m):
        os.environ['ISDF_PANEL_CLUSTER'] = str(m)
    yield set_mode
    if old is None:
        os.environ.pop('ISDF_PANEL_CLUSTER', None)
    else:
        os.environ['ISDF_PANEL_CLUSTER'] = old


@pytest.mark.parametrize('mode', [1, 2])
def test_cluster_forms_bit_exact(handle, panel_mode, mode):
    rng = np.random.default_rng(3)
    phi = rng.standard_normal((24, 5000))
    ip_ref, resid_ref, Lt_ref = c_oracle.select_ip_chain(phi, 300, return_factor=True)
    panel_mode(mode)
    ip, resid, Lt = _select(handle, phi, 300)
    used = handle.lib.isdf_select_last_panel_mode()
    assert np.array_equal(ip, ip_ref)
    assert np.array_equal(resid, resid_ref)
    assert np.array_equal(Lt, Lt_ref)
    assert used == mode, 'form %d was requested but form %d ran (cluster not schedulable?)' % (mode, used)
    # symmetric cell: exact ties, lowest grid index wins; several panels, certificate ends some early
    cell = diamond_c8([12, 12, 12], basis='gth-szv')
    phi = oracle_phi(cell)
    ip_ref, resid_ref = c_oracle.select_ip_chain(phi, 200)
    ip, resid, _ = _select(handle, phi, 200)
    assert np.array_equal(ip, ip_ref)
    assert np.array_equal(resid, resid_ref)
    # rank-deficient input: the tolerance stop inside the panel
    small = rng.standard_normal((3, 700))
    ip_ref, _ = c_oracle.select_ip_chain(small, 50)
    ip, _, _ = _select(handle, small, 50)
    assert np.array_equal(ip, ip_ref)
