"""The panel-blocked, threaded C oracle of the interpolation-point selection (oracle/c/isdf_oracle_blocked.c)
against the textbook loop (oracle/c/isdf_oracle.c): pivots, residuals and factor bit for bit -- generic data,
a symmetric cell (tied residuals: the certificate fails and the fall-back scan runs), rank-deficient data
(tolerance stop), fewer grid points than candidates, ragged last panel."""
import numpy as np
import pytest

from oracle import c_oracle
from util import diamond_c8, oracle_phi


def _same(phi, npts, tol=-1.0):
    a = c_oracle.select_ip_chain(phi, npts, tol, return_factor=True)
    b = c_oracle.select_ip_blocked(phi, npts, tol, return_factor=True)
    assert len(a[0]) == len(b[0])
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    return len(a[0])


@pytest.mark.parametrize('nao,ng,npts', [(10, 500, 60), (20, 5000, 300), (7, 2500, 28 * 7 // 2 + 5), (16, 3001, 130)])
def test_blocked_oracle_equals_loop_on_random_data(nao, ng, npts):
    rng = np.random.default_rng(nao)
    phi = rng.standard_normal((nao, ng)) * np.exp(-3 * rng.random(ng))
    rank = _same(phi, npts)
    assert rank <= nao * (nao + 1) // 2 + 1


def test_blocked_oracle_on_a_symmetric_cell_with_tied_residuals():
    cell = diamond_c8([12, 12, 12], basis='gth-szv', jitter=0.0)     # full point group: many equal residuals
    phi = oracle_phi(cell)
    _same(phi, 6 * cell.nao)
    assert _same(phi, 40, tol=1e-3) <= 40                              # user tolerance stop


def test_blocked_oracle_on_a_jittered_dzvp_cell():
    cell = diamond_c8([13, 12, 11], basis='gth-dzvp', jitter=0.15)
    _same(oracle_phi(cell), 4 * cell.nao)
