"""CPU: PROTOTYPE oracle of the exact-arithmetic selection proposed for round 2 (DESIGN.md section 8
item 3; oracle/c/isdf_oracle_exact.c).  Not a parity test of the product -- the product still runs the
FMA-chain arithmetic -- but the checks that make the proposal credible:
  * R(S) (integer -> FP64 through centred mixed-radix digits + FMA Horner) is exact below 2^53, accurate
    to 2^-50 above, and equal bit for bit to the device routine of csrc/crt_core.cuh compiled for the host;
  * on generic data the exact-arithmetic selection picks the same points as the FMA-chain oracle and its
    factor agrees to rounding level;
  * its long sums do not depend on the summation order: permuting the AO index gives identical bits
    (which the FMA chain cannot offer, and what makes split-K / any tiling / any rank count reproducible).
"""
import ctypes
import numpy as np
import pytest
from oracle import c_oracle
from test_crt_emul_cpu import crt  # noqa: F401  (fixture: the device routine compiled with g++)


def test_recon_exact_and_equal_to_device_routine(crt):  # noqa: F811
    rng = np.random.default_rng(2)
    p = [crt.crt_modulus(k) for k in range(16)]
    arr = (ctypes.c_int * 16)()
    vals = [0, 1, -1, 2**53, -(2**53) + 1, 2**123, -(2**123) + 12345]
    for bits in (10, 40, 53, 54, 80, 106, 121):
        vals += [int(rng.integers(0, 2**62)) * int(rng.integers(0, 2**62)) % (1 << bits) * (1 if rng.random() < .5 else -1)
                 for _ in range(40)]
    for x in vals:
        got = c_oracle.recon_int(x)
        if abs(x) <= 2**53:
            assert got == float(x)
        else:
            assert abs(got - float(x)) <= 2.0**-50 * abs(float(x))
        for k, q in enumerate(p):
            arr[k] = x % q
        assert got == crt.crt_reconstruct(arr, 16)            # same digits, same FMA Horner: same bits


def test_same_points_as_the_chain_oracle_on_generic_data():
    rng = np.random.default_rng(3)
    phi = rng.standard_normal((24, 1500))
    npts = 300                                              # crosses two 128-column block boundaries
    ip0, resid0, L0 = c_oracle.select_ip_chain(phi, npts, return_factor=True)
    ip1, resid1, L1 = c_oracle.select_ip_exact(phi, npts, return_factor=True)
    assert np.array_equal(ip0, ip1)
    assert abs(resid1 - resid0).max() <= 1e-10 * resid0.max()
    assert abs(L1 - L0).max() <= 1e-9 * abs(L0).max()
    assert not np.array_equal(L0, L1)                        # a different arithmetic, not the same bits


def test_long_sums_are_order_independent():
    rng = np.random.default_rng(4)
    phi = rng.standard_normal((17, 600)) * np.exp(rng.standard_normal((17, 1)))
    perm = rng.permutation(17)
    a = c_oracle.select_ip_exact(phi, 140, return_factor=True)
    b = c_oracle.select_ip_exact(phi[perm], 140, return_factor=True)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    # the FMA-chain arithmetic does depend on the order (same points here, different bits)
    c = c_oracle.select_ip_chain(phi, 140, return_factor=True)
    e = c_oracle.select_ip_chain(phi[perm], 140, return_factor=True)
    assert not np.array_equal(c[2], e[2])
