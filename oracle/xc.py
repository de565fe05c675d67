"""CPU oracle (test infrastructure): exchange-correlation on the uniform grid --
``NumInt.nr_rks`` of the reference (pyscf/pbc/dft/numint.py:283-381: density and
its gradient from the AO values :96-188, functional, ``weight * vxc``, ``_vxc_mat``
:805-827 with the 0.5 of :987-994, ``vmat + vmat^H`` :374) for closed shells.

The functionals themselves come from libxc 6 (a dependency of the reference that is
not in its tree; the reference calls it through ``pyscf/dft/libxc.py:eval_xc``), so
their published closed forms are restated here as energy densities ``f(rho, sigma)``
per unit volume and differentiated by the complex-step rule (exact to rounding for
these analytic expressions):
  Slater exchange; the semilocal parts of PBE0 (0.75 PBE exchange + PBE correlation) and of libxc's B3LYP
  (0.08 Slater + 0.72 B88 + 0.81 LYP + 0.19 VWN-RPA); VWN correlation (form V, Can. J. Phys. 58, 1200; libxc's lda_c_vwn, the
  'vwn' of the reference's default 'LDA,VWN'); Becke 88 exchange (Phys. Rev. A 38, 3098); LYP correlation in the
  form of Miehlich et al. (Chem. Phys. Lett. 157, 200); PBE exchange and correlation
  (Phys. Rev. Lett. 77, 3865) on the PW92 correlation with libxc's ``pw_mod`` constants.

Pinning: ``nr_rks(..., 'blyp', ...)`` reproduces the reference's fixed numbers of
pyscf/pbc/dft/test/test_numint.py:183-206 (electron count, exc, fingerprint of vmat at a
k-point), which pins the density, gradient and matrix assembly and the B88 and LYP
restatements; the AO gradients reproduce test_numint.py:77-100.  Slater + VWN is pinned through the reference's
Gamma-point RKS energies of diamond (pyscf/pbc/dft/test/test_krks.py:90-118, tests/test_scf_cpu.py).  The
reference holds no fixed number for PBE: its formulas are **unpinned** beyond the published limits checked in
tests/test_xc_cpu.py.
"""
import numpy as np

H = 1e-30          # complex step, relative to the argument


def f_slater(rho, sigma):
    return -0.75 * (3 / np.pi)**(1 / 3.) * rho**(4 / 3.)


def f_b88(rho, sigma):
    beta = 0.0042
    ra = 0.5 * rho                                    # one spin channel
    ga = 0.5 * np.sqrt(sigma)
    r43 = ra**(4 / 3.)
    x = ga / r43
    cx = 1.5 * (3 / (4 * np.pi))**(1 / 3.)
    return 2 * (-cx * r43 - beta * r43 * x * x / (1 + 6 * beta * x * np.arcsinh(x)))


def f_lyp(rho, sigma):
    a, b, c, d = 0.04918, 0.132, 0.2533, 0.349
    cf = 0.3 * (3 * np.pi**2)**(2 / 3.)
    r13 = rho**(-1 / 3.)
    den = 1 + d * r13
    omega = np.exp(-c * r13) / den * rho**(-11 / 3.)
    delta = c * r13 + d * r13 / den
    ra = 0.5 * rho
    saa = 0.25 * sigma                                # |grad rho_a|^2 = |grad rho_b|^2
    t1 = -a * 4 / den * ra * ra / rho
    brace = ra * ra * (2**(11 / 3.) * cf * 2 * ra**(8 / 3.)
                       + (47 / 18. - 7 * delta / 18.) * sigma
                       - (2.5 - delta / 18.) * 2 * saa
                       - (delta - 11) / 9. * 2 * (ra / rho) * saa) \
        - 2 / 3. * rho**2 * sigma + 2 * (2 / 3. * rho**2 - ra * ra) * saa
    return t1 - a * b * omega * brace


def f_pbe_x(rho, sigma):
    kappa, mu = 0.804, 0.2195149727645171
    kf = (3 * np.pi**2 * rho)**(1 / 3.)
    s2 = sigma / (2 * kf * rho)**2
    return f_slater(rho, sigma) * (1 + kappa - kappa / (1 + mu * s2 / kappa))


def eps_pw92(rs):
    A, a1, b1, b2, b3, b4 = 0.0310906908696549, 0.21370, 7.5957, 3.5876, 1.6382, 0.49294
    q = 2 * A * (b1 * np.sqrt(rs) + b2 * rs + b3 * rs**1.5 + b4 * rs * rs)
    return -2 * A * (1 + a1 * rs) * np.log(1 + 1 / q)


def f_pbe_c(rho, sigma):
    beta, gamma = 0.06672455060314922, (1 - np.log(2.)) / np.pi**2
    rs = (3 / (4 * np.pi * rho))**(1 / 3.)
    ec = eps_pw92(rs)
    kf = (3 * np.pi**2 * rho)**(1 / 3.)
    ks2 = 4 * kf / np.pi
    t2 = sigma / (4 * ks2 * rho * rho)
    A = beta / gamma / (np.exp(-ec / gamma) - 1)
    At2 = A * t2
    return rho * (ec + gamma * np.log(1 + beta / gamma * t2 * (1 + At2) / (1 + At2 + At2 * At2)))


def f_vwn5(rho, sigma, params=(0.0310907, 3.72744, 12.9352, -0.10498)):
    A, b, c, x0 = params
    x = np.sqrt((3 / (4 * np.pi * rho))**(1 / 3.))
    X = x * x + b * x + c
    X0 = x0 * x0 + b * x0 + c
    Q = np.sqrt(4 * c - b * b)
    at = np.arctan(Q / (2 * x + b))
    ec = A * (np.log(x * x / X) + 2 * b / Q * at
              - b * x0 / X0 * (np.log((x - x0)**2 / X) + 2 * (b + 2 * x0) / Q * at))
    return rho * ec


def f_vwn_rpa(rho, sigma):
    """VWN fit to the RPA correlation energy (libxc lda_c_vwn_rpa): the local correlation of libxc's B3LYP."""
    return f_vwn5(rho, sigma, (0.0310907, 13.0720, 42.7198, -0.409286))


def _scaled(fn, w):
    return lambda rho, sigma: w * fn(rho, sigma)


# semilocal parts of the two hybrids (the exact-exchange fraction is the caller's: K from get_jk)
HYBRID = {'pbe0': 0.25, 'b3lyp': 0.20}

FUNCTIONALS = {
    'pbe0': (_scaled(f_pbe_x, 0.75), f_pbe_c),
    'b3lyp': (_scaled(f_slater, 0.08), _scaled(f_b88, 0.72), _scaled(f_lyp, 0.81), _scaled(f_vwn_rpa, 0.19)),
    'lda,vwn': (f_slater, f_vwn5),
    ',vwn': (f_vwn5,),
    'lda,': (f_slater,),
    'b88,': (f_b88,),
    ',lyp': (f_lyp,),
    'blyp': (f_b88, f_lyp),
    'pbe': (f_pbe_x, f_pbe_c),
    'pbe,': (f_pbe_x,),
    ',pbe': (f_pbe_c,),
}
RHO_CUT = 1e-15     # libxc's density threshold: zero below it


def eval_xc(code, rho, sigma):
    """exc (per particle), vrho, vsigma for a closed-shell density (libxc conventions,
    pyscf/dft/libxc.py:eval_xc with spin=0, deriv=1)."""
    rho = np.asarray(rho, dtype=float)
    sigma = np.asarray(sigma, dtype=float)
    ok = rho > RHO_CUT
    r = np.where(ok, rho, 1.0)
    s = np.where(ok, np.maximum(sigma, 1e-40), 1.0)
    f = np.zeros_like(r)
    vr = np.zeros_like(r)
    vs = np.zeros_like(r)
    for fn in FUNCTIONALS[code.lower()]:
        f += fn(r, s)
        vr += fn(r * (1 + 1j * H), s + 0j).imag / (r * H)
        vs += fn(r + 0j, s * (1 + 1j * H)).imag / (s * H)
    return np.where(ok, f / r, 0.0), np.where(ok, vr, 0.0), np.where(ok, vs, 0.0)


def eval_rho(ao, dm, gga=True):
    """rho[4, ng] (density, gradient) for a hermitian dm: numint.py:138-145 (hermi=1)."""
    c0 = ao[0].dot(dm)
    rho = np.empty((4 if gga else 1, ao.shape[1]))
    for i in range(len(rho)):
        rho[i] = np.einsum('pi,pi->p', ao[i].conj(), c0).real
    rho[1:] *= 2
    return rho


def nr_rks(ao, weight, code, dm):
    """(nelec, exc, vmat) for AO values with gradients ao[4, ng, nao] (real at Gamma, complex at a
    k-point) and a hermitian density matrix; GGA branch of numint.py:346-378."""
    rho = eval_rho(ao, dm)
    sigma = np.einsum('xg,xg->g', rho[1:], rho[1:])
    exc, vrho, vsigma = eval_xc(code, rho[0], sigma)
    den = rho[0] * weight
    wv = np.empty_like(rho)
    wv[0] = weight * vrho * .5                       # numint.py:990
    wv[1:] = weight * 2 * vsigma * rho[1:]           # transform_xc: d/d(grad rho) = 2 vsigma grad rho
    aow = np.einsum('npi,np->pi', ao, wv)
    mat = ao[0].conj().T.dot(aow)
    return den.sum(), den.dot(exc), mat + mat.conj().T
