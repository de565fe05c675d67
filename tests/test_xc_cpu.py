"""CPU: the exchange-correlation oracle (oracle/xc.py, AO gradients of oracle/eval_ao.py) against the
reference's fixed numbers (pyscf/pbc/dft/test/test_numint.py) and against published limits."""
import numpy as np
import oracle
from oracle import xc
from pyscf_isdf_b200.cell import Cell

# basis data of the reference's fixtures (pyscf/gto/basis/cc-pvdz.dat, He and C entries)
HE_CCPVDZ = [[0, (38.36, 0.023809), (5.77, 0.154891), (1.24, 0.469987)], [0, (0.2976, 1.0)], [1, (1.275, 1.0)]]
C_CCPVDZ = [[0, (6665.0, 0.000692, -0.000146), (1000.0, 0.005329, -0.001154), (228.0, 0.027077, -0.005725),
             (64.71, 0.101718, -0.023312), (21.06, 0.27474, -0.063955), (7.495, 0.448564, -0.149981),
             (2.797, 0.285074, -0.127262), (0.5215, 0.015204, 0.544529)],
            [0, (0.1596, 1.0)],
            [1, (9.439, 0.038109), (2.002, 0.20948), (0.5456, 0.508557)],
            [1, (0.1517, 1.0)],
            [2, (0.55, 1.0)]]


def _cell(sym, basis):
    return Cell(atom=[[sym, (1., .8, 1.9)], [sym, (.1, .2, .3)]], a=np.eye(3) * 2.5, basis={sym: basis},
                pseudo=None, mesh=[21] * 3, precision=1e-11).build()


def test_ao_gradients_reproduce_the_reference_fingerprints():
    # pyscf/pbc/dft/test/test_numint.py:77-100 (C2, cc-pVDZ: s, p and d shells, two contractions)
    cell = _cell('C', C_CCPVDZ)
    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, deriv=1)
    assert ao.shape == (4, 21**3, 28)
    assert abs(oracle.fp(ao[0]) - (-0.54069672246407219)) < 1e-8
    assert abs(oracle.fp(ao) - 8.8004405892746433) < 1e-8
    # and are the derivatives of the values (central differences, one grid block)
    r = cell.get_uniform_grids()[:56]
    h = 1e-5
    for ax in range(3):
        e = np.zeros(3)
        e[ax] = h
        d = (oracle.eval_ao(cell._atm, cell._bas, cell._env, r + e, Ls, rcut * 2)
             - oracle.eval_ao(cell._atm, cell._bas, cell._env, r - e, Ls, rcut * 2)) / (2 * h)
        assert abs(d - ao[1 + ax, :56]).max() < 1e-6 * max(1.0, abs(d).max())


def test_nr_rks_blyp_reproduces_the_reference_numbers_at_a_kpoint():
    # pyscf/pbc/dft/test/test_numint.py:183-206
    cell = _cell('He', HE_CCPVDZ)
    nao = cell.nao
    rs = np.random.RandomState(1)
    kpts = rs.random_sample((2, 3))
    dms = rs.random_sample((2, nao, nao))
    dms = (dms + dms.transpose(0, 2, 1)) * .5
    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts[:1], deriv=1)[0]
    ne, exc, vmat = xc.nr_rks(ao, cell.vol / ao.shape[1], 'blyp', dms[0])
    assert abs(ne - 5.0499199224525153) < 1e-8
    assert abs(exc - (-3.8870579114663886)) < 1e-8
    assert abs(oracle.fp(vmat) - (0.42538491159934377 + 0.14139753327162483j)) < 1e-8
    assert abs(vmat - vmat.conj().T).max() < 1e-14


def test_functional_derivatives_and_published_limits():
    rng = np.random.default_rng(3)
    rho = 10.0**rng.uniform(-4, 1, 200)
    sigma = rho**(8 / 3.) * 10.0**rng.uniform(-3, 2, 200)
    for code in ('blyp', 'pbe', 'lda,', 'lda,vwn'):
        exc, vr, vs = xc.eval_xc(code, rho, sigma)
        f = lambda r, s: xc.eval_xc(code, r, s)[0] * r            # noqa: E731
        hr, hs = 1e-5 * rho, 1e-5 * sigma
        assert abs((f(rho + hr, sigma) - f(rho - hr, sigma)) / (2 * hr) - vr).max() < 1e-6 * abs(vr).max()
        dvs = (f(rho, sigma + hs) - f(rho, sigma - hs)) / (2 * hs)
        assert (abs(dvs - vs) < 1e-5 * abs(vs) + 1e-10 * abs(f(rho, sigma)) / hs).all()      # second term: rounding of the difference quotient
    # PBE (no fixed number in the reference): uniform-gas limit, the enhancement factor at s = 1 and
    # its Lieb-Oxford ceiling 1 + kappa, correlation switched off by a large gradient
    r = np.array([0.3])
    kf = (3 * np.pi**2 * r)**(1 / 3.)
    assert abs(xc.f_pbe_x(r, 0 * r) - xc.f_slater(r, 0 * r)) < 1e-15
    assert abs(xc.f_pbe_x(r, (2 * kf * r)**2) / xc.f_slater(r, r) - (1.804 - 0.804 / (1 + 0.2195149727645171 / 0.804))) < 1e-14
    assert abs(xc.f_pbe_x(r, 1e12 * r) / xc.f_slater(r, r) - 1.804) < 1e-6
    rs_ = (3 / (4 * np.pi * r))**(1 / 3.)
    assert abs(xc.f_pbe_c(r, 1e-30 * r) - r * xc.eps_pw92(rs_)) < 1e-14
    assert abs(xc.f_pbe_c(r, 1e14 * r)) < 1e-6
    # PW92 at rs = 1 and 2 (Perdew & Wang 1992, table: -0.0598, -0.0448 Ha to the digits given)
    assert abs(xc.eps_pw92(np.array([1.0, 2.0])) - np.array([-0.0598, -0.0448])).max() < 1.5e-4
    # VWN (form V) against PW92: two fits of the same Monte-Carlo data, within 6e-4 Ha over rs = 0.5 .. 20;
    # and its own published high-density coefficient A ln rs: d eps / d ln rs -> A = 0.0310907 as rs -> 0
    rs_ = np.array([0.5, 1.0, 2.0, 5.0, 10.0, 20.0])
    rr = 3 / (4 * np.pi * rs_**3)
    assert abs(xc.f_vwn5(rr, 0 * rr) / rr - xc.eps_pw92(rs_)).max() < 6e-4
    e = lambda x: xc.f_vwn5(np.array([3 / (4 * np.pi * x**3)]), 0.0)[0] / (3 / (4 * np.pi * x**3))      # noqa: E731
    assert abs((e(2e-6) - e(1e-6)) / np.log(2.0) - 0.0310907) < 2e-4
    # VWN fit to the RPA correlation energy (the local correlation of libxc's B3LYP) against the RPA values
    # of the electron gas at rs = 1, 2, 5, 10 (-0.157, -0.124, -0.0844, -0.0614 Ry)
    rpa = np.array([-0.0785, -0.0619, -0.0422, -0.0307])
    r4 = 3 / (4 * np.pi * np.array([1.0, 2.0, 5.0, 10.0])**3)
    assert abs(xc.f_vwn_rpa(r4, 0 * r4) / r4 - rpa).max() < 1.5e-3
    # the hybrids' semilocal parts are the stated mixtures
    rr_, ss_ = np.array([0.2, 1.1]), np.array([0.05, 2.0])
    e_pbe0 = xc.eval_xc('pbe0', rr_, ss_)[0]
    assert abs(e_pbe0 - (0.75 * xc.eval_xc('pbe,', rr_, ss_)[0] + xc.eval_xc(',pbe', rr_, ss_)[0])).max() < 1e-15
    assert xc.HYBRID == {'pbe0': 0.25, 'b3lyp': 0.20}
    # densities below the threshold give zeros, not NaNs
    e0, v0, s0 = xc.eval_xc('pbe', np.array([0.0, 1e-20]), np.array([0.0, 1e-30]))
    assert (e0 == 0).all() and (v0 == 0).all() and (s0 == 0).all()


def test_kmesh_nr_rks_through_the_supercell_matches_the_kpoint_sum():
    """ISDF.nr_rks on a uniform k-point mesh (KNumInt.nr_rks, pyscf/pbc/dft/numint.py:1090-1204: rho = 1/Nk sum_k
    rho_k, electrons and energy per cell, vmat_k = phi_k^H aow): served by the supercell's Gamma-point nr_rks and
    folded back (pyscf_isdf_b200/kmesh.py).  Here the supercell object is a stand-in that runs the oracle, so the
    product's own routing, Bloch transformations and 1/Nk factors are what is checked."""
    from pyscf_isdf_b200.isdf import ISDF
    cell = _cell('He', HE_CCPVDZ)
    cell = Cell(atom=cell.atom, a=cell.a, basis=cell.basis, pseudo=None, unit=cell.unit, mesh=[9, 9, 9]).build()
    nao = cell.nao
    kpts = cell.make_kpts([1, 3, 1])
    rs = np.random.RandomState(4)
    dms = rs.random_sample((3, nao, nao)) + 1j * rs.random_sample((3, nao, nao))
    dms = dms + dms.conj().transpose(0, 2, 1)

    mydf = ISDF(cell, kpts=kpts)
    mydf.check_sanity()
    sc = mydf._kmesh.supercell
    rcut, Ls = sc.ao_eval_params()
    ao_sc = oracle.eval_ao(sc._atm, sc._bas, sc._env, sc.get_uniform_grids(), Ls, rcut, deriv=1)

    class SupercellStandIn:
        def nr_rks(self, code, dm, kpts=None, max_bytes=None):
            assert dm.dtype == np.float64 and abs(dm - dm.T).max() < 1e-12
            return xc.nr_rks(ao_sc, sc.vol / ao_sc.shape[1], code, dm)

    mydf._sc = SupercellStandIn()
    ne, exc, vmat = mydf.nr_rks('pbe', dms, kpts)

    rcut, Ls = cell.ao_eval_params()
    ao = oracle.eval_ao(cell._atm, cell._bas, cell._env, cell.get_uniform_grids(), Ls, rcut, kpts=kpts, deriv=1)
    w = cell.vol / ao.shape[2]
    rho = sum(xc.eval_rho(ao[k], dms[k]) for k in range(3)) / 3
    sigma = np.einsum('xg,xg->g', rho[1:], rho[1:])
    e, vrho, vsigma = xc.eval_xc('pbe', rho[0], sigma)
    wv = np.empty_like(rho)
    wv[0] = w * vrho * .5
    wv[1:] = w * 2 * vsigma * rho[1:]
    assert abs(ne - w * rho[0].sum()) < 1e-9 and abs(exc - w * rho[0].dot(e)) < 1e-9
    for k in range(3):
        mat = ao[k][0].conj().T.dot(np.einsum('npi,np->pi', ao[k], wv))
        assert abs(vmat[k] - (mat + mat.conj().T)).max() < 1e-8
