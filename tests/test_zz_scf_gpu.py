"""GPU: SCF energies through the product path.  The reference's Gamma-point RHF fixture
(pyscf/pbc/scf/test/test_hf.py:30-43) run with ISDF.get_jk and ISDF.get_nuc inside the test harness's
Roothaan loop: with n_ip >= the rank of the pair space (21) ISDF is exact, so the energies must equal the
reference's fixed numbers (test_hf.py:94-95, :129-132).  Tolerance 1e-7 Ha (the reference asserts 7 places;
the oracle chain itself sits 1.1e-8 from them, tests/test_scf_cpu.py); overlap and kinetic matrices come
from pyscf_isdf_b200/int1e.py (in the reference they come from libcint, outside the ISDF path)."""
import numpy as np
import pytest
from pyscf_isdf_b200 import int1e
from pyscf_isdf_b200.isdf import ISDF
from scf_harness import rhf_gamma
from test_scf_cpu import E_EXX_EWALD, E_EXX_NONE, he2_cell

pytestmark = pytest.mark.gpu


def test_rhf_energies_with_isdf_in_the_exact_limit():
    cell = he2_cell()
    S, T = int1e.overlap_kinetic(cell)
    mydf = ISDF(cell, n_ip=60).build()
    assert mydf.n_ip <= 22
    hcore = T + mydf.get_nuc()
    enuc = cell.ewald()
    for exx, gold in ((None, E_EXX_NONE), ('ewald', E_EXX_EWALD)):
        e, dm, ncyc = rhf_gamma(hcore, S, cell.nelectron, enuc, lambda dm: mydf.get_jk(dm, exxdiv=exx))
        assert abs(e - gold) < 1e-7, (exx, e - gold)


def _rks_diamond(xc_on_device):
    """RKS 'lda,vwn' on the reference's cubic diamond fixture (pyscf/pbc/dft/test/test_krks.py:90-97) with the
    product's get_pp (local part on the device, projectors on the host) and exact-density J on the device;
    exchange-correlation from ISDF.nr_rks on the device, or from the oracle."""
    from oracle import xc
    from scf_harness import rks_gamma
    from test_scf_cpu import E_LDA_C8_CUBIC, diamond_cubic, oracle_rks_pieces
    cell = diamond_cubic([17] * 3)
    ao4 = oracle_rks_pieces(cell)[2]
    S, T = int1e.overlap_kinetic(cell)
    mydf = ISDF(cell, c=10, exact_j=True).build()
    hcore = T + mydf.get_pp()
    w = cell.vol / ao4.shape[1]
    nr_rks = (lambda dm: mydf.nr_rks('lda,vwn', dm)) if xc_on_device else (lambda dm: xc.nr_rks(ao4, w, 'lda,vwn', dm))
    e, dm, ncyc = rks_gamma(hcore, S, cell.nelectron, cell.ewald(), lambda dm: mydf.get_j(dm), nr_rks)
    assert abs(e - E_LDA_C8_CUBIC) < 1e-7, e - E_LDA_C8_CUBIC


def test_rks_lda_energy_of_diamond_with_device_hcore_and_j():
    _rks_diamond(False)


def test_rks_lda_energy_of_diamond_all_on_device():
    _rks_diamond(True)


def test_rhf_energy_of_diamond_with_pseudopotential_through_the_plugin():
    """pyscf/pbc/scf/test/test_newton.py:52-59: RHF, exxdiv='ewald', diamond primitive cell, gth-szv/gth-pade.
    ISDF with n_ip above the pair rank (36) is exact; get_pp supplies local + non-local parts."""
    from test_scf_cpu import E_RHF_C2_PRIMITIVE, diamond_primitive_bohr
    cell = diamond_primitive_bohr([19] * 3)
    S, T = int1e.overlap_kinetic(cell)
    mydf = ISDF(cell, n_ip=60).build()
    assert mydf.n_ip <= 36
    e = rhf_gamma(T + mydf.get_pp(), S, cell.nelectron, cell.ewald(), lambda dm: mydf.get_jk(dm, exxdiv='ewald'))[0]
    assert abs(e - E_RHF_C2_PRIMITIVE) < 1e-7, e - E_RHF_C2_PRIMITIVE
