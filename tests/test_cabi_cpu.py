"""CPU: the C-ABI library builds, loads and exports every declared symbol;
the host-side plugin surface behaves like the reference's with_df objects."""
import ctypes
import os
import re
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_header_symbol():
    from pyscf_isdf_b200 import _lib
    lib = _lib.load_library()
    hdr = open(os.path.join(ROOT, 'include', 'isdf_b200.h')).read()
    names = set(re.findall(r'\b(isdf_[a-z_A-Z0-9]+)\s*\(', hdr))
    assert names, 'no prototypes parsed'
    for n in sorted(names):
        assert hasattr(lib, n), 'missing export %s' % n
        assert n in _lib.SIGNATURES, 'no ctypes prototype for %s' % n
    # workspace queries need no GPU
    assert lib.isdf_get_jk_ws_bytes(104, 1560) > 8 * 1560 * 1560
    mesh = (ctypes.c_int * 3)(35, 35, 35)
    assert lib.isdf_build_W_ws_bytes(10, mesh) >= 16 * 10 * 35 * 35 * 18
    assert lib.isdf_select_ip_ws_bytes(104, 42875, 1560) > 0
    assert lib.isdf_fit_aux_ws_bytes(104, 42875, 1560) > 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from pyscf_isdf_b200._lib import Handle
    with pytest.raises(RuntimeError):
        Handle(0)
    from util import diamond_c8
    from pyscf_isdf_b200.isdf import ISDF
    mydf = ISDF(diamond_c8([9, 9, 9], basis='gth-szv'), c=3)
    with pytest.raises(RuntimeError):
        mydf.get_jk(np.eye(mydf.cell.nao))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'pyscf_isdf_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r'^\s*(from|import)\s+oracle', src, re.M), fn


def test_plugin_surface_errors():
    from util import diamond_c8
    from pyscf_isdf_b200.isdf import ISDF
    cell = diamond_c8([9, 9, 9], basis='gth-szv')
    mydf = ISDF(cell, c=3)
    assert list(mydf.mesh) == [9, 9, 9]
    assert mydf.kpts.shape == (1, 3)
    dm = np.eye(cell.nao)
    with pytest.raises(NotImplementedError):
        mydf.get_jk(dm, omega=0.3)
    with pytest.raises(NotImplementedError):
        mydf.get_jk(dm, kpts_band=np.zeros(3))
    with pytest.raises(NotImplementedError):
        mydf.get_jk(dm, kpts=np.array([0.1, 0, 0]))
    with pytest.raises(NotImplementedError):
        mydf.get_jk(dm, exxdiv='vcut_sph')
    with pytest.raises(NotImplementedError):
        ISDF(cell, select='qr').check_sanity()

    class FakeMF:
        def __init__(self):
            self._eri = 1
            self.with_df = None

        def copy(self):
            import copy
            return copy.copy(self)

        def _is_mem_enough(self):
            return True
    mf = mydf.update_mf(FakeMF())
    assert mf.with_df is mydf and mf._eri is None and mf._is_mem_enough() is False
    # the rest of FFTDF's surface (pyscf/pbc/df/fft.py:246-272, :317-359) that pbc.scf can reach on with_df
    # (hf.py:676-679 get_ao_eri when the in-core shortcut is not patched; post-SCF ao2mo; molecular DF.loop):
    # a reason, not an AttributeError
    for name, args in (('get_ao_eri', ()), ('get_eri', ()), ('ao2mo', (np.eye(cell.nao),)), ('get_mo_eri', (dm,)),
                       ('ao2mo_7d', (dm,)), ('get_ao_pairs_G', ()), ('get_ao_pairs', ()), ('get_mo_pairs_G', (dm,)),
                       ('get_mo_pairs', (dm,)), ('aoR_loop', ()), ('loop', ()), ('range_coulomb', (0.3,)),
                       ('get_jk_e1', (dm,)), ('get_j_e1', (dm,)), ('get_k_e1', (dm,))):
        with pytest.raises(NotImplementedError, match='ISDF'):
            getattr(mydf, name)(*args)
    with pytest.raises(NotImplementedError):
        mydf.grids
    assert mydf.get_naoaux() == 3 * cell.nao
    with pytest.raises(NotImplementedError):
        mydf.get_nuc(np.array([0.1, 0.0, 0.0]))
    with pytest.raises(NotImplementedError):
        mydf.nr_rks('pbe', dm, kpts=np.array([0.1, 0.0, 0.0]))
    with pytest.raises(RuntimeError):
        mydf.save('/tmp/never_written.npz')            # nothing built yet
    # tagged density matrices: which ones take the occupied-orbital form
    C = np.linalg.qr(np.random.default_rng(0).standard_normal((cell.nao, cell.nao)))[0]
    occ = np.zeros(cell.nao); occ[:4] = 2
    tags = ISDF._mo_tags
    class T(np.ndarray):
        pass
    t = dm.view(T); t.mo_coeff, t.mo_occ = C, occ
    assert tags(t, dm, 1, cell.nao) is not None
    assert tags(t, dm, 2, cell.nao) is None                              # a set of matrices
    assert tags(dm, dm, 1, cell.nao) is None                             # no tags
    t.mo_coeff = C.astype(complex)
    assert tags(t, dm, 1, cell.nao) is None                              # complex orbitals
    t.mo_coeff, t.mo_occ = C[None], occ[None]                            # k-point style leading axis of length 1
    assert tags(t, dm, 1, cell.nao) is not None


def test_cell_model_matches_reference_conventions():
    from util import diamond_c8
    from pyscf_isdf_b200.cell import madelung, super_cell
    cell = diamond_c8([35, 35, 35])
    assert cell.nao == 104 and cell.nelectron == 32 and cell.nbas == 24
    # normalised contracted GTOs: self-overlap of every contraction is 1
    from pyscf_isdf_b200.cell import gaussian_int
    for ib in range(3):
        l, es, cs = cell.bas_angular(ib), cell.bas_exp(ib), cell._libcint_ctr_coeff(ib)
        ee = gaussian_int(2 * l + 2, es[:, None] + es[None, :])
        assert abs(np.einsum('pi,pq,qi->i', cs, ee, cs) - 1).max() < 1e-12
    # Madelung constant of a simple cubic lattice: 2.837297479 / L
    L = cell.lattice_vectors()[0, 0]
    assert abs(madelung(cell) - 2.837297479480619 / L) < 1e-7
    sup = super_cell(cell, [2, 2, 2])
    assert sup.natm == 64 and sup.nao == 832 and list(sup.mesh) == [71, 71, 71]
    coords = cell.get_uniform_grids([4, 4, 4])
    assert coords.shape == (64, 3) and abs(coords[1] - [0, 0, L / 4]).max() < 1e-12


def test_reference_arm_prints_the_contract_line():
    """bench.py --impl reference (the CPU oracle arm) on the smoke workload: one JSON line with the keys
    the driver reads, a bounded sample described in cpu_baseline, no GPU involved."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, 'bench.py'), '--impl', 'reference', '--workload', 'c8-small',
                          '--steps', '1', '--warmup', '0', '--c', '6'], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ('impl', 'metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better',
                'scaling', 'vs_baseline', 'dtype', 'data', 'config', 'cpu_baseline', 'e2e'):
        assert key in line, key
    assert line['impl'] == 'reference' and line['higher_is_better'] is False and line['value'] > 0
    assert line['cpu_baseline']['kind'] == 'port' and line['cpu_baseline']['cores'] >= 1
    assert line['e2e']['h2d_bytes_per_step'] == 0 and 'workload' in line['config']
