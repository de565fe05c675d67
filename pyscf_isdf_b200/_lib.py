"""ctypes binding of libisdf_b200.so (the C-ABI in include/isdf_b200.h).

Mirrors how the reference reaches its native AO kernel
(numpy.ctypeslib.load_library + raw pointers,
pyscf/pbc/gto/eval_gto.py:139-151, pyscf/lib/misc.py:91-103).  There is no
CPU fallback: if the CUDA library is missing or no B200 is visible, every
entry point raises.
"""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libisdf_b200.so')

c_int, c_ll, c_dbl, c_vp, c_sz = (ctypes.c_int, ctypes.c_longlong, ctypes.c_double,
                                  ctypes.c_void_p, ctypes.c_size_t)

# name -> (restype, argtypes); must list every symbol declared in include/isdf_b200.h
SIGNATURES = {
    'isdf_create': (c_int, [c_int, ctypes.POINTER(c_vp)]),
    'isdf_destroy': (c_int, [c_vp]),
    'isdf_set_stream': (c_int, [c_vp, c_vp]),
    'isdf_last_error': (ctypes.c_char_p, []),
    'isdf_num_sms': (c_int, [c_vp]),
    'isdf_launch_count': (c_ll, [c_vp, c_int]),
    'isdf_stage_time_ms': (c_int, [c_vp, c_int, ctypes.POINTER(ctypes.c_float)]),
    'isdf_eval_ao': (c_int, [c_vp, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_vp,
                             c_vp, c_int, c_vp, c_ll, c_ll, c_vp, c_ll]),
    'isdf_select_ip_ws_bytes': (c_sz, [c_int, c_ll, c_int]),
    'isdf_select_ip': (c_int, [c_vp, c_vp, c_int, c_ll, c_ll, c_int, c_dbl, c_vp, c_ll,
                               c_vp, c_vp, c_vp, c_vp, c_sz]),
    'isdf_select_max_rows': (c_int, []),
    'isdf_select_panel_width': (c_int, []),
    'isdf_select_last_panel_mode': (c_int, []),
    'isdf_select_row_len': (c_int, [c_int, c_int]),
    'isdf_select_first_quota': (c_int, [c_int]),
    'isdf_select_next_quota': (c_int, [c_int, c_int, c_int, c_int]),
    'isdf_select_create': (c_int, [c_vp, c_int, c_ll, c_ll, c_ll, c_int, c_vp, c_sz, ctypes.POINTER(c_vp)]),
    'isdf_select_destroy': (c_int, [c_vp]),
    'isdf_select_begin': (c_int, [c_vp, c_vp, c_ll, ctypes.POINTER(c_dbl)]),
    'isdf_select_set_tol': (c_int, [c_vp, c_dbl]),
    'isdf_select_candidates': (c_int, [c_vp, c_vp, c_ll, c_vp, c_ll, c_int, c_int, c_vp]),
    'isdf_select_panel': (c_int, [c_vp, c_vp, c_int, c_int, c_int, c_int, c_vp, c_ll,
                                  ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    'isdf_select_flush': (c_int, [c_vp, c_vp, c_ll, c_vp, c_ll, c_int, c_int]),
    'isdf_select_finish': (c_int, [c_vp, c_int, c_vp, c_vp]),
    'isdf_fit_from_U_ws_bytes': (c_sz, [c_ll, c_int]),
    'isdf_fit_aux_from_U': (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_ll, c_ll, c_vp, c_sz]),
    'isdf_fft_zy': (c_int, [c_vp, c_vp, c_ll, c_int, c_int, c_vp, c_vp, c_ll]),
    'isdf_fft_pack': (c_int, [c_vp, c_vp, c_ll, c_int, c_int, c_int, c_vp, c_int, c_vp]),
    'isdf_fft_unpack': (c_int, [c_vp, c_vp, c_int, c_vp, c_int, c_int, c_vp, c_ll]),
    'isdf_fft_x_coulomb': (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_vp, c_int, c_int]),
    'isdf_symmetrize_upper': (c_int, [c_vp, c_vp, c_ll, c_int]),
    'isdf_syrk_i8_ws_bytes': (c_sz, [c_int, c_ll, c_int]),
    'isdf_syrk_i8_last_ops': (c_dbl, []),
    'isdf_syrk_i8': (c_int, [c_vp, c_int, c_ll, c_vp, c_ll, c_vp, c_ll, c_dbl, c_int, c_vp, c_sz]),
    'isdf_grid_j_ws_bytes': (c_sz, [c_int, c_ll]),
    'isdf_density_on_grid': (c_int, [c_vp, c_vp, c_ll, c_int, c_ll, c_vp, c_int, c_vp, c_ll, c_vp, c_sz]),
    'isdf_j_from_potential': (c_int, [c_vp, c_vp, c_ll, c_int, c_ll, c_vp, c_ll, c_int, c_dbl, c_vp, c_vp, c_sz]),
    'isdf_eval_ao_deriv1': (c_int, [c_vp, c_vp, c_int, c_vp, c_int, c_vp, c_int, c_vp, c_vp,
                                    c_vp, c_int, c_vp, c_ll, c_ll, c_vp, c_ll, c_ll]),
    'isdf_xc_code': (c_int, [ctypes.c_char_p]),
    'isdf_xc_hybrid_fraction': (c_dbl, [c_int]),
    'isdf_xc_ws_bytes': (c_sz, [c_int, c_ll]),
    'isdf_xc_density': (c_int, [c_vp, c_vp, c_ll, c_ll, c_int, c_ll, c_vp, c_vp, c_ll, c_vp, c_sz]),
    'isdf_xc_potential': (c_int, [c_vp, c_int, c_vp, c_ll, c_ll, c_dbl, c_vp, c_vp, c_vp, c_sz]),
    'isdf_xc_matrix': (c_int, [c_vp, c_vp, c_ll, c_ll, c_int, c_ll, c_vp, c_ll, c_vp, c_vp, c_sz]),
    'isdf_get_jk_block_ws_bytes': (c_sz, [c_int, c_int, c_int]),
    'isdf_get_jk_block': (c_int, [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_vp, c_int, c_int, c_int,
                                  c_int, c_int, c_vp, c_vp, c_vp, c_sz]),
    'isdf_fit_aux_ws_bytes': (c_sz, [c_int, c_ll, c_int]),
    'isdf_fit_aux': (c_int, [c_vp, c_vp, c_int, c_ll, c_ll, c_vp, c_int, c_vp, c_ll, c_vp, c_sz]),
    'isdf_fit_aux_from_factor': (c_int, [c_vp, c_vp, c_ll, c_ll, c_vp, c_int, c_vp, c_sz]),
    'isdf_build_W_ws_bytes': (c_sz, [c_int, c_vp]),
    'isdf_build_W': (c_int, [c_vp, c_vp, c_int, c_ll, c_vp, c_vp, c_vp, c_vp, c_ll, c_vp, c_sz]),
    'isdf_coulomb_apply': (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_vp, c_vp, c_ll, c_vp, c_sz]),
    'isdf_get_jk_ws_bytes': (c_sz, [c_int, c_int]),
    'isdf_get_jk': (c_int, [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int,
                            c_vp, c_vp, c_vp, c_sz]),
    'isdf_get_jk_block_occ': (c_int, [c_vp, c_vp, c_vp, c_ll, c_int, c_int, c_vp, c_int, c_int, c_int,
                                      c_int, c_int, c_vp, c_vp, c_vp, c_sz]),
    'isdf_get_jk_static_max_bytes': (c_sz, []),
    'isdf_get_jk_static': (c_int, [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    'isdf_phase_expand': (c_int, [c_vp, c_vp, c_vp, c_ll, c_int, c_vp, c_ll, c_vp, c_dbl, c_vp, c_vp, c_ll]),
    'isdf_congruence': (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_ll, c_vp, c_sz]),
    'isdf_transpose': (c_int, [c_vp, c_vp, c_ll, c_int, c_int, c_vp, c_ll]),
    'isdf_dgemm': (c_int, [c_vp, c_int, c_int, c_int, c_vp, c_ll, c_int, c_vp, c_ll, c_int,
                           c_vp, c_ll, c_vp, c_ll, c_vp, c_ll, c_dbl, c_int, c_int, c_int]),
    'isdf_set_gemm_tile': (c_int, [c_vp, c_int]),
    'isdf_gemm_profile': (c_int, [c_vp, c_int]),
    'isdf_gemm_profile_read': (c_int, [c_vp, ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl),
                                       ctypes.POINTER(c_ll)]),
    'isdf_gather_cols': (c_int, [c_vp, c_vp, c_ll, c_int, c_vp, c_int, c_vp, c_ll]),
}

STAGE_AO, STAGE_SELECT, STAGE_FIT, STAGE_COULOMB, STAGE_JK, STAGE_FFT, STAGE_GEMM = range(7)

_lib = None


def load_library():
    """Load libisdf_b200.so and declare prototypes; raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            'libisdf_b200.so not found at %s: build it with '
            '`python -c "import __graft_entry__ as g; g.build()"` '
            '(there is no CPU fallback for the ISDF hot path)' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load_library().isdf_last_error()
        raise RuntimeError('libisdf_b200 error %d: %s' % (rc, msg.decode() if msg else '?'))


def hptr(a):
    """host pointer of a C-contiguous numpy array"""
    assert a.flags.c_contiguous
    return a.ctypes.data_as(c_vp)


def pad8(n):
    return (int(n) + 7) // 8 * 8


class Handle:
    """Owns one isdf_handle bound to one GPU (one process per GPU)."""

    def __init__(self, device=None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError('pyscf_isdf_b200 needs a CUDA device (B200, sm_100a); '
                               'there is no CPU fallback')
        self.lib = load_library()
        if device is None:
            device = torch.cuda.current_device()
        self.device = int(device)
        self.torch_device = torch.device('cuda', self.device)
        h = c_vp()
        check(self.lib.isdf_create(self.device, ctypes.byref(h)))
        self.h = h
        with torch.cuda.device(self.device):
            self.stream = torch.cuda.current_stream()
        check(self.lib.isdf_set_stream(self.h, c_vp(self.stream.cuda_stream)))
        if os.environ.get('ISDF_GEMM_TILE'):          # tuning knob, see include/isdf_b200.h
            self.set_gemm_tile(int(os.environ['ISDF_GEMM_TILE']))

    def __del__(self):
        try:
            if getattr(self, 'h', None):
                self.lib.isdf_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # -- helpers ---------------------------------------------------------------
    def empty(self, *shape, dtype=None):
        import torch
        return torch.empty(*shape, dtype=dtype or torch.float64, device=self.torch_device)

    def zeros(self, *shape, dtype=None):
        import torch
        return torch.zeros(*shape, dtype=dtype or torch.float64, device=self.torch_device)

    def empty_complex(self, *shape):
        import torch
        return torch.empty(*shape, dtype=torch.complex128, device=self.torch_device)

    @staticmethod
    def as_real(t):
        """complex128 [.., n] -> float64 view [.., 2n] (same memory)"""
        import torch
        v = torch.view_as_real(t)
        return v.reshape(*t.shape[:-1], 2 * t.shape[-1])

    @staticmethod
    def as_complex(t):
        """float64 [m, 2n] (contiguous rows) -> complex128 view [m, n] (same memory)"""
        import torch
        return torch.view_as_complex(t.view(t.shape[0], t.shape[1] // 2, 2))

    def from_host(self, arr):
        """float64 host array -> device tensor (through pinned memory, asynchronous on the current stream)."""
        import torch
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64))
        return t.pin_memory().to(self.torch_device, non_blocking=True)

    def index(self, idx):
        import torch
        return torch.as_tensor(np.asarray(idx, dtype=np.int64), device=self.torch_device)

    def synchronize(self):
        import torch
        torch.cuda.synchronize(self.torch_device)

    def timer(self):
        return StageTimer(self)

    def device_mem_bytes(self):
        import torch
        return int(torch.cuda.get_device_properties(self.device).total_memory)

    def workspace(self, nbytes):
        import torch
        return torch.empty(int(nbytes), dtype=torch.uint8, device=self.torch_device)

    def stage_ms(self, stage):
        out = ctypes.c_float()
        check(self.lib.isdf_stage_time_ms(self.h, stage, ctypes.byref(out)))
        return out.value

    def set_gemm_tile(self, which):
        check(self.lib.isdf_set_gemm_tile(self.h, int(which)))

    def gemm_profile(self, enable):
        check(self.lib.isdf_gemm_profile(self.h, int(enable)))

    def gemm_profile_read(self):
        """(algorithmic flops, summed kernel ms, launches) since gemm_profile(True)"""
        fl, ms, n = c_dbl(), c_dbl(), c_ll()
        check(self.lib.isdf_gemm_profile_read(self.h, ctypes.byref(fl), ctypes.byref(ms), ctypes.byref(n)))
        return fl.value, ms.value, n.value

    def launches(self):
        return (self.lib.isdf_launch_count(self.h, 0), self.lib.isdf_launch_count(self.h, 1))

    # -- stage wrappers (thin; shapes documented in include/isdf_b200.h) ------
    def eval_ao(self, atm, bas, env, a, mesh, Ls, rcut, g0, g1, phi, ldphi):
        atm = np.ascontiguousarray(atm, dtype=np.int32)
        bas = np.ascontiguousarray(bas, dtype=np.int32)
        env = np.ascontiguousarray(env, dtype=np.float64)
        a = np.ascontiguousarray(a, dtype=np.float64)
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        Ls = np.ascontiguousarray(Ls, dtype=np.float64)
        rcut = np.ascontiguousarray(rcut, dtype=np.float64)
        check(self.lib.isdf_eval_ao(self.h, hptr(atm), len(atm), hptr(bas), len(bas), hptr(env), len(env),
                                    hptr(a), hptr(mesh), hptr(Ls), len(Ls), hptr(rcut),
                                    g0, g1, c_vp(phi.data_ptr()), ldphi))

    def eval_ao_deriv1(self, atm, bas, env, a, mesh, Ls, rcut, g0, g1, phi4, ldphi, comp_stride):
        """phi4: device tensor [4, nao, ldphi] (value, d/dx, d/dy, d/dz planes)."""
        atm = np.ascontiguousarray(atm, dtype=np.int32)
        bas = np.ascontiguousarray(bas, dtype=np.int32)
        env = np.ascontiguousarray(env, dtype=np.float64)
        a = np.ascontiguousarray(a, dtype=np.float64)
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        Ls = np.ascontiguousarray(Ls, dtype=np.float64)
        rcut = np.ascontiguousarray(rcut, dtype=np.float64)
        check(self.lib.isdf_eval_ao_deriv1(self.h, hptr(atm), len(atm), hptr(bas), len(bas), hptr(env), len(env),
                                           hptr(a), hptr(mesh), hptr(Ls), len(Ls), hptr(rcut),
                                           g0, g1, c_vp(phi4.data_ptr()), ldphi, comp_stride))

    def xc_code(self, name):
        code = self.lib.isdf_xc_code(str(name).encode())
        if code < 0:
            raise NotImplementedError('xc functional %r (have: lda, / lda,vwn / blyp / pbe / pbe0 / b3lyp)' % (name,))
        return code

    def xc_hybrid_fraction(self, name):
        return float(self.lib.isdf_xc_hybrid_fraction(self.xc_code(name)))

    def xc_density(self, phi4, ldphi, comp_stride, nao, ngl, dm, rho4, ldr, ws=None):
        nbytes = self.lib.isdf_xc_ws_bytes(nao, ldphi)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_xc_density(self.h, c_vp(phi4.data_ptr()), ldphi, comp_stride, nao, ngl, c_vp(dm.data_ptr()),
                                       c_vp(rho4.data_ptr()), ldr, c_vp(ws.data_ptr()), nbytes))
        return ws

    def xc_potential(self, code, rho4, ldr, ngl, weight, wv4, sums, ws):
        check(self.lib.isdf_xc_potential(self.h, int(code), c_vp(rho4.data_ptr()), ldr, ngl, float(weight),
                                         c_vp(wv4.data_ptr()), c_vp(sums.data_ptr()), c_vp(ws.data_ptr()),
                                         ws.numel() * ws.element_size()))

    def xc_matrix(self, phi4, ldphi, comp_stride, nao, ngl, wv4, ldr, vmat, ws):
        check(self.lib.isdf_xc_matrix(self.h, c_vp(phi4.data_ptr()), ldphi, comp_stride, nao, ngl,
                                      c_vp(wv4.data_ptr()), ldr, c_vp(vmat.data_ptr()), c_vp(ws.data_ptr()),
                                      ws.numel() * ws.element_size()))

    def select_ip(self, phi, nao, ng, ldphi, npts, tol, Lt, ldl):
        nbytes = self.lib.isdf_select_ip_ws_bytes(nao, ng, npts)
        ws = self.workspace(nbytes)
        ip = np.zeros(npts, dtype=np.int64)
        resid = np.zeros(npts, dtype=np.float64)
        rank = c_int(0)
        check(self.lib.isdf_select_ip(self.h, c_vp(phi.data_ptr()), nao, ng, ldphi, npts, float(tol),
                                      c_vp(Lt.data_ptr()), ldl, hptr(ip), hptr(resid),
                                      ctypes.byref(rank), c_vp(ws.data_ptr()), nbytes))
        r = rank.value
        return ip[:r].copy(), resid[:r].copy()

    def fit_aux(self, phi, nao, ng, ldphi, ip, zeta, ldz):
        ip = np.ascontiguousarray(ip, dtype=np.int64)
        nbytes = self.lib.isdf_fit_aux_ws_bytes(nao, ng, len(ip))
        ws = self.workspace(nbytes)
        check(self.lib.isdf_fit_aux(self.h, c_vp(phi.data_ptr()), nao, ng, ldphi, hptr(ip), len(ip),
                                    c_vp(zeta.data_ptr()), ldz, c_vp(ws.data_ptr()), nbytes))

    def fit_aux_from_factor(self, Lt, ldl, ng, ip):
        ip = np.ascontiguousarray(ip, dtype=np.int64)
        nbytes = self.lib.isdf_fit_aux_ws_bytes(1, ng, len(ip))
        ws = self.workspace(nbytes)
        check(self.lib.isdf_fit_aux_from_factor(self.h, c_vp(Lt.data_ptr()), ldl, ng, hptr(ip), len(ip),
                                                c_vp(ws.data_ptr()), nbytes))

    def build_W(self, zeta, npts, ldz, mesh, a, W, V=None, ldv=0):
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        a = np.ascontiguousarray(a, dtype=np.float64)
        nbytes = self.lib.isdf_build_W_ws_bytes(npts, hptr(mesh))
        ws = self.workspace(nbytes)
        check(self.lib.isdf_build_W(self.h, c_vp(zeta.data_ptr()), npts, ldz, hptr(mesh), hptr(a),
                                    c_vp(W.data_ptr()), c_vp(V.data_ptr()) if V is not None else None, ldv,
                                    c_vp(ws.data_ptr()), nbytes))

    def coulomb_apply(self, vin, ldin, nvec, mesh, a, vout, ldout, ws_bytes=None):
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        a = np.ascontiguousarray(a, dtype=np.float64)
        ngh = int(mesh[0]) * int(mesh[1]) * (int(mesh[2]) // 2 + 1)
        if ws_bytes is None:
            ws_bytes = 16 * ngh * min(nvec, 64)
        ws = self.workspace(ws_bytes)
        check(self.lib.isdf_coulomb_apply(self.h, c_vp(vin.data_ptr()), ldin, nvec, hptr(mesh), hptr(a),
                                          c_vp(vout.data_ptr()), ldout, c_vp(ws.data_ptr()), ws_bytes))

    def get_jk(self, X, W, dm, nset, nao, npts, with_j, with_k, vj, vk, ws=None):
        nbytes = self.lib.isdf_get_jk_ws_bytes(nao, npts)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_get_jk(self.h, c_vp(X.data_ptr()), c_vp(W.data_ptr()), c_vp(dm.data_ptr()),
                                   nset, nao, npts, int(with_j), int(with_k),
                                   c_vp(vj.data_ptr()) if vj is not None else None,
                                   c_vp(vk.data_ptr()) if vk is not None else None,
                                   c_vp(ws.data_ptr()), ws.numel()))

    def phase_expand(self, tab_re, tab_im, ldt, nrow, Gv, ng, R, sign, out_re, out_im, ldo):
        R = np.ascontiguousarray(R, dtype=np.float64)
        check(self.lib.isdf_phase_expand(self.h, c_vp(tab_re.data_ptr()),
                                         c_vp(tab_im.data_ptr()) if tab_im is not None else None, ldt, nrow,
                                         c_vp(Gv.data_ptr()), ng, hptr(R), float(sign),
                                         c_vp(out_re.data_ptr()), c_vp(out_im.data_ptr()), ldo))

    def congruence(self, U, ldu, npts, M, ldm):
        """M <- U^-1 M U^-T in place (symmetric M, upper-triangular U), csrc/fit.cu."""
        nbytes = self.lib.isdf_fit_from_U_ws_bytes(npts, npts)
        ws = self.workspace(nbytes)
        check(self.lib.isdf_congruence(self.h, c_vp(U.data_ptr()), ldu, npts, c_vp(M.data_ptr()), ldm,
                                       c_vp(ws.data_ptr()), nbytes))

    def transpose(self, src, ldin, rows, cols, out, ldout):
        """out[cols, rows] = src[rows, cols]^T on the handle's stream (csrc/fit.cu)."""
        check(self.lib.isdf_transpose(self.h, c_vp(src.data_ptr()), ldin, rows, cols, c_vp(out.data_ptr()), ldout))

    def get_jk_static(self, X, W, dm, nset, nao, npts, with_j, with_k, vj, vk):
        """isdf_get_jk with the library's built-in scratch array: allocates nothing (include/isdf_b200.h)."""
        check(self.lib.isdf_get_jk_static(self.h, c_vp(X.data_ptr()), c_vp(W.data_ptr()), c_vp(dm.data_ptr()),
                                          nset, nao, npts, int(with_j), int(with_k),
                                          c_vp(vj.data_ptr()) if vj is not None else None,
                                          c_vp(vk.data_ptr()) if vk is not None else None))

    def dgemm(self, M, N, K, A, lda, a_kmajor, B, ldb, b_kmajor, C, ldc,
              Cin=None, ldcin=0, E=None, lde=0, alpha=1.0, neg=0, epi=0, upper_only=0):
        check(self.lib.isdf_dgemm(self.h, M, N, K, c_vp(A.data_ptr()), lda, int(a_kmajor),
                                  c_vp(B.data_ptr()), ldb, int(b_kmajor), c_vp(C.data_ptr()), ldc,
                                  c_vp(Cin.data_ptr()) if Cin is not None else None, ldcin,
                                  c_vp(E.data_ptr()) if E is not None else None, lde,
                                  float(alpha), int(neg), int(epi), int(upper_only)))

    def gather_cols(self, src, ldin, nrows, idx, out, ldout):
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        check(self.lib.isdf_gather_cols(self.h, c_vp(src.data_ptr()), ldin, nrows, hptr(idx), len(idx),
                                        c_vp(out.data_ptr()), ldout))


    # -- stepwise stage 2-5 pieces used by the multi-GPU driver (pyscf_isdf_b200/dist.py) ---------
    def select_create(self, nao, ngl, g_offset, ng_total, npts, ws=None):
        """ws: optional caller-kept uint8 device buffer of at least select_ws_bytes(nao, ngl, npts)."""
        nbytes = self.lib.isdf_select_ip_ws_bytes(nao, ngl, npts)
        if ws is None or ws.numel() < nbytes:
            ws = self.workspace(nbytes)
        plan = c_vp()
        check(self.lib.isdf_select_create(self.h, nao, ngl, g_offset, ng_total, npts,
                                          c_vp(ws.data_ptr()), nbytes, ctypes.byref(plan)))
        return SelectPlan(self, plan, ws)

    def select_ws_bytes(self, nao, ngl, npts):
        return int(self.lib.isdf_select_ip_ws_bytes(nao, ngl, npts))

    def syrk_i8_ws_bytes(self, n, K, nslices=8):
        return int(self.lib.isdf_syrk_i8_ws_bytes(n, K, nslices))

    def fit_aux_from_U(self, U, ldu, npts, Z, ldz, ngl):
        nbytes = self.lib.isdf_fit_from_U_ws_bytes(ngl, npts)
        ws = self.workspace(nbytes)
        check(self.lib.isdf_fit_aux_from_U(self.h, c_vp(U.data_ptr()), ldu, npts, c_vp(Z.data_ptr()), ldz, ngl,
                                           c_vp(ws.data_ptr()), nbytes))

    def fft_zy(self, vin, ldin, nvec, nxl, mesh, spec, lds):
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        check(self.lib.isdf_fft_zy(self.h, c_vp(vin.data_ptr()), ldin, nvec, nxl, hptr(mesh),
                                   c_vp(spec.data_ptr()), lds))

    def fft_pack(self, spec, lds, nvec, nxl, ncols, col_off, send):
        col_off = np.ascontiguousarray(col_off, dtype=np.int32)
        check(self.lib.isdf_fft_pack(self.h, c_vp(spec.data_ptr()), lds, nvec, nxl, ncols, hptr(col_off),
                                     len(col_off) - 1, c_vp(send.data_ptr())))

    def fft_unpack(self, recv, nvec, x_off, ncl, Y, ldy):
        x_off = np.ascontiguousarray(x_off, dtype=np.int32)
        check(self.lib.isdf_fft_unpack(self.h, c_vp(recv.data_ptr()), nvec, hptr(x_off), len(x_off) - 1, ncl,
                                       c_vp(Y.data_ptr()), ldy))

    def fft_x_coulomb(self, Y, ldy, nvec, mesh, a, col0, ncl):
        mesh = np.ascontiguousarray(mesh, dtype=np.int32)
        a = np.ascontiguousarray(a, dtype=np.float64)
        check(self.lib.isdf_fft_x_coulomb(self.h, c_vp(Y.data_ptr()), ldy, nvec, hptr(mesh), hptr(a), col0, ncl))

    def symmetrize_upper(self, W, ld, n):
        check(self.lib.isdf_symmetrize_upper(self.h, c_vp(W.data_ptr()), ld, n))

    def syrk_i8(self, n, K, A, lda, C, ldc, alpha=1.0, nslices=8, ws=None):
        """Upper triangle of C = alpha * A A^T on the int8 tensor cores (csrc/syrk_i8.cu).
        ws: optional caller-kept uint8 device buffer of at least syrk_i8_ws_bytes(n, K, nslices)."""
        nbytes = self.lib.isdf_syrk_i8_ws_bytes(n, K, nslices)
        if ws is None or ws.numel() < nbytes:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_syrk_i8(self.h, n, K, c_vp(A.data_ptr()), lda, c_vp(C.data_ptr()), ldc,
                                    float(alpha), int(nslices), c_vp(ws.data_ptr()), nbytes))
        self.syrk_i8_ops = float(self.lib.isdf_syrk_i8_last_ops())
        return ws

    def density_on_grid(self, phi, ldphi, nao, ngl, dm, nset, rho, ldr, ws=None):
        nbytes = self.lib.isdf_grid_j_ws_bytes(nao, ldphi)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_density_on_grid(self.h, c_vp(phi.data_ptr()), ldphi, nao, ngl, c_vp(dm.data_ptr()), nset,
                                            c_vp(rho.data_ptr()), ldr, c_vp(ws.data_ptr()), nbytes))
        return ws

    def j_from_potential(self, phi, ldphi, nao, ngl, vR, ldv, nset, weight, vj, ws=None):
        nbytes = self.lib.isdf_grid_j_ws_bytes(nao, ldphi)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_j_from_potential(self.h, c_vp(phi.data_ptr()), ldphi, nao, ngl, c_vp(vR.data_ptr()), ldv,
                                             nset, float(weight), c_vp(vj.data_ptr()), c_vp(ws.data_ptr()), nbytes))
        return ws

    def get_jk_block(self, X, Wrows, ldw, p0, npl, dm, nset, nao, npts, with_j, with_k, vj, vk, ws=None):
        nbytes = self.lib.isdf_get_jk_block_ws_bytes(nao, npts, npl)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_get_jk_block(self.h, c_vp(X.data_ptr()), c_vp(Wrows.data_ptr()), ldw, p0, npl,
                                         c_vp(dm.data_ptr()), nset, nao, npts, int(with_j), int(with_k),
                                         c_vp(vj.data_ptr()) if vj is not None else None,
                                         c_vp(vk.data_ptr()) if vk is not None else None,
                                         c_vp(ws.data_ptr()), ws.numel()))
        return ws


    def get_jk_block_occ(self, X, Wrows, ldw, p0, npl, Cs, nocc, nao, npts, with_j, with_k, vj, vk, ws=None):
        """J/K of D = C diag(n) C^T from Cs[nocc, nao] = sqrt(n) C^T (include/isdf_b200.h)."""
        nbytes = self.lib.isdf_get_jk_block_ws_bytes(nao, npts, npl)
        if ws is None:
            ws = self.workspace(nbytes)
        check(self.lib.isdf_get_jk_block_occ(self.h, c_vp(X.data_ptr()), c_vp(Wrows.data_ptr()), ldw, p0, npl,
                                             c_vp(Cs.data_ptr()), nocc, nao, npts, int(with_j), int(with_k),
                                             c_vp(vj.data_ptr()) if vj is not None else None,
                                             c_vp(vk.data_ptr()) if vk is not None else None,
                                             c_vp(ws.data_ptr()), ws.numel()))
        return ws


class StageTimer:
    """CUDA-event stopwatch on the handle's stream: mark(name) closes the interval since the previous mark."""

    def __init__(self, handle):
        import torch
        self._torch = torch
        self._dev = handle.torch_device
        self._marks = []
        self._last = torch.cuda.Event(enable_timing=True)
        self._last.record()
        self._times = None

    def mark(self, name):
        e = self._torch.cuda.Event(enable_timing=True)
        e.record()
        self._marks.append((name, self._last, e))
        self._last = e
        self._times = None

    @property
    def times(self):
        if self._times is None:
            self._torch.cuda.synchronize(self._dev)
            self._times = {n: a.elapsed_time(b) for n, a, b in self._marks}
        return self._times


class SelectPlan:
    """One stage-2 run on one slab of grid points (include/isdf_b200.h, isdf_select_*)."""

    def __init__(self, handle, plan, ws):
        self.handle, self.plan, self.ws = handle, plan, ws
        self.lib = handle.lib

    def close(self):
        if self.plan:
            self.lib.isdf_select_destroy(self.plan)
            self.plan = None

    def begin(self, phi, ldphi):
        dmax = c_dbl()
        check(self.lib.isdf_select_begin(self.plan, c_vp(phi.data_ptr()), ldphi, ctypes.byref(dmax)))
        return dmax.value

    def set_tol(self, tol):
        check(self.lib.isdf_select_set_tol(self.plan, float(tol)))

    def candidates(self, phi, ldphi, Lt, ldl, k0, quota, block):
        check(self.lib.isdf_select_candidates(self.plan, c_vp(phi.data_ptr()), ldphi, c_vp(Lt.data_ptr()), ldl,
                                              k0, quota, c_vp(block.data_ptr())))

    def panel(self, blocks, nranks, quota, k0, nbmax, U, ldu):
        nb, stop = c_int(), c_int()
        check(self.lib.isdf_select_panel(self.plan, c_vp(blocks.data_ptr()), nranks, quota, k0, nbmax,
                                         c_vp(U.data_ptr()) if U is not None else None, ldu,
                                         ctypes.byref(nb), ctypes.byref(stop)))
        return nb.value, bool(stop.value)

    def flush(self, phi, ldphi, Lt, ldl, k0, nb):
        check(self.lib.isdf_select_flush(self.plan, c_vp(phi.data_ptr()), ldphi, c_vp(Lt.data_ptr()), ldl, k0, nb))

    def finish(self, rank):
        ip = np.zeros(max(rank, 1), dtype=np.int64)
        resid = np.zeros(max(rank, 1), dtype=np.float64)
        check(self.lib.isdf_select_finish(self.plan, rank, hptr(ip), hptr(resid)))
        return ip[:rank].copy(), resid[:rank].copy()
