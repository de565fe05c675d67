// CPU harness for pyscf_isdf_b200/csrc/crt_core.cuh (residues / Garner reconstruction of the CRT form of
// the int8 metric product).  TEST INFRASTRUCTURE ONLY; used by tests/test_crt_emul_cpu.py.
#define __host__
#define __device__
#include <cstdint>
#include <cmath>
using std::ldexp;
using std::llrint;
using std::rint;
using std::fma;
#include "../../pyscf_isdf_b200/csrc/crt_core.cuh"

extern "C" int crt_modulus(int k) { return isdf_crt::modulus(k); }
extern "C" int crt_input_bits(int nm) { return isdf_crt::input_bits(nm); }
extern "C" void crt_residues(long long x, signed char* out, int nm) { isdf_crt::residues(x, (int8_t*)out, nm); }
extern "C" int crt_centred_mod_rt(int v, int p) { return isdf_crt::centred_mod_rt(v, p, 1.0 / p); }
extern "C" double crt_reconstruct(const int* r, int nm)
{
    int rr[isdf_crt::NM_MAX] = {0};
    for (int k = 0; k < nm; ++k) rr[k] = r[k];
    return isdf_crt::reconstruct(rr, nm);
}

// the two element-wise kernels, item by item (what i8_split_crt_kernel / crt_reconstruct_kernel launch)
extern "C" void crt_split_all(const double* A, long long lda, int n, int nrows_pad, long long k0, int kc, int kcpad,
                              const double* sc, signed char* planes, long long rb_stride, long long plane_stride,
                              int nm, int bits)
{
    const long long items = (long long)nrows_pad * (kcpad / 16);
    for (long long idx = 0; idx < items + 7; ++idx)       // a few items past the end, like a rounded-up grid
        isdf_crt::split_item(idx, A, lda, n, nrows_pad, k0, kc, kcpad, sc, (int8_t*)planes, rb_stride, plane_stride, nm, bits);
}

extern "C" void crt_reconstruct_all(const signed char* R, long long ldr, long long res_stride, int nm, int n,
                                    const double* sc, double f, double* C, long long ldc)
{
    for (int row = 0; row < n; ++row)
        for (int col = 0; col < ((n + 127) / 128) * 128; ++col)
            isdf_crt::reconstruct_item(row, col, (const int8_t*)R, ldr, res_stride, nm, n, sc, f, C, ldc);
}
