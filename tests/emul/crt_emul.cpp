// CPU harness for pyscf_isdf_b200/csrc/crt_core.cuh (residues / Garner reconstruction of the CRT form of
// the int8 metric product).  TEST INFRASTRUCTURE ONLY; used by tests/test_crt_emul_cpu.py.
#define __host__
#define __device__
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
