// CPU harness for pyscf_isdf_b200/csrc/xc_core.cuh (dual-number exchange-correlation functionals).
// TEST INFRASTRUCTURE ONLY; used by tests/test_xc_emul_cpu.py.
#define __host__
#define __device__
#define __forceinline__ inline
#include <cmath>
using std::pow; using std::sqrt; using std::exp; using std::log; using std::asinh;
#include "../../pyscf_isdf_b200/csrc/xc_core.cuh"

extern "C" int xc_eval(int code, long long n, const double* rho, const double* sigma,
                       double* exc, double* vrho, double* vsigma)
{
    for (long long i = 0; i < n; ++i)
        if (!isdf_xc::eval_point(code, rho[i], sigma[i], exc + i, vrho + i, vsigma + i)) return 1;
    return 0;
}
