// CPU harness for pyscf_isdf_b200/csrc/xc_core.cuh (dual-number exchange-correlation functionals).
// TEST INFRASTRUCTURE ONLY; used by tests/test_xc_emul_cpu.py.
#define __host__
#define __device__
#define __forceinline__ inline
#include <cmath>
using std::pow; using std::sqrt; using std::exp; using std::log; using std::asinh; using std::fma; using std::atan;
#include "../../pyscf_isdf_b200/csrc/xc_core.cuh"

extern "C" int xc_eval(int code, long long n, const double* rho, const double* sigma,
                       double* exc, double* vrho, double* vsigma)
{
    for (long long i = 0; i < n; ++i)
        if (!isdf_xc::eval_point(code, rho[i], sigma[i], exc + i, vrho + i, vsigma + i)) return 1;
    return 0;
}

// the three streaming kernels of xc.cu, item by item (grid rounded up past ngl like a launch; the aow
// kernel's grid.y stride emulated with ystep "blocks")
extern "C" void xc_stream_items(int code, int nao, long long ngl, long long ld, const double* phi4, const double* T,
                                double weight, double* rho4, double* wv4, double* aow, double* sums, int ystep)
{
    const long long cstride = (long long)nao * ld;
    double n = 0.0, e = 0.0;
    for (long long g = 0; g < ngl; ++g) isdf_xc::rho_item(g, phi4, ld, cstride, T, nao, rho4, ld);
    for (long long g = 0; g < ngl; ++g) {
        double den, ex;
        isdf_xc::point_item(code, g, rho4, ld, weight, wv4, &den, &ex);
        n += den; e += ex;
    }
    for (int y = 0; y < ystep; ++y)
        for (long long g = 0; g < ngl; ++g) isdf_xc::aow_item(g, y, ystep, phi4, ld, cstride, wv4, ld, nao, aow);
    sums[0] = n; sums[1] = e;
}
