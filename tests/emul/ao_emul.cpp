// CPU harness for pyscf_isdf_b200/csrc/ao_core.cuh (solid harmonics with gradients, radial part of the
// AO gradient kernel).  TEST INFRASTRUCTURE ONLY; used by tests/test_ao_emul_cpu.py.
#define __host__
#define __device__
#include <cmath>
using std::exp;
#include "../../pyscf_isdf_b200/csrc/ao_core.cuh"

// values and gradients of one contracted shell (one contraction) at n displaced points: out[4][2l+1][n]
extern "C" void ao_shell_deriv1(int l, int nprim, const double* ex, const double* cf, double fac,
                                long long n, const double* dr, double* out)
{
    const int deg = 2 * l + 1;
    for (long long i = 0; i < n; ++i) {
        const double x = dr[3 * i], y = dr[3 * i + 1], z = dr[3 * i + 2];
        double ang[7], dang[3][7], e0, e1;
        isdf_ao::ang_grad(l, x, y, z, ang, dang);
        isdf_ao::radial01(ex, cf, nprim, x * x + y * y + z * z, fac, &e0, &e1);
        const double d[3] = {x, y, z};
        for (int m = 0; m < deg; ++m) {
            out[(0 * deg + m) * n + i] = e0 * ang[m];
            for (int a = 0; a < 3; ++a) out[((1 + a) * deg + m) * n + i] = e0 * dang[a][m] + e1 * d[a] * ang[m];
        }
    }
}
