// CPU harness for pyscf_isdf_b200/csrc/ao_core.cuh (solid harmonics with gradients, radial part of the
// AO gradient kernel).  TEST INFRASTRUCTURE ONLY; used by tests/test_ao_emul_cpu.py.
#define __host__
#define __device__
#include <cmath>
using std::exp; using std::sqrt;
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

// Whole-grid driver around isdf_ao::shell_point_deriv1 with the kernel's block structure: per block of 56
// grid points and per atom, the images whose block-minimum distance is below the atom's largest shell
// cutoff are listed with that distance (what eval_ao_deriv1_kernel holds in shared memory), then every
// shell / contraction / point is accumulated by the shared routine.  out[4][nao][ng].
#include <vector>
#include <algorithm>
extern "C" void ao_grid_deriv1(const int* atm, int natm, const int* bas, int nbas, const double* env,
                               const double* rcut, const double* Ls, int nimg,
                               long long ng, const double* coords, double* out)
{
    const int BLK = 56;
    std::vector<int> ao_off(nbas + 1, 0);
    for (int ib = 0; ib < nbas; ++ib) ao_off[ib + 1] = ao_off[ib] + bas[ib * 8 + 3] * (2 * bas[ib * 8 + 1] + 1);
    const long long nao = ao_off[nbas];
    for (long long b0 = 0; b0 < ng; b0 += BLK) {
        const int npt = (int)std::min<long long>(BLK, ng - b0);
        for (int ia = 0; ia < natm; ++ia) {
            const double* r = env + atm[ia * 6 + 1];
            double rcmax = 0.0;
            for (int ib = 0; ib < nbas; ++ib) if (bas[ib * 8] == ia) rcmax = std::max(rcmax, rcut[ib]);
            std::vector<int> img; std::vector<double> mind;
            for (int m = 0; m < nimg; ++m) {
                double d2min = 1e300;
                for (int i = 0; i < npt; ++i) {
                    const double dx = coords[3 * (b0 + i)] - (r[0] + Ls[3 * m]);
                    const double dy = coords[3 * (b0 + i) + 1] - (r[1] + Ls[3 * m + 1]);
                    const double dz = coords[3 * (b0 + i) + 2] - (r[2] + Ls[3 * m + 2]);
                    d2min = std::min(d2min, dx * dx + dy * dy + dz * dz);
                }
                if (std::sqrt(d2min) < rcmax) { img.push_back(m); mind.push_back(std::sqrt(d2min)); }
            }
            for (int ib = 0; ib < nbas; ++ib) {
                if (bas[ib * 8] != ia) continue;
                const int l = bas[ib * 8 + 1], nprim = bas[ib * 8 + 2], nctr = bas[ib * 8 + 3];
                const double fac = l == 0 ? 0.282094791773878143 : (l == 1 ? 0.488602511902919921 : 1.0);
                const int deg = 2 * l + 1;
                for (int c = 0; c < nctr; ++c)
                    for (int i = 0; i < npt; ++i) {
                        double acc[4][7];
                        isdf_ao::shell_point_deriv1(l, nprim, env + bas[ib * 8 + 5], env + bas[ib * 8 + 6] + c * nprim, fac,
                                                    rcut[ib], img.data(), mind.data(), (int)img.size(), Ls,
                                                    r[0], r[1], r[2], coords[3 * (b0 + i)], coords[3 * (b0 + i) + 1],
                                                    coords[3 * (b0 + i) + 2], acc);
                        for (int k = 0; k < 4; ++k)
                            for (int d = 0; d < deg; ++d)
                                out[(k * nao + ao_off[ib] + c * deg + d) * ng + b0 + i] = acc[k][d];
                    }
            }
        }
    }
}
