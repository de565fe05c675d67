// Real solid harmonics r^l Y_lm (l <= 3) of the AO evaluator with their Cartesian gradients, in the
// normalisation and order of eval_ao.cu (libcint's CINTc2s_ket_sph1 after CINTcommon_fac_sp; l = 1 as
// x, y, z).  Shared by eval_ao_deriv1.cu and, compiled with g++, by tests/emul/ao_emul.cpp, where the
// gradients are checked against the oracle's cartesian -> spherical route.
#pragma once

namespace isdf_ao {

// ang[d], dang[axis][d] for d < 2l+1; the l = 0, 1 prefactors 1/(2 sqrt(pi)), sqrt(3/(4 pi)) are NOT
// included (the caller folds them into the radial part, as the reference does).
__host__ __device__ inline void ang_grad(int l, double x, double y, double z, double* ang, double (*dang)[7])
{
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int d = 0; d < 7; ++d) dang[a][d] = 0.0;
    if (l == 0) {
        ang[0] = 1.0;
    } else if (l == 1) {
        ang[0] = x; ang[1] = y; ang[2] = z;
        dang[0][0] = 1.0; dang[1][1] = 1.0; dang[2][2] = 1.0;
    } else if (l == 2) {
        const double c1 = 1.092548430592079070, c0 = 0.315391565252520002, c2 = 0.546274215296039535;
        ang[0] = c1 * (x * y);  dang[0][0] = c1 * y;  dang[1][0] = c1 * x;
        ang[1] = c1 * (y * z);  dang[1][1] = c1 * z;  dang[2][1] = c1 * y;
        ang[2] = 2.0 * c0 * z * z - c0 * (x * x + y * y);
        dang[0][2] = -2.0 * c0 * x;  dang[1][2] = -2.0 * c0 * y;  dang[2][2] = 4.0 * c0 * z;
        ang[3] = c1 * (x * z);  dang[0][3] = c1 * z;  dang[2][3] = c1 * x;
        ang[4] = c2 * (x * x - y * y);  dang[0][4] = 2.0 * c2 * x;  dang[1][4] = -2.0 * c2 * y;
    } else {
        const double a = 1.770130769779930531, b = 0.590043589926643510, c = 2.890611442640554055;
        const double e = 1.828183197857862944, f = 0.457045799464465739;
        const double p = 0.746352665180230782, q = 1.119528997770346170, g = 1.445305721320277020;
        const double xx = x * x, yy = y * y, zz = z * z;
        ang[0] = y * (a * xx - b * yy);
        dang[0][0] = 2.0 * a * x * y;  dang[1][0] = a * xx - 3.0 * b * yy;
        ang[1] = c * (x * y * z);
        dang[0][1] = c * y * z;  dang[1][1] = c * x * z;  dang[2][1] = c * x * y;
        ang[2] = y * (e * zz - f * (xx + yy));
        dang[0][2] = -2.0 * f * x * y;  dang[1][2] = e * zz - f * xx - 3.0 * f * yy;  dang[2][2] = 2.0 * e * y * z;
        ang[3] = z * (p * zz - q * (xx + yy));
        dang[0][3] = -2.0 * q * x * z;  dang[1][3] = -2.0 * q * y * z;  dang[2][3] = 3.0 * p * zz - q * (xx + yy);
        ang[4] = x * (e * zz - f * (xx + yy));
        dang[0][4] = e * zz - 3.0 * f * xx - f * yy;  dang[1][4] = -2.0 * f * x * y;  dang[2][4] = 2.0 * e * x * z;
        ang[5] = g * z * (xx - yy);
        dang[0][5] = 2.0 * g * x * z;  dang[1][5] = -2.0 * g * y * z;  dang[2][5] = g * (xx - yy);
        ang[6] = x * (b * xx - a * yy);
        dang[0][6] = 3.0 * b * xx - a * yy;  dang[1][6] = -2.0 * a * x * y;
    }
}

// One contracted radial function and its derivative factor at r^2 = rr:
//   e0 = fac sum_j c_j exp(-a_j rr),   e1 = fac sum_j c_j (-2 a_j) exp(-a_j rr)
// so that  d/dx [S(r) e0(r^2)] = dS/dx e0 + S x e1   (GTOshell_eval_grid_cart_deriv1, pyscf/lib/gto/deriv1.c:167-204)
__host__ __device__ inline void radial01(const double* ex, const double* cf, int nprim, double rr, double fac,
                                         double* e0, double* e1)
{
    double s0 = 0.0, s1 = 0.0;
    for (int j = 0; j < nprim; ++j) {
        const double e = exp(-ex[j] * rr) * fac * cf[j];
        s0 += e;
        s1 += e * (-2.0 * ex[j]);
    }
    *e0 = s0; *e1 = s1;
}

// Running sums of one contraction of one shell at one grid point over the screened image list of the
// point's block: acc[k][d], k = value, d/dx, d/dy, d/dz.  img/mind: images kept for the atom and their
// block-minimum distances; an image contributes to this shell iff mind < rcut (the reference's rule,
// pyscf/lib/pbc/grid_ao.c:344-365).  ex/cf: exponents and the coefficients of THIS contraction.
__host__ __device__ inline void shell_point_deriv1(int l, int nprim, const double* ex, const double* cf, double fac,
                                                   double rcut, const int* img, const double* mind, int cnt,
                                                   const double* Ls, double ax, double ay, double az,
                                                   double px, double py, double pz, double (*acc)[7])
{
    const int deg = 2 * l + 1;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int d = 0; d < 7; ++d) acc[k][d] = 0.0;
    for (int q = 0; q < cnt; ++q) {
        if (!(mind[q] < rcut)) continue;
        const int m = img[q];
        const double dx = px - (ax + Ls[3 * m]);
        const double dy = py - (ay + Ls[3 * m + 1]);
        const double dz = pz - (az + Ls[3 * m + 2]);
        double e0, e1, ang[7], dang[3][7];
        radial01(ex, cf, nprim, dx * dx + dy * dy + dz * dz, fac, &e0, &e1);
        ang_grad(l, dx, dy, dz, ang, dang);
#pragma unroll
        for (int d = 0; d < 7; ++d) {
            if (d < deg) {
                const double r1 = e1 * ang[d];
                acc[0][d] += e0 * ang[d];
                acc[1][d] += e0 * dang[0][d] + r1 * dx;
                acc[2][d] += e0 * dang[1][d] + r1 * dy;
                acc[3][d] += e0 * dang[2][d] + r1 * dz;
            }
        }
    }
}

}  // namespace isdf_ao
