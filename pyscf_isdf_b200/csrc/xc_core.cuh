// Closed-shell exchange-correlation functionals at one grid point, f(rho, sigma) per unit volume with
// its two first derivatives carried along by forward-mode dual numbers (value, d/d rho, d/d sigma), so
// that each functional is written once, as its published energy expression, and no derivative is
// derived by hand.  What the reference gets from libxc through pyscf/dft/libxc.py:eval_xc (spin = 0,
// deriv = 1): exc per particle, vrho, vsigma.
//   Slater exchange; VWN correlation, form V (Can. J. Phys. 58, 1200); Becke 88 exchange (Phys. Rev. A 38, 3098); LYP correlation in the closed-shell
//   form of Miehlich et al. (Chem. Phys. Lett. 157, 200); PBE exchange and correlation (Phys. Rev.
//   Lett. 77, 3865) over PW92 with libxc's pw_mod constants.
// Like fft_core.cuh and crt_core.cuh this header also compiles with g++ (tests/emul/xc_emul.cpp), where
// it is checked point by point against oracle/xc.py (complex-step derivatives of the same formulas,
// pinned to the reference's BLYP numbers).
#pragma once
#include <cmath>

namespace isdf_xc {

constexpr int XC_LDA_X = 1;      // 'lda,'
constexpr int XC_BLYP = 2;       // 'blyp' = b88 + lyp
constexpr int XC_PBE = 3;        // 'pbe'  = pbe exchange + pbe correlation
constexpr int XC_PBE0 = 5;       // semilocal part of 'pbe0':  0.75 pbe exchange + pbe correlation (exact-exchange fraction 0.25)
constexpr int XC_B3LYP = 6;      // semilocal part of libxc's 'b3lyp': 0.08 slater + 0.72 b88 + 0.81 lyp + 0.19 vwn-rpa (fraction 0.20)
constexpr int XC_LDA_VWN = 4;    // 'lda,vwn' = slater + vwn (form V): the default xc of the reference's pbc RKS
constexpr double RHO_CUT = 1e-15;
constexpr double SIGMA_FLOOR = 1e-40;
constexpr double PI = 3.14159265358979323846;

struct Dual { double v, r, s; };

__host__ __device__ __forceinline__ Dual cst(double a) { return Dual{a, 0.0, 0.0}; }
__host__ __device__ __forceinline__ Dual operator+(Dual a, Dual b) { return Dual{a.v + b.v, a.r + b.r, a.s + b.s}; }
__host__ __device__ __forceinline__ Dual operator-(Dual a, Dual b) { return Dual{a.v - b.v, a.r - b.r, a.s - b.s}; }
__host__ __device__ __forceinline__ Dual operator-(Dual a) { return Dual{-a.v, -a.r, -a.s}; }
__host__ __device__ __forceinline__ Dual operator*(Dual a, Dual b)
{
    return Dual{a.v * b.v, a.r * b.v + a.v * b.r, a.s * b.v + a.v * b.s};
}
__host__ __device__ __forceinline__ Dual operator/(Dual a, Dual b)
{
    const double q = a.v / b.v, ib = 1.0 / b.v;
    return Dual{q, (a.r - q * b.r) * ib, (a.s - q * b.s) * ib};
}
__host__ __device__ __forceinline__ Dual operator+(Dual a, double b) { return Dual{a.v + b, a.r, a.s}; }
__host__ __device__ __forceinline__ Dual operator+(double b, Dual a) { return Dual{a.v + b, a.r, a.s}; }
__host__ __device__ __forceinline__ Dual operator-(Dual a, double b) { return Dual{a.v - b, a.r, a.s}; }
__host__ __device__ __forceinline__ Dual operator-(double b, Dual a) { return Dual{b - a.v, -a.r, -a.s}; }
__host__ __device__ __forceinline__ Dual operator*(Dual a, double b) { return Dual{a.v * b, a.r * b, a.s * b}; }
__host__ __device__ __forceinline__ Dual operator*(double b, Dual a) { return Dual{a.v * b, a.r * b, a.s * b}; }
__host__ __device__ __forceinline__ Dual operator/(Dual a, double b) { return a * (1.0 / b); }
__host__ __device__ __forceinline__ Dual operator/(double b, Dual a) { return cst(b) / a; }
__host__ __device__ __forceinline__ Dual chain(Dual a, double fv, double dfv) { return Dual{fv, dfv * a.r, dfv * a.s}; }
__host__ __device__ __forceinline__ Dual dpow(Dual a, double p) { const double v = pow(a.v, p); return chain(a, v, p * v / a.v); }
__host__ __device__ __forceinline__ Dual dsqrt(Dual a) { const double v = sqrt(a.v); return chain(a, v, 0.5 / v); }
__host__ __device__ __forceinline__ Dual dexp(Dual a) { const double v = exp(a.v); return chain(a, v, v); }
__host__ __device__ __forceinline__ Dual dlog(Dual a) { return chain(a, log(a.v), 1.0 / a.v); }
__host__ __device__ __forceinline__ Dual dasinh(Dual a) { return chain(a, asinh(a.v), 1.0 / sqrt(1.0 + a.v * a.v)); }

__host__ __device__ __forceinline__ Dual datan(Dual a) { return chain(a, atan(a.v), 1.0 / (1.0 + a.v * a.v)); }

__host__ __device__ inline Dual f_vwn(Dual rho, double A, double b, double c, double x0)
{
    const Dual x = dsqrt(dpow(3.0 / (4.0 * PI) / rho, 1.0 / 3.0));
    const Dual X = x * x + b * x + c;
    const double X0 = x0 * x0 + b * x0 + c;
    const double Q = sqrt(4.0 * c - b * b);
    const Dual at = datan(Q / (2.0 * x + b));
    const Dual xm = x - x0;
    const Dual ec = A * (dlog(x * x / X) + 2.0 * b / Q * at
                         - b * x0 / X0 * (dlog(xm * xm / X) + 2.0 * (b + 2.0 * x0) / Q * at));
    return rho * ec;
}
__host__ __device__ inline Dual f_vwn5(Dual rho) { return f_vwn(rho, 0.0310907, 3.72744, 12.9352, -0.10498); }
__host__ __device__ inline Dual f_vwn_rpa(Dual rho) { return f_vwn(rho, 0.0310907, 13.0720, 42.7198, -0.409286); }

__host__ __device__ inline Dual f_slater(Dual rho)
{
    return -0.75 * pow(3.0 / PI, 1.0 / 3.0) * dpow(rho, 4.0 / 3.0);
}

__host__ __device__ inline Dual f_b88(Dual rho, Dual sigma)
{
    const double beta = 0.0042;
    const Dual ra = 0.5 * rho;                         // one spin channel
    const Dual ga = 0.5 * dsqrt(sigma);
    const Dual r43 = dpow(ra, 4.0 / 3.0);
    const Dual x = ga / r43;
    const double cx = 1.5 * pow(3.0 / (4.0 * PI), 1.0 / 3.0);
    return 2.0 * (-cx * r43 - beta * r43 * x * x / (1.0 + 6.0 * beta * x * dasinh(x)));
}

__host__ __device__ inline Dual f_lyp(Dual rho, Dual sigma)
{
    const double a = 0.04918, b = 0.132, c = 0.2533, d = 0.349;
    const double cf = 0.3 * pow(3.0 * PI * PI, 2.0 / 3.0);
    const Dual r13 = dpow(rho, -1.0 / 3.0);
    const Dual den = 1.0 + d * r13;
    const Dual omega = dexp(-c * r13) / den * dpow(rho, -11.0 / 3.0);
    const Dual delta = c * r13 + d * r13 / den;
    const Dual ra = 0.5 * rho;
    const Dual saa = 0.25 * sigma;                     // |grad rho_a|^2 = |grad rho_b|^2
    const Dual t1 = -a * 4.0 / den * ra * ra / rho;
    const Dual brace = ra * ra * (pow(2.0, 11.0 / 3.0) * cf * 2.0 * dpow(ra, 8.0 / 3.0)
                                  + (47.0 / 18.0 - 7.0 / 18.0 * delta) * sigma
                                  - (2.5 - delta / 18.0) * 2.0 * saa
                                  - (delta - 11.0) / 9.0 * 2.0 * (ra / rho) * saa)
                       - 2.0 / 3.0 * rho * rho * sigma + 2.0 * (2.0 / 3.0 * rho * rho - ra * ra) * saa;
    return t1 - a * b * omega * brace;
}

__host__ __device__ inline Dual f_pbe_x(Dual rho, Dual sigma)
{
    const double kappa = 0.804, mu = 0.2195149727645171;
    const Dual kf = dpow(3.0 * PI * PI * rho, 1.0 / 3.0);
    const Dual d = 2.0 * kf * rho;
    const Dual s2 = sigma / (d * d);
    return f_slater(rho) * (1.0 + kappa - kappa / (1.0 + mu / kappa * s2));
}

__host__ __device__ inline Dual eps_pw92(Dual rs)
{
    const double A = 0.0310906908696549, a1 = 0.21370, b1 = 7.5957, b2 = 3.5876, b3 = 1.6382, b4 = 0.49294;
    const Dual sq = dsqrt(rs);
    const Dual q = 2.0 * A * (b1 * sq + b2 * rs + b3 * rs * sq + b4 * rs * rs);
    return -2.0 * A * (1.0 + a1 * rs) * dlog(1.0 + 1.0 / q);
}

__host__ __device__ inline Dual f_pbe_c(Dual rho, Dual sigma)
{
    const double beta = 0.06672455060314922;
    const double gamma = (1.0 - 0.69314718055994530942) / (PI * PI);
    const Dual rs = dpow(3.0 / (4.0 * PI) / rho, 1.0 / 3.0);
    const Dual ec = eps_pw92(rs);
    const Dual kf = dpow(3.0 * PI * PI * rho, 1.0 / 3.0);
    const Dual ks2 = 4.0 / PI * kf;
    const Dual t2 = sigma / (4.0 * ks2 * rho * rho);
    const Dual A = beta / gamma / (dexp(-ec / gamma) - 1.0);
    const Dual At2 = A * t2;
    return rho * (ec + gamma * dlog(1.0 + beta / gamma * t2 * (1.0 + At2) / (1.0 + At2 + At2 * At2)));
}

// exact-exchange fraction that goes with a functional code (the caller adds -fraction/2 K)
__host__ __device__ inline double hybrid_fraction(int code) { return code == XC_PBE0 ? 0.25 : (code == XC_B3LYP ? 0.20 : 0.0); }

// exc per particle, vrho = df/drho, vsigma = df/dsigma; zeros below the density threshold.
// Returns false for an unknown functional code.
__host__ __device__ inline bool eval_point(int code, double rho, double sigma, double* exc, double* vrho, double* vsigma)
{
    *exc = 0.0; *vrho = 0.0; *vsigma = 0.0;
    if (code < XC_LDA_X || code > XC_B3LYP) return false;
    if (!(rho > RHO_CUT)) return true;
    const Dual r{rho, 1.0, 0.0};
    const Dual s{sigma > SIGMA_FLOOR ? sigma : SIGMA_FLOOR, 0.0, 1.0};
    Dual f;
    if (code == XC_LDA_X) f = f_slater(r);
    else if (code == XC_LDA_VWN) f = f_slater(r) + f_vwn5(r);
    else if (code == XC_BLYP) f = f_b88(r, s) + f_lyp(r, s);
    else if (code == XC_PBE0) f = 0.75 * f_pbe_x(r, s) + f_pbe_c(r, s);
    else if (code == XC_B3LYP) f = 0.08 * f_slater(r) + 0.72 * f_b88(r, s) + 0.81 * f_lyp(r, s) + 0.19 * f_vwn_rpa(r);
    else f = f_pbe_x(r, s) + f_pbe_c(r, s);
    *exc = f.v / rho; *vrho = f.r; *vsigma = f.s;
    return true;
}

// ---- per-grid-point bodies of the streaming kernels of xc.cu (shared with tests/emul/xc_emul.cpp) ----

// rho4[c][g] = f_c sum_mu phi_c[mu,g] T[mu,g],  f = 1, 2, 2, 2   (eval_rho, pyscf/pbc/dft/numint.py:138-145)
__host__ __device__ inline void rho_item(long long g, const double* phi4, long long ldphi, long long cstride,
                                         const double* T, int nao, double* rho4, long long ldr)
{
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (int mu = 0; mu < nao; ++mu) {
        const long long o = (long long)mu * ldphi + g;
        const double t = T[o];
        s0 = fma(phi4[o], t, s0);
        s1 = fma(phi4[o + cstride], t, s1);
        s2 = fma(phi4[o + 2 * cstride], t, s2);
        s3 = fma(phi4[o + 3 * cstride], t, s3);
    }
    rho4[g] = s0;
    rho4[ldr + g] = 2.0 * s1;
    rho4[2 * ldr + g] = 2.0 * s2;
    rho4[3 * ldr + g] = 2.0 * s3;
}

// functional at one point: wv4 = weight (vrho/2, 2 vsigma grad rho) (numint.py:366, :987-994);
// den = rho weight, ex = den exc (numint.py:361-365)
__host__ __device__ inline void point_item(int code, long long g, const double* rho4, long long ldr, double weight,
                                           double* wv4, double* den, double* ex)
{
    const double rho = rho4[g], gx = rho4[ldr + g], gy = rho4[2 * ldr + g], gz = rho4[3 * ldr + g];
    double exc, vrho, vsigma;
    eval_point(code, rho, gx * gx + gy * gy + gz * gz, &exc, &vrho, &vsigma);
    *den = rho * weight;
    *ex = *den * exc;
    const double f = 2.0 * weight * vsigma;
    wv4[g] = 0.5 * weight * vrho;
    wv4[ldr + g] = f * gx;
    wv4[2 * ldr + g] = f * gy;
    wv4[3 * ldr + g] = f * gz;
}

// aow[mu,g] = sum_c wv4[c][g] phi_c[mu,g] for mu = mu0, mu0 + step, ...   (_scale_ao, numint.py:812-814)
__host__ __device__ inline void aow_item(long long g, int mu0, int step, const double* phi4, long long ldphi,
                                         long long cstride, const double* wv4, long long ldr, int nao, double* aow)
{
    const double w0 = wv4[g], w1 = wv4[ldr + g], w2 = wv4[2 * ldr + g], w3 = wv4[3 * ldr + g];
    for (int mu = mu0; mu < nao; mu += step) {
        const long long o = (long long)mu * ldphi + g;
        aow[o] = fma(w3, phi4[o + 3 * cstride], fma(w2, phi4[o + 2 * cstride], fma(w1, phi4[o + cstride], w0 * phi4[o])));
    }
}

}  // namespace isdf_xc
