// Line FFTs in shared memory: the device routines every FFT kernel of fft.cu calls, and the host
// code that builds their tables.  Kept in a header so that the same source compiles twice:
// by nvcc into fft.cu, and by g++ into the CPU emulation harness tests/emul/fft_emul.cpp (a CTA
// is emulated by real threads and a barrier), which checks the index logic against numpy.fft
// without a GPU.  Nothing here launches a kernel or touches the CUDA runtime.
#pragma once
#include <vector>
#include <cmath>
#include <cstring>

// plain types shared by fft.cu and fft_fused.cu (kernel parameters, launcher signatures)
namespace isdf_fft {

constexpr int FFT_THREADS = 256;
constexpr int MAX_RAD = 16;
constexpr int BIG_RADIX_MAX = 256;

struct Radix {
    int n, nrad;
    int rad[MAX_RAD];
    const double2* tw;      // exp(-2 pi i k / n), k < n
};

// A prime length n > 7 is transformed with Rader's algorithm: the n-1 non-zero
// indices, permuted by powers of a generator of (Z/n)*, turn the DFT into a
// cyclic convolution of length n-1 (composite), done with two sub-FFTs.
struct FftPlan {
    Radix main;
    int rader;              // 1: main.n is prime, use sub / perm / bhat;  2: the same through rader_fused_smem (sub
                            //    factorised by factorize_wide);  3: composite n factorised by factorize_wide.  Plans
                            //    with rader >= 2 may only run on the FUSED kernels (fft_fused.cu)
    Radix sub;              // length n-1
    const int* perm;        // g^p mod n,      p < n-1
    const int* iperm;       // g^-q mod n,     q < n-1
    const double2* bhat_f;  // FFT_{n-1}(w^(g^-m)) / (n-1) for the forward transform (w = e^{-2 pi i/n})
    const double2* bhat_b;  // same for the inverse transform (w = e^{+2 pi i/n})
    __host__ __device__ int n() const { return main.n; }
    __host__ __device__ int lalloc() const { return main.n + (rader ? 2 : 0); }   // smem slots per line
};

struct AxisOffsets { size_t tw, sub, bf, bb, perm, iperm; };   // offsets of one axis' tables inside the shared arrays

}  // namespace isdf_fft
using namespace isdf_fft;

namespace {            // internal linkage, as when this code lived inside fft.cu (keeps nvcc's inlining)

__device__ __forceinline__ double2 cmul(double2 a, double2 b)
{
    return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }

template <int SIGN>
__device__ __forceinline__ double2 twid(const double2* tw, int idx)
{
    double2 w = tw[idx];
    if (SIGN > 0) w.y = -w.y;
    return w;
}

template <int R1, int R2, int SIGN>
__device__ __forceinline__ void dft_pair(double2* u, const double2* tw, int n);

// In-place small DFT of u[0..R), sign SIGN: hand-written butterflies for R = 2, 3, 4, 5, 7, 10, two-level
// butterflies (dft_pair) for R = 6, 8, 9, 12 (only the FUSED kernels use those), a direct sum otherwise
template <int R, int SIGN>
__device__ __forceinline__ void dft_small(double2* u, const double2* tw, int n)
{
    if (R == 2) {
        const double2 a = u[0], b = u[1];
        u[0] = cadd(a, b); u[1] = csub(a, b);
    } else if (R == 4) {
        const double2 s02 = cadd(u[0], u[2]), d02 = csub(u[0], u[2]);
        const double2 s13 = cadd(u[1], u[3]), d13 = csub(u[1], u[3]);
        // W = SIGN * i ;  W*d13 = SIGN * (-d13.y, d13.x)
        const double2 wd = make_double2(-SIGN * d13.y, SIGN * d13.x);
        u[0] = cadd(s02, s13); u[2] = csub(s02, s13);
        u[1] = cadd(d02, wd);  u[3] = csub(d02, wd);
    } else if (R == 3) {
        const double2 t = cadd(u[1], u[2]);
        const double2 m = make_double2(u[0].x - 0.5 * t.x, u[0].y - 0.5 * t.y);
        const double c = 0.86602540378443864676 * SIGN;
        const double2 d = csub(u[1], u[2]);
        const double2 s = make_double2(-c * d.y, c * d.x);       // SIGN * i * (sqrt3/2) * d
        u[0] = cadd(u[0], t); u[1] = cadd(m, s); u[2] = csub(m, s);
    } else if (R == 5) {
        // X1,4 = m1 +- i SIGN n1, X2,3 = m2 +- i SIGN n2 (constants cos/sin(2 pi k/5))
        const double c1 = 0.30901699437494742, c2 = -0.80901699437494742;
        const double s1 = 0.95105651629515357 * SIGN, s2 = 0.58778525229247313 * SIGN;
        const double2 t1 = cadd(u[1], u[4]), t2 = cadd(u[2], u[3]);
        const double2 t3 = csub(u[1], u[4]), t4 = csub(u[2], u[3]);
        const double2 m1 = make_double2(u[0].x + c1 * t1.x + c2 * t2.x, u[0].y + c1 * t1.y + c2 * t2.y);
        const double2 m2 = make_double2(u[0].x + c2 * t1.x + c1 * t2.x, u[0].y + c2 * t1.y + c1 * t2.y);
        // i * (s1 t3 + s2 t4), i * (s2 t3 - s1 t4)
        const double2 n1 = make_double2(-(s1 * t3.y + s2 * t4.y), s1 * t3.x + s2 * t4.x);
        const double2 n2 = make_double2(-(s2 * t3.y - s1 * t4.y), s2 * t3.x - s1 * t4.x);
        u[0] = cadd(u[0], cadd(t1, t2));
        u[1] = cadd(m1, n1); u[4] = csub(m1, n1);
        u[2] = cadd(m2, n2); u[3] = csub(m2, n2);
    } else if (R == 7) {
        const double c1 = 0.62348980185873353, c2 = -0.22252093395631440, c3 = -0.90096886790241913;
        const double s1 = 0.78183148246802981 * SIGN, s2 = 0.97492791218182361 * SIGN,
                     s3 = 0.43388373911755812 * SIGN;
        const double2 t1 = cadd(u[1], u[6]), t2 = cadd(u[2], u[5]), t3 = cadd(u[3], u[4]);
        const double2 d1 = csub(u[1], u[6]), d2 = csub(u[2], u[5]), d3 = csub(u[3], u[4]);
        const double2 m1 = make_double2(u[0].x + c1 * t1.x + c2 * t2.x + c3 * t3.x, u[0].y + c1 * t1.y + c2 * t2.y + c3 * t3.y);
        const double2 m2 = make_double2(u[0].x + c2 * t1.x + c3 * t2.x + c1 * t3.x, u[0].y + c2 * t1.y + c3 * t2.y + c1 * t3.y);
        const double2 m3 = make_double2(u[0].x + c3 * t1.x + c1 * t2.x + c2 * t3.x, u[0].y + c3 * t1.y + c1 * t2.y + c2 * t3.y);
        // i * (s1 d1 + s2 d2 + s3 d3), i * (s2 d1 - s3 d2 - s1 d3), i * (s3 d1 - s1 d2 + s2 d3)
        const double2 n1 = make_double2(-(s1 * d1.y + s2 * d2.y + s3 * d3.y), s1 * d1.x + s2 * d2.x + s3 * d3.x);
        const double2 n2 = make_double2(-(s2 * d1.y - s3 * d2.y - s1 * d3.y), s2 * d1.x - s3 * d2.x - s1 * d3.x);
        const double2 n3 = make_double2(-(s3 * d1.y - s1 * d2.y + s2 * d3.y), s3 * d1.x - s1 * d2.x + s2 * d3.x);
        u[0] = cadd(cadd(u[0], t1), cadd(t2, t3));
        u[1] = cadd(m1, n1); u[6] = csub(m1, n1);
        u[2] = cadd(m2, n2); u[5] = csub(m2, n2);
        u[3] = cadd(m3, n3); u[4] = csub(m3, n3);
    } else if (R == 6) {
        dft_pair<2, 3, SIGN>(u, tw, n);
    } else if (R == 8) {
        dft_pair<2, 4, SIGN>(u, tw, n);
    } else if (R == 9) {
        dft_pair<3, 3, SIGN>(u, tw, n);
    } else if (R == 12) {
        dft_pair<3, 4, SIGN>(u, tw, n);
    } else if (R == 10) {
        // 10 = 2 x 5: two radix-5 butterflies on the even / odd inputs, then X[k], X[k+5] = E[k] +- W10^k O[k]
        double2 e[5], o[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) { e[j] = u[2 * j]; o[j] = u[2 * j + 1]; }
        dft_small<5, SIGN>(e, tw, n);
        dft_small<5, SIGN>(o, tw, n);
        // W10^k = (cos(2 pi k/10), SIGN sin(2 pi k/10)), k = 1..4
        const double2 t0 = o[0];
        const double2 t1 = cmul(o[1], make_double2(0.80901699437494742, SIGN * 0.58778525229247313));
        const double2 t2 = cmul(o[2], make_double2(0.30901699437494742, SIGN * 0.95105651629515357));
        const double2 t3 = cmul(o[3], make_double2(-0.30901699437494742, SIGN * 0.95105651629515357));
        const double2 t4 = cmul(o[4], make_double2(-0.80901699437494742, SIGN * 0.58778525229247313));
        u[0] = cadd(e[0], t0); u[5] = csub(e[0], t0);
        u[1] = cadd(e[1], t1); u[6] = csub(e[1], t1);
        u[2] = cadd(e[2], t2); u[7] = csub(e[2], t2);
        u[3] = cadd(e[3], t3); u[8] = csub(e[3], t3);
        u[4] = cadd(e[4], t4); u[9] = csub(e[4], t4);
    } else {
        double2 v[R];
        const int step = n / R;
#pragma unroll
        for (int q = 0; q < R; ++q) {
            double2 acc = u[0];
#pragma unroll
            for (int s = 1; s < R; ++s) acc = cadd(acc, cmul(u[s], twid<SIGN>(tw, ((q * s) % R) * step)));
            v[q] = acc;
        }
#pragma unroll
        for (int q = 0; q < R; ++q) u[q] = v[q];
    }
}

// DFT of size R = R1 R2 in registers (Cooley-Tukey): with s = R1 j + s1 and k = k2 + R2 k1,
//   X[k2 + R2 k1] = sum_s1 W_R1^(s1 k1) [ W_R^(s1 k2) sum_j u[R1 j + s1] W_R2^(j k2) ]
// i.e. R1 butterflies of size R2, a twiddle from the stage's table (R divides n), R2 butterflies of size R1.
template <int R1, int R2, int SIGN>
__device__ __forceinline__ void dft_pair(double2* u, const double2* tw, int n)
{
    constexpr int R = R1 * R2;
    double2 y[R1][R2];
#pragma unroll
    for (int s1 = 0; s1 < R1; ++s1) {
#pragma unroll
        for (int j = 0; j < R2; ++j) y[s1][j] = u[R1 * j + s1];
        dft_small<R2, SIGN>(y[s1], tw, n);
    }
    const int step = n / R;
#pragma unroll
    for (int k2 = 0; k2 < R2; ++k2) {
        double2 v[R1];
#pragma unroll
        for (int s1 = 0; s1 < R1; ++s1)
            v[s1] = (s1 == 0 || k2 == 0) ? y[s1][k2] : cmul(y[s1][k2], twid<SIGN>(tw, s1 * k2 * step));
        dft_small<R1, SIGN>(v, tw, n);
#pragma unroll
        for (int k1 = 0; k1 < R1; ++k1) u[k2 + R2 * k1] = v[k1];
    }
}

// exact a / b for 0 <= a < 2^24, b > 0, with a precomputed float reciprocal
__device__ __forceinline__ int fast_div(int a, int b, float invb)
{
    int q = __float2int_rz(__int2float_rn(a) * invb);
    const int r = a - q * b;
    if (r < 0) --q;
    else if (r >= b) ++q;
    return q;
}

template <int R, int SIGN>
__device__ __forceinline__ void stage_small(const double2* __restrict__ in, double2* __restrict__ out,
                                            const Radix& pl, int p, int nlines, int line_stride,
                                            int elem_stride, bool line_fast)
{
    const int n = pl.n, t = n / R;
    const int tws = n / (p * R);
    const float inv_nl = 1.0f / (float)nlines, inv_t = 1.0f / (float)t, inv_p = 1.0f / (float)p;
    for (int idx = threadIdx.x; idx < nlines * t; idx += blockDim.x) {
        int line, i;
        if (line_fast) { i = fast_div(idx, nlines, inv_nl); line = idx - i * nlines; }
        else { line = fast_div(idx, t, inv_t); i = idx - line * t; }
        const int k = i - fast_div(i, p, inv_p) * p;
        const int j = (i - k) * R + k;
        const double2* src = in + (long long)line * line_stride;
        double2* dst = out + (long long)line * line_stride;
        double2 u[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            double2 x = src[(i + q * t) * elem_stride];
            if (q > 0 && k > 0) x = cmul(x, twid<SIGN>(pl.tw, q * k * tws));
            u[q] = x;
        }
        dft_small<R, SIGN>(u, pl.tw, n);
#pragma unroll
        for (int q = 0; q < R; ++q) dst[(j + q * p) * elem_stride] = u[q];
    }
}

// (TAG separates the copies called from the FUSED kernels, so that the number of call sites -- and with
// it nvcc's inlining of the default kernels -- is what it was before the fused path existed)
template <int SIGN, int TAG = 0>
__device__ void stage_big(const double2* __restrict__ in, double2* __restrict__ out, const Radix& pl,
                          int r, int p, int nlines, int line_stride, int elem_stride, bool line_fast)
{
    const int n = pl.n, t = n / r;
    const int tws = n / (p * r), step = n / r;
    // one output element per work item: direct O(r^2) DFT for a large prime factor
    for (int idx = threadIdx.x; idx < nlines * t * r; idx += blockDim.x) {
        int line, rest;
        if (line_fast) { rest = idx / nlines; line = idx - rest * nlines; }
        else { line = idx / (t * r); rest = idx - line * (t * r); }
        const int q = rest / t, i = rest - q * t;
        const int k = i % p;
        const int j = (i - k) * r + k;
        const double2* src = in + (long long)line * line_stride;
        double2 acc = make_double2(0.0, 0.0);
        for (int s = 0; s < r; ++s) {
            double2 x = src[(i + s * t) * elem_stride];
            // twiddle w^(s*k*tws) * W_r^(q*s)
            const int e = (int)(((long long)s * k * tws + (long long)((q * s) % r) * step) % n);
            x = cmul(x, twid<SIGN>(pl.tw, e));
            acc = cadd(acc, x);
        }
        out[(long long)line * line_stride + (j + q * p) * elem_stride] = acc;
    }
}

// FFT of `nlines` lines living in shared memory; returns the buffer holding the result.
template <int SIGN, int TAG = 0>
__device__ double2* fft_radix_smem(double2* a, double2* b, const Radix& pl, int nlines,
                                   int line_stride, int elem_stride, bool line_fast)
{
    int p = 1;
    for (int s = 0; s < pl.nrad; ++s) {
        const int r = pl.rad[s];
        switch (r) {
        case 2: stage_small<2, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 3: stage_small<3, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 4: stage_small<4, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 5: stage_small<5, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 7: stage_small<7, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        default: stage_big<SIGN, TAG>(a, b, pl, r, p, nlines, line_stride, elem_stride, line_fast); break;
        }
        __syncthreads();
        double2* tmp = a; a = b; b = tmp;
        p *= r;
    }
    return a;
}

// The same driver with the wider radix set (6, 8, 10 on top of 2, 3, 4, 5, 7): composite lengths under the
// FUSED kernels (plans with rader == 3, built by factorize_wide): 96 = 8 x 6 x 2 and 108 = 6 x 6 x 3 are three
// passes over shared memory instead of four, 64 = 8 x 8 two instead of three.
template <int SIGN>
__device__ double2* fft_radix_wide_smem(double2* a, double2* b, const Radix& pl, int nlines,
                                        int line_stride, int elem_stride, bool line_fast)
{
    int p = 1;
    for (int s = 0; s < pl.nrad; ++s) {
        const int r = pl.rad[s];
        switch (r) {
        case 2: stage_small<2, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 3: stage_small<3, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 4: stage_small<4, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 5: stage_small<5, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 6: stage_small<6, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 7: stage_small<7, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 8: stage_small<8, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        case 10: stage_small<10, SIGN>(a, b, pl, p, nlines, line_stride, elem_stride, line_fast); break;
        default: stage_big<SIGN, 1>(a, b, pl, r, p, nlines, line_stride, elem_stride, line_fast); break;
        }
        __syncthreads();
        double2* tmp = a; a = b; b = tmp;
        p *= r;
    }
    return a;
}

// Rader transform of prime-length lines (slots n and n+1 of buffer `a` hold x[0] and X[0]).
template <int SIGN>
__device__ double2* rader_smem(double2* a, double2* b, const FftPlan& pl, int nlines,
                               int ls, int es, bool line_fast)
{
    const int n = pl.main.n, m = n - 1;
    const double2* bhat = SIGN < 0 ? pl.bhat_f : pl.bhat_b;
    for (int idx = threadIdx.x; idx < nlines * m; idx += blockDim.x) {
        int line, q;
        if (line_fast) { q = idx / nlines; line = idx - q * nlines; }
        else { line = idx / m; q = idx - line * m; }
        b[(long long)line * ls + q * es] = a[(long long)line * ls + pl.perm[q] * es];
    }
    __syncthreads();
    for (int line = threadIdx.x; line < nlines; line += blockDim.x)
        a[(long long)line * ls + n * es] = a[(long long)line * ls];
    __syncthreads();
    double2* r = fft_radix_smem<-1>(b, a, pl.sub, nlines, ls, es, line_fast);
    double2* o = (r == a) ? b : a;
    for (int idx = threadIdx.x; idx < nlines * m; idx += blockDim.x) {
        int line, q;
        if (line_fast) { q = idx / nlines; line = idx - q * nlines; }
        else { line = idx / m; q = idx - line * m; }
        const double2 v = r[(long long)line * ls + q * es];
        if (q == 0) a[(long long)line * ls + (n + 1) * es] = cadd(a[(long long)line * ls + n * es], v);
        r[(long long)line * ls + q * es] = cmul(v, bhat[q]);
    }
    __syncthreads();
    double2* r2 = fft_radix_smem<+1>(r, o, pl.sub, nlines, ls, es, line_fast);
    double2* o2 = (r2 == a) ? b : a;
    for (int idx = threadIdx.x; idx < nlines * m; idx += blockDim.x) {
        int line, q;
        if (line_fast) { q = idx / nlines; line = idx - q * nlines; }
        else { line = idx / m; q = idx - line * m; }
        o2[(long long)line * ls + pl.iperm[q] * es] =
            cadd(a[(long long)line * ls + n * es], r2[(long long)line * ls + q * es]);
    }
    for (int line = threadIdx.x; line < nlines; line += blockDim.x)
        o2[(long long)line * ls] = a[(long long)line * ls + (n + 1) * es];
    __syncthreads();
    return o2;
}

// ---- Rader with the data movement folded into the sub-FFT stages (FftPlan::rader == 2) ------------------
// rader_smem above makes 3 + 2*nrad passes over the lines in shared memory (gather by g^q, nrad
// forward stages, multiply by the kernel spectrum, nrad inverse stages, scatter by g^-q).  Here the
// gather is the load of the first forward stage (IN = 1), the multiply is the store of the last forward
// stage (OUT = 1), the scatter is the store of the last inverse stage (OUT = 2), and n-1 = 70 is done as
// 10 x 7 instead of 2 x 5 x 7: 4 passes instead of 9 for n = 71.  The arithmetic per element is the
// same as in rader_smem except for the association inside the radix-10 butterfly.
//   a0: the buffer the lines arrived in; its slots n and n+1 keep x[0] and X[0] = x[0] + sum x[j].
template <int R, int DIR, int SIGN, int IN, int OUT>
__device__ __forceinline__ void stage_fused(const double2* __restrict__ in, double2* __restrict__ out,
                                            double2* __restrict__ a0, const FftPlan& fp, int p,
                                            int nlines, int ls, int es, bool line_fast)
{
    const Radix& pl = fp.sub;
    const int m = pl.n, t = m / R, n = fp.main.n;
    const int tws = m / (p * R);
    const double2* bhat = SIGN < 0 ? fp.bhat_f : fp.bhat_b;      // OUT == 1 only (forward sub-FFT of either direction)
    const float inv_nl = 1.0f / (float)nlines, inv_t = 1.0f / (float)t, inv_p = 1.0f / (float)p;
    for (int idx = threadIdx.x; idx < nlines * t; idx += blockDim.x) {
        int line, i;
        if (line_fast) { i = fast_div(idx, nlines, inv_nl); line = idx - i * nlines; }
        else { line = fast_div(idx, t, inv_t); i = idx - line * t; }
        const int k = i - fast_div(i, p, inv_p) * p;
        const int j = (i - k) * R + k;
        const long long lo = (long long)line * ls;
        double2 u[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int e = i + q * t;
            double2 x = IN == 1 ? in[lo + (long long)fp.perm[e] * es] : in[lo + (long long)e * es];
            if (q > 0 && k > 0) x = cmul(x, twid<DIR>(pl.tw, q * k * tws));
            u[q] = x;
        }
        if (IN == 1 && i == 0) a0[lo + (long long)n * es] = a0[lo];               // keep x[0]
        dft_small<R, DIR>(u, pl.tw, m);
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int jo = j + q * p;
            if (OUT == 0) out[lo + (long long)jo * es] = u[q];
            else if (OUT == 1) {
                if (jo == 0) a0[lo + (long long)(n + 1) * es] = cadd(a0[lo + (long long)n * es], u[q]);
                out[lo + (long long)jo * es] = cmul(u[q], bhat[jo]);
            } else {
                out[lo + (long long)fp.iperm[jo] * es] = cadd(a0[lo + (long long)n * es], u[q]);
                if (jo == 0) out[lo] = a0[lo + (long long)(n + 1) * es];
            }
        }
    }
}

// one sub-FFT of the fused path: DIR = -1 forward (first = gather, last = multiply), +1 inverse (last = scatter)
template <int DIR, int SIGN>
__device__ double2* subfft_fused(double2* a, double2* b, double2* a0, const FftPlan& fp, int nlines,
                                 int ls, int es, bool line_fast)
{
    const Radix& pl = fp.sub;
    int p = 1;
    for (int s = 0; s < pl.nrad; ++s) {
        const int r = pl.rad[s];
        const bool first = s == 0, last = s == pl.nrad - 1;
#define ISDF_FUSED_STAGE(R_)                                                                                  \
        do {                                                                                                  \
            if (DIR < 0) {                                                                                    \
                if (first && last) stage_fused<R_, -1, SIGN, 1, 1>(a, b, a0, fp, p, nlines, ls, es, line_fast); \
                else if (first) stage_fused<R_, -1, SIGN, 1, 0>(a, b, a0, fp, p, nlines, ls, es, line_fast);  \
                else if (last) stage_fused<R_, -1, SIGN, 0, 1>(a, b, a0, fp, p, nlines, ls, es, line_fast);   \
                else stage_fused<R_, -1, SIGN, 0, 0>(a, b, a0, fp, p, nlines, ls, es, line_fast);             \
            } else {                                                                                          \
                if (last) stage_fused<R_, +1, SIGN, 0, 2>(a, b, a0, fp, p, nlines, ls, es, line_fast);        \
                else stage_fused<R_, +1, SIGN, 0, 0>(a, b, a0, fp, p, nlines, ls, es, line_fast);             \
            }                                                                                                 \
        } while (0)
        switch (r) {
        case 2: ISDF_FUSED_STAGE(2); break;
        case 3: ISDF_FUSED_STAGE(3); break;
        case 4: ISDF_FUSED_STAGE(4); break;
        case 5: ISDF_FUSED_STAGE(5); break;
        case 6: ISDF_FUSED_STAGE(6); break;
        case 7: ISDF_FUSED_STAGE(7); break;
        case 8: ISDF_FUSED_STAGE(8); break;
        default: ISDF_FUSED_STAGE(10); break;
        }
#undef ISDF_FUSED_STAGE
        __syncthreads();
        double2* tmp = a; a = b; b = tmp;
        p *= r;
    }
    return a;
}

template <int SIGN>
__device__ double2* rader_fused_smem(double2* a, double2* b, const FftPlan& pl, int nlines,
                                     int ls, int es, bool line_fast)
{
    // forward sub-FFT reads `a` through the permutation (first stage a -> b), so x[0] in a's slot 0
    // survives until that stage has copied it to slot n
    double2* r = subfft_fused<-1, SIGN>(a, b, a, pl, nlines, ls, es, line_fast);
    double2* o = (r == a) ? b : a;
    return subfft_fused<+1, SIGN>(r, o, a, pl, nlines, ls, es, line_fast);
}

// FUSED is a compile-time flag of the calling kernel: the default kernels (FUSED = false) contain none
// of the fused code, so their registers and occupancy are those measured in round 1.
template <int SIGN, bool FUSED>
__device__ __forceinline__ double2* fft_smem(double2* a, double2* b, const FftPlan& pl, int nlines,
                                             int line_stride, int elem_stride, bool line_fast)
{
    if constexpr (FUSED) {
        if (pl.rader == 2) return rader_fused_smem<SIGN>(a, b, pl, nlines, line_stride, elem_stride, line_fast);
        return fft_radix_wide_smem<SIGN>(a, b, pl.main, nlines, line_stride, elem_stride, line_fast);
    }
    if (pl.rader) return rader_smem<SIGN>(a, b, pl, nlines, line_stride, elem_stride, line_fast);
    return fft_radix_smem<SIGN>(a, b, pl.main, nlines, line_stride, elem_stride, line_fast);
}

// ---- host side: tables of one axis ----------------------------------------------------------------
inline int factorize(int n, int* rad)
{
    int k = 0;
    const int small[5] = {4, 2, 3, 5, 7};
    for (int f : small) while (n % f == 0 && k < MAX_RAD) { rad[k++] = f; n /= f; }
    for (int p = 11; n > 1; p += 2) while (n % p == 0) { if (k >= MAX_RAD || p > BIG_RADIX_MAX) return -1; rad[k++] = p; n /= p; }
    return k;
}

// Factorisation for the FUSED kernels: fewest stages with radices from {10, 8, 7, 6, 5, 4, 3, 2} (the
// two-level butterflies 9 and 12 exist in dft_small but cost 100-128 registers per thread; without them the
// FUSED kernels fit 3 CTAs per SM like the default ones)
// (ties: the smaller largest radix, i.e. fewer registers), larger primes <= BIG_RADIX_MAX appended for the
// direct-DFT stage.  Returns the number of stages, or -1.  `small_only`: fail instead of appending primes.
inline int factorize_wide(int n, int* rad, bool small_only)
{
    int big[MAX_RAD], nbig = 0;
    int m = n;
    for (int p = 11; p <= m; p += 2) {
        bool prime = true;
        for (int q = 3; q * q <= p; q += 2) if (p % q == 0) { prime = false; break; }
        if (!prime) continue;
        while (m % p == 0) { if (small_only || p > BIG_RADIX_MAX || nbig >= MAX_RAD) return -1; big[nbig++] = p; m /= p; }
    }
    // m is now 2^a 3^b 5^c 7^d: depth-first search over the radix set (m <= 1024: tiny)
    const int cand[8] = {10, 8, 7, 6, 5, 4, 3, 2};
    int best[MAX_RAD], nbest = MAX_RAD + 1, bestmax = 1 << 30;
    int cur[MAX_RAD];
    struct Rec {
        static void go(int m, int depth, int from, int curmax, int* cur, const int* cand, int* best, int& nbest, int& bestmax)
        {
            if (m == 1) {
                if (depth < nbest || (depth == nbest && curmax < bestmax)) {
                    nbest = depth; bestmax = curmax;
                    for (int i = 0; i < depth; ++i) best[i] = cur[i];
                }
                return;
            }
            if (depth + 1 > nbest || depth >= MAX_RAD) return;
            for (int c = from; c < 8; ++c)
                if (m % cand[c] == 0) {
                    cur[depth] = cand[c];
                    go(m / cand[c], depth + 1, c, curmax > cand[c] ? curmax : cand[c], cur, cand, best, nbest, bestmax);
                }
        }
    };
    if (m > 1) Rec::go(m, 0, 0, 0, cur, cand, best, nbest, bestmax);
    else nbest = 0;
    if (nbest > MAX_RAD || nbest + nbig > MAX_RAD) return -1;
    int k = 0;
    for (int i = 0; i < nbest; ++i) rad[k++] = best[i];
    for (int i = 0; i < nbig; ++i) rad[k++] = big[i];
    return k;
}

inline bool is_prime(int n)
{
    if (n < 2) return false;
    for (int p = 2; p * p <= n; ++p) if (n % p == 0) return false;
    return true;
}

inline int powmod(long long b, int e, int n)
{
    long long r = 1; b %= n;
    for (; e > 0; e >>= 1) { if (e & 1) r = r * b % n; b = b * b % n; }
    return (int)r;
}

inline int primitive_root(int n)
{
    std::vector<int> fac;
    int m = n - 1;
    for (int p = 2; p * p <= m; ++p) if (m % p == 0) { fac.push_back(p); while (m % p == 0) m /= p; }
    if (m > 1) fac.push_back(m);
    for (int g = 2; g < n; ++g) {
        bool ok = true;
        for (int f : fac) if (powmod(g, (n - 1) / f, n) == 1) { ok = false; break; }
        if (ok) return g;
    }
    return -1;
}

inline void push_twiddles(std::vector<double2>& tw, int n)
{
    for (int k = 0; k < n; ++k) {
        const long double ang = -2.0L * 3.141592653589793238462643383279502884L * k / n;
        tw.push_back(make_double2((double)cosl(ang), (double)sinl(ang)));
    }
}


// Appends the tables of an n-point axis to tw / iv and fills everything of `pl` except the pointers.
// Returns 0, or -1 if n has a prime factor the kernels do not support.  `fused` requests the plans of
// the FUSED kernels where they apply: prime n > 7 whose n-1 is 2^a 3^b 5^c 7^d (rader = 2), composite n
// that needs fewer passes with the wide radix set (rader = 3).
inline int build_axis_tables(int n, int fused, FftPlan& pl, std::vector<double2>& tw, std::vector<int>& iv, AxisOffsets& off)
{
    memset(&pl, 0, sizeof(pl));
    memset(&off, 0, sizeof(off));
    pl.main.n = n;
    off.tw = tw.size();
    push_twiddles(tw, n);
    int sub_rad[MAX_RAD];
    const int sub_nrad = (n > 7 && is_prime(n)) ? factorize(n - 1, sub_rad) : -1;
    if (sub_nrad >= 0) {
        // Rader: DFT_n as a cyclic convolution of length m = n-1
        const int m = n - 1, g = primitive_root(n), ginv = powmod(g, n - 2, n);
        pl.rader = 1;
        pl.main.nrad = 0;
        pl.sub.n = m; pl.sub.nrad = sub_nrad;
        for (int i = 0; i < sub_nrad; ++i) pl.sub.rad[i] = sub_rad[i];
        int frad[MAX_RAD];
        const int fn = fused ? factorize_wide(m, frad, true) : -1;
        if (fn > 0) {
            pl.rader = 2;
            pl.sub.nrad = fn;
            for (int i = 0; i < fn; ++i) pl.sub.rad[i] = frad[i];
        }
        off.sub = tw.size();
        push_twiddles(tw, m);
        off.perm = iv.size();
        for (int q = 0; q < m; ++q) iv.push_back(powmod(g, q, n));
        off.iperm = iv.size();
        for (int q = 0; q < m; ++q) iv.push_back(powmod(ginv, q, n));
        const long double PI2 = 2.0L * 3.141592653589793238462643383279502884L;
        for (int sgn = -1; sgn <= 1; sgn += 2) {
            (sgn < 0 ? off.bf : off.bb) = tw.size();
            for (int k = 0; k < m; ++k) {
                // bhat[k] = (1/m) sum_j w^(ginv^j) e^{-2 pi i j k / m},  w = e^{sgn 2 pi i / n}
                long double re = 0, im = 0;
                for (int j = 0; j < m; ++j) {
                    const long double ang = sgn * PI2 * powmod(ginv, j, n) / n - PI2 * ((long long)j * k % m) / m;
                    re += cosl(ang); im += sinl(ang);
                }
                tw.push_back(make_double2((double)(re / m), (double)(im / m)));
            }
        }
    } else {
        pl.main.nrad = factorize(n, pl.main.rad);
        if (pl.main.nrad < 0) return -1;
        if (fused) {
            int frad[MAX_RAD];
            const int fn = factorize_wide(n, frad, false);
            if (fn >= 0 && fn < pl.main.nrad) {          // fewer passes with the wide radix set
                pl.rader = 3;
                pl.main.nrad = fn;
                for (int i = 0; i < fn; ++i) pl.main.rad[i] = frad[i];
            }
        }
    }
    return 0;
}

// Points the plan at its tables once they sit at their final address (device memory, or the host
// vectors themselves in the emulation harness).
inline void bind_axis_tables(FftPlan& pl, const double2* tw_base, const int* iv_base, const AxisOffsets& off)
{
    pl.main.tw = tw_base + off.tw;
    if (pl.rader) {
        pl.sub.tw = tw_base + off.sub;
        pl.bhat_f = tw_base + off.bf;
        pl.bhat_b = tw_base + off.bb;
        pl.perm = iv_base + off.perm;
        pl.iperm = iv_base + off.iperm;
    }
}

}  // namespace
