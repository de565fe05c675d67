// Stage 4: batched 3-D real FFT of the auxiliary functions fused with the
// Coulomb kernel 4 pi/|G|^2, and the Coulomb metric W.
//
// Conventions restated from the reference: forward DFT unnormalised, inverse
// carries 1/N (pyscf/pbc/tools/pbc.py:149-211); G vectors in FFT index order
// (pyscf/pbc/gto/cell.py:523-596); coulG = 4 pi/|G|^2 with G=0 -> 0
// (pyscf/pbc/tools/pbc.py:304,352-356).
//
// Hand-written mixed-radix Stockham FFTs in shared memory (radix 4/2/3/5/7 and
// a direct DFT for any other prime factor, so odd and prime meshes such as
// 35^3 or 71^3 work).  A 3-D transform is three passes, each reading and
// writing coalesced row segments:
//   z: real lines -> half spectrum (nz/2+1)          [R2C]
//   y: columns of every (x) plane
//   x: columns of the whole [nx, ny*nzh] matrix, with the Coulomb-kernel
//      multiply fused (and, for V, the inverse x transform in the same kernel).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <vector>
#include <cstdlib>
#include <cmath>
#include <algorithm>

namespace {

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
    int rader;              // 1: main.n is prime, use sub / perm / bhat
    Radix sub;              // length n-1
    const int* perm;        // g^p mod n,      p < n-1
    const int* iperm;       // g^-q mod n,     q < n-1
    const double2* bhat_f;  // FFT_{n-1}(w^(g^-m)) / (n-1) for the forward transform (w = e^{-2 pi i/n})
    const double2* bhat_b;  // same for the inverse transform (w = e^{+2 pi i/n})
    __host__ __device__ int n() const { return main.n; }
    __host__ __device__ int lalloc() const { return main.n + (rader ? 2 : 0); }   // smem slots per line
};

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

// In-place small DFT of u[0..R) (R <= 7), sign SIGN
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

template <int SIGN>
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
template <int SIGN>
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
        default: stage_big<SIGN>(a, b, pl, r, p, nlines, line_stride, elem_stride, line_fast); break;
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

template <int SIGN>
__device__ __forceinline__ double2* fft_smem(double2* a, double2* b, const FftPlan& pl, int nlines,
                                             int line_stride, int elem_stride, bool line_fast)
{
    if (pl.rader) return rader_smem<SIGN>(a, b, pl, nlines, line_stride, elem_stride, line_fast);
    return fft_radix_smem<SIGN>(a, b, pl.main, nlines, line_stride, elem_stride, line_fast);
}

struct CoulGeom {
    int mesh[3];
    int nzh;
    double b[9];        // reciprocal vectors (rows)
    double scale;       // multiplies the kernel (1/ng for V)
    int col0;           // first (y,z)-column of this rank's slab in the x pass (0 when not sharded)
};

__device__ __forceinline__ double coulG_raw(const CoulGeom& g, int fx, int fy, int fz)
{
    const double gx = fx * g.b[0] + fy * g.b[3] + fz * g.b[6];
    const double gy = fx * g.b[1] + fy * g.b[4] + fz * g.b[7];
    const double gz = fx * g.b[2] + fy * g.b[5] + fz * g.b[8];
    const double g2 = gx * gx + gy * gy + gz * gz;
    return (fx == 0 && fy == 0 && fz == 0) ? 0.0 : 4.0 * 3.14159265358979323846 / g2;
}

__device__ __forceinline__ int freq_of(int i, int n) { return i < (n + 1) / 2 ? i : i - n; }

// Coulomb kernel at FFT index (ix,iy,iz).  The reference multiplies the full
// complex spectrum by coulG(G) and keeps the real part of the inverse FFT
// (pyscf/pbc/df/fft_jk.py:72-74), which is the same as using the average of
// coulG at G and at the index-negated -G.  The two differ only on the Nyquist
// planes of even meshes in non-orthogonal cells.
__device__ __forceinline__ double coulG_at(const CoulGeom& g, int ix, int iy, int iz)
{
    const int fx = freq_of(ix, g.mesh[0]), fy = freq_of(iy, g.mesh[1]), fz = freq_of(iz, g.mesh[2]);
    const int mx = freq_of((g.mesh[0] - ix) % g.mesh[0], g.mesh[0]);
    const int my = freq_of((g.mesh[1] - iy) % g.mesh[1], g.mesh[1]);
    const int mz = freq_of((g.mesh[2] - iz) % g.mesh[2], g.mesh[2]);
    const double c0 = coulG_raw(g, fx, fy, fz);
    if (mx == -fx && my == -fy && mz == -fz) return c0;
    return 0.5 * (c0 + coulG_raw(g, mx, my, mz));
}

// ---- z pass: real lines -> half spectrum ------------------------------------
// Two real lines share one complex transform: z = x_even + i x_odd, then
// X_even[k] = (Z[k] + conj Z[n-k]) / 2 and X_odd[k] = -i (Z[k] - conj Z[n-k]) / 2.
// LB = complex (paired) lines per CTA, i.e. 2 LB real lines.
__global__ void __launch_bounds__(FFT_THREADS)
fft_z_r2c_kernel(const double* __restrict__ in, long long ldin, double2* __restrict__ out, long long ldout,
                 FftPlan pl, int nlines, int LB)
{
    extern __shared__ double2 sm[];
    const int nz = pl.n(), nzh = nz / 2 + 1, LS = pl.lalloc();
    double2* a = sm; double2* b = sm + LB * LS;
    const int line0 = blockIdx.x * 2 * LB;
    const int nl = min(2 * LB, nlines - line0);          // real lines of this CTA
    const int np = (nl + 1) >> 1;                        // complex lines
    const double* src = in + (long long)blockIdx.y * ldin + (long long)line0 * nz;
    const float inv_nz = 1.0f / (float)nz, inv_nzh = 1.0f / (float)nzh;
    for (int idx = threadIdx.x; idx < np * nz; idx += blockDim.x) {
        const int p = fast_div(idx, nz, inv_nz), e = idx - p * nz;
        const double re = src[(long long)(2 * p) * nz + e];
        const double im = (2 * p + 1 < nl) ? src[(long long)(2 * p + 1) * nz + e] : 0.0;
        a[p * LS + e] = make_double2(re, im);
    }
    __syncthreads();
    double2* r = fft_smem<-1>(a, b, pl, np, LS, 1, false);
    double2* dst = out + (long long)blockIdx.y * ldout + (long long)line0 * nzh;
    for (int idx = threadIdx.x; idx < nl * nzh; idx += blockDim.x) {
        const int line = fast_div(idx, nzh, inv_nzh), e = idx - line * nzh;
        const double2* rp = r + (line >> 1) * LS;
        const double2 zk = rp[e];
        double2 zc = rp[e ? nz - e : 0];
        zc.y = -zc.y;
        dst[idx] = (line & 1) ? make_double2(0.5 * (zk.y - zc.y), -0.5 * (zk.x - zc.x))
                              : make_double2(0.5 * (zk.x + zc.x), 0.5 * (zk.y + zc.y));
    }
}

// ---- z pass: half spectrum -> real lines (unnormalised inverse) ----------------
__global__ void __launch_bounds__(FFT_THREADS)
fft_z_c2r_kernel(const double2* __restrict__ in, long long ldin, double* __restrict__ out, long long ldout,
                 FftPlan pl, int nlines, int LB)
{
    extern __shared__ double2 sm[];
    const int nz = pl.n(), nzh = nz / 2 + 1, LS = pl.lalloc();
    double2* a = sm; double2* b = sm + LB * LS;
    const int line0 = blockIdx.x * LB;
    const int nl = min(LB, nlines - line0);
    const double2* src = in + (long long)blockIdx.y * ldin + (long long)line0 * nzh;
    for (int idx = threadIdx.x; idx < nl * nz; idx += blockDim.x) {
        const int line = idx / nz, e = idx - line * nz;
        double2 v;
        if (e < nzh) v = src[line * nzh + e];
        else { v = src[line * nzh + (nz - e)]; v.y = -v.y; }
        a[line * LS + e] = v;
    }
    __syncthreads();
    double2* r = fft_smem<+1>(a, b, pl, nl, LS, 1, false);
    double* dst = out + (long long)blockIdx.y * ldout + (long long)line0 * nz;
    for (int idx = threadIdx.x; idx < nl * nz; idx += blockDim.x) {
        const int line = idx / nz, e = idx - line * nz;
        dst[idx] = r[line * LS + e].x;
    }
}

// ---- column pass over a row-major [n, ncols] complex matrix --------------------
// MODE 0: forward            MODE 3: inverse (unnormalised)
// MODE 1: forward, then multiply by sqrt(w * coulG)   (x pass of the W build)
// MODE 2: forward, multiply by coulG*scale, inverse   (x pass of the V build)
template <int MODE>
__global__ void __launch_bounds__(FFT_THREADS)
fft_cols_kernel(double2* __restrict__ data, long long vec_stride, long long outer_stride,
                FftPlan pl, int ncols, int CH, CoulGeom geo)
{
    extern __shared__ double2 sm[];
    const int n = pl.n();
    double2* a = sm; double2* b = sm + pl.lalloc() * CH;
    const int c0 = blockIdx.x * CH;
    const int nc = min(CH, ncols - c0);
    double2* base = data + (long long)blockIdx.z * vec_stride + (long long)blockIdx.y * outer_stride + c0;
    const float inv_nc = 1.0f / (float)nc, inv_nzh = 1.0f / (float)geo.nzh;
    for (int idx = threadIdx.x; idx < n * nc; idx += blockDim.x) {
        const int i = fast_div(idx, nc, inv_nc), c = idx - i * nc;
        a[i * CH + c] = base[(long long)i * ncols + c];
    }
    __syncthreads();
    double2* r;
    if (MODE == 3) r = fft_smem<+1>(a, b, pl, nc, 1, CH, true);
    else r = fft_smem<-1>(a, b, pl, nc, 1, CH, true);
    if (MODE == 1 || MODE == 2) {
        for (int idx = threadIdx.x; idx < n * nc; idx += blockDim.x) {
            const int i = fast_div(idx, nc, inv_nc), c = idx - i * nc;
            const int col = geo.col0 + c0 + c;
            const int iy = fast_div(col, geo.nzh, inv_nzh), iz = col - iy * geo.nzh;
            double f = coulG_at(geo, i, iy, iz);
            if (MODE == 1) {
                const bool self = (iz == 0) || (2 * iz == geo.mesh[2]);
                f = sqrt(self ? f : 2.0 * f);
            } else f *= geo.scale;
            double2 v = r[i * CH + c];
            v.x *= f; v.y *= f;
            r[i * CH + c] = v;
        }
        __syncthreads();
        if (MODE == 2) {
            double2* other = (r == a) ? b : a;
            r = fft_smem<+1>(r, other, pl, nc, 1, CH, true);
        }
    }
    for (int idx = threadIdx.x; idx < n * nc; idx += blockDim.x) {
        const int i = fast_div(idx, nc, inv_nc), c = idx - i * nc;
        base[(long long)i * ncols + c] = r[i * CH + c];
    }
}

__global__ void symmetrize_upper_kernel(double* __restrict__ W, long long ld, int n)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i < n && j < n && j < i) W[(long long)i * ld + j] = W[(long long)j * ld + i];
}

// ---- host side -----------------------------------------------------------------
int factorize(int n, int* rad)
{
    int k = 0;
    const int small[5] = {4, 2, 3, 5, 7};
    for (int f : small) while (n % f == 0 && k < MAX_RAD) { rad[k++] = f; n /= f; }
    for (int p = 11; n > 1; p += 2) while (n % p == 0) { if (k >= MAX_RAD || p > BIG_RADIX_MAX) return -1; rad[k++] = p; n /= p; }
    return k;
}

struct Plans {
    FftPlan ax[3];
    double2* d_tw;      // one allocation: twiddles, Rader kernels, permutations
};

bool is_prime(int n)
{
    if (n < 2) return false;
    for (int p = 2; p * p <= n; ++p) if (n % p == 0) return false;
    return true;
}

int powmod(long long b, int e, int n)
{
    long long r = 1; b %= n;
    for (; e > 0; e >>= 1) { if (e & 1) r = r * b % n; b = b * b % n; }
    return (int)r;
}

int primitive_root(int n)
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

void push_twiddles(std::vector<double2>& tw, int n)
{
    for (int k = 0; k < n; ++k) {
        const long double ang = -2.0L * 3.141592653589793238462643383279502884L * k / n;
        tw.push_back(make_double2((double)cosl(ang), (double)sinl(ang)));
    }
}

int make_plans(isdf_handle* h, const int* mesh, Plans& P)
{
    std::vector<double2> tw;          // complex tables, offsets in double2 units
    std::vector<int> iv;              // integer tables
    size_t off_tw[3], off_sub[3] = {0, 0, 0}, off_bf[3] = {0, 0, 0}, off_bb[3] = {0, 0, 0};
    size_t off_perm[3] = {0, 0, 0}, off_iperm[3] = {0, 0, 0};
    for (int d = 0; d < 3; ++d) {
        const int n = mesh[d];
        ARG_CHECK(n >= 1 && n <= 1024);
        FftPlan& pl = P.ax[d];
        memset(&pl, 0, sizeof(pl));
        pl.main.n = n;
        off_tw[d] = tw.size();
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
            off_sub[d] = tw.size();
            push_twiddles(tw, m);
            off_perm[d] = iv.size();
            for (int q = 0; q < m; ++q) iv.push_back(powmod(g, q, n));
            off_iperm[d] = iv.size();
            for (int q = 0; q < m; ++q) iv.push_back(powmod(ginv, q, n));
            const long double PI2 = 2.0L * 3.141592653589793238462643383279502884L;
            for (int sgn = -1; sgn <= 1; sgn += 2) {
                (sgn < 0 ? off_bf[d] : off_bb[d]) = tw.size();
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
            if (pl.main.nrad < 0) { isdf_set_error("mesh size %d has an unsupported prime factor", n); return ISDF_ERR_ARG; }
        }
    }
    const size_t bytes_tw = align_up(sizeof(double2) * tw.size(), 256);
    const size_t bytes_iv = sizeof(int) * std::max<size_t>(iv.size(), 1);
    CUDA_TRY(cudaMallocAsync((void**)&P.d_tw, bytes_tw + bytes_iv, h->stream));
    int* d_iv = (int*)((char*)P.d_tw + bytes_tw);
    CUDA_TRY(cudaMemcpyAsync(P.d_tw, tw.data(), sizeof(double2) * tw.size(), cudaMemcpyHostToDevice, h->stream));
    if (!iv.empty())
        CUDA_TRY(cudaMemcpyAsync(d_iv, iv.data(), sizeof(int) * iv.size(), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    for (int d = 0; d < 3; ++d) {
        FftPlan& pl = P.ax[d];
        pl.main.tw = P.d_tw + off_tw[d];
        if (pl.rader) {
            pl.sub.tw = P.d_tw + off_sub[d];
            pl.bhat_f = P.d_tw + off_bf[d];
            pl.bhat_b = P.d_tw + off_bb[d];
            pl.perm = d_iv + off_perm[d];
            pl.iperm = d_iv + off_iperm[d];
        }
    }
    return ISDF_OK;
}

template <typename K>
int set_smem(K kern, size_t smem)
{
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return ISDF_OK;
}

void fill_geom(CoulGeom& g, const int* mesh, const double* a, double scale)
{
    // reciprocal vectors b = 2 pi inv(a^T): rows b_i with a_i . b_j = 2 pi delta_ij
    const double a00 = a[0], a01 = a[1], a02 = a[2], a10 = a[3], a11 = a[4], a12 = a[5], a20 = a[6], a21 = a[7], a22 = a[8];
    const double det = a00 * (a11 * a22 - a12 * a21) - a01 * (a10 * a22 - a12 * a20) + a02 * (a10 * a21 - a11 * a20);
    const double f = 2.0 * 3.14159265358979323846 / det;
    // b_0 = 2pi (a1 x a2)/det, etc.
    g.b[0] = f * (a11 * a22 - a12 * a21); g.b[1] = f * (a12 * a20 - a10 * a22); g.b[2] = f * (a10 * a21 - a11 * a20);
    g.b[3] = f * (a21 * a02 - a22 * a01); g.b[4] = f * (a22 * a00 - a20 * a02); g.b[5] = f * (a20 * a01 - a21 * a00);
    g.b[6] = f * (a01 * a12 - a02 * a11); g.b[7] = f * (a02 * a10 - a00 * a12); g.b[8] = f * (a00 * a11 - a01 * a10);
    for (int d = 0; d < 3; ++d) g.mesh[d] = mesh[d];
    g.nzh = mesh[2] / 2 + 1;
    g.scale = scale;
    g.col0 = 0;
}

// shared memory per CTA of the FFT kernels (two ping-pong line buffers); smaller = more CTAs per SM.
// Tuning knob: environment ISDF_FFT_SMEM_KB.
// shared memory per CTA: column passes (y, x) want long contiguous row segments (wide CTAs), the z
// pass reads whole lines and prefers more resident CTAs (tuning knobs ISDF_FFT_SMEM_KB / _Z_KB)
size_t fft_smem_budget(bool zpass = false)
{
    static size_t v[2] = {0, 0};
    if (v[0] == 0) {
        const char* e = getenv("ISDF_FFT_SMEM_KB");
        const char* ez = getenv("ISDF_FFT_SMEM_Z_KB");
        const long kb = e ? atol(e) : 64, kbz = ez ? atol(ez) : 64;
        v[0] = (size_t)std::min(200L, std::max(8L, kb)) * 1024;
        v[1] = (size_t)std::min(200L, std::max(8L, kbz)) * 1024;
    }
    return v[zpass ? 1 : 0];
}

// lines (or columns) per CTA such that two ping-pong buffers of `nalloc` slots per line fit the budget
int pick_lb(int nalloc, size_t budget) { return (int)std::max<size_t>(1, std::min<size_t>(64, budget / (2 * sizeof(double2) * nalloc))); }

// forward 3-D transform of nvec real vectors into half spectra (spec[vec*lds + ...]);
// xmode selects what the x pass fuses (0 plain, 1 sqrt(w coulG), 2 coulG*scale + inverse x)
int forward_passes(isdf_handle* h, const Plans& P, const int* mesh, const double* in, long long ldin,
                   int nvec, double2* spec, long long lds, int xmode, const CoulGeom& geo)
{
    const int nx = mesh[0], ny = mesh[1], nz = mesh[2], nzh = nz / 2 + 1;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    {
        const int LB = pick_lb(P.ax[2].lalloc(), fft_smem_budget(true));
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        int rc; if ((rc = set_smem(fft_z_r2c_kernel, smem))) return rc;
        dim3 grid((nx * ny + 2 * LB - 1) / (2 * LB), nvec);
        fft_z_r2c_kernel<<<grid, FFT_THREADS, smem, s>>>(in, ldin, spec, lds, P.ax[2], nx * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        int rc; if ((rc = set_smem(fft_cols_kernel<0>, smem))) return rc;
        dim3 grid((nzh + CH - 1) / CH, nx, nvec);
        fft_cols_kernel<0><<<grid, FFT_THREADS, smem, s>>>(spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    {
        const int ncols = ny * nzh;
        const int CH = std::min(pick_lb(P.ax[0].lalloc(), budget), ncols);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[0].lalloc();
        dim3 grid((ncols + CH - 1) / CH, 1, nvec);
        int rc;
        if (xmode == 1) {
            if ((rc = set_smem(fft_cols_kernel<1>, smem))) return rc;
            fft_cols_kernel<1><<<grid, FFT_THREADS, smem, s>>>(spec, lds, 0, P.ax[0], ncols, CH, geo);
        } else if (xmode == 2) {
            if ((rc = set_smem(fft_cols_kernel<2>, smem))) return rc;
            fft_cols_kernel<2><<<grid, FFT_THREADS, smem, s>>>(spec, lds, 0, P.ax[0], ncols, CH, geo);
        } else {
            if ((rc = set_smem(fft_cols_kernel<0>, smem))) return rc;
            fft_cols_kernel<0><<<grid, FFT_THREADS, smem, s>>>(spec, lds, 0, P.ax[0], ncols, CH, geo);
        }
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 3;
    return ISDF_OK;
}

// inverse y and z passes (x already inverted): spectra -> real vectors
int inverse_yz(isdf_handle* h, const Plans& P, const int* mesh, double2* spec, long long lds,
               int nvec, double* out, long long ldout, const CoulGeom& geo)
{
    const int nx = mesh[0], ny = mesh[1], nz = mesh[2], nzh = nz / 2 + 1;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        int rc; if ((rc = set_smem(fft_cols_kernel<3>, smem))) return rc;
        dim3 grid((nzh + CH - 1) / CH, nx, nvec);
        fft_cols_kernel<3><<<grid, FFT_THREADS, smem, s>>>(spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    {
        const int LB = pick_lb(P.ax[2].lalloc(), budget);
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        int rc; if ((rc = set_smem(fft_z_c2r_kernel, smem))) return rc;
        dim3 grid((nx * ny + LB - 1) / LB, nvec);
        fft_z_c2r_kernel<<<grid, FFT_THREADS, smem, s>>>(spec, lds, out, ldout, P.ax[2], nx * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 2;
    return ISDF_OK;
}

}  // namespace

extern "C" size_t isdf_build_W_ws_bytes(int np, const int* h_mesh)
{
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    return sizeof(double2) * ngh * (size_t)np + 4096;
}

extern "C" int isdf_coulomb_apply(isdf_handle* h, const double* in, long long ldin, int nvec,
                                  const int* h_mesh, const double* h_a,
                                  double* out, long long ldout, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && in && out && h_mesh && h_a && ws && nvec > 0);
    const long long ng = (long long)h_mesh[0] * h_mesh[1] * h_mesh[2];
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    ARG_CHECK(ldin >= ng && ldout >= ng);
    const size_t per_vec = sizeof(double2) * ngh;
    const int batch = (int)std::min<size_t>(nvec, ws_bytes / per_vec);
    if (batch < 1) { isdf_set_error("coulomb_apply workspace too small: need >= %zu", per_vec); return ISDF_ERR_WS; }
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0 / (double)ng);
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FFT], h->stream));
    for (int v0 = 0; v0 < nvec; v0 += batch) {
        const int nv = std::min(batch, nvec - v0);
        if ((rc = forward_passes(h, P, h_mesh, in + (long long)v0 * ldin, ldin, nv, (double2*)ws, (long long)ngh, 2, geo))) return rc;
        if ((rc = inverse_yz(h, P, h_mesh, (double2*)ws, (long long)ngh, nv, out + (long long)v0 * ldout, ldout, geo))) return rc;
    }
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FFT], h->stream));
    h->ev_used[ISDF_STAGE_FFT] = true;
    CUDA_TRY(cudaFreeAsync(P.d_tw, h->stream));
    return ISDF_OK;
}

extern "C" int isdf_build_W(isdf_handle* h, const double* zeta, int np, long long ldz,
                            const int* h_mesh, const double* h_a,
                            double* W, double* V, long long ldv,
                            void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && zeta && h_mesh && h_a && W && ws && np > 0);
    const long long ng = (long long)h_mesh[0] * h_mesh[1] * h_mesh[2];
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    ARG_CHECK(ldz >= ng);
    const size_t need = isdf_build_W_ws_bytes(np, h_mesh);
    if (need > ws_bytes) { isdf_set_error("build_W workspace too small: need %zu have %zu", need, ws_bytes); return ISDF_ERR_WS; }
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_COULOMB], s));
    if (V != nullptr) {
        ARG_CHECK(ldv >= ng);
        if ((rc = isdf_coulomb_apply(h, zeta, ldz, np, h_mesh, h_a, V, ldv, ws, ws_bytes))) return rc;
    }
    Plans P;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    const double det = h_a[0] * (h_a[4] * h_a[8] - h_a[5] * h_a[7]) - h_a[1] * (h_a[3] * h_a[8] - h_a[5] * h_a[6]) +
                       h_a[2] * (h_a[3] * h_a[7] - h_a[4] * h_a[6]);
    const double vol = fabs(det);
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0);
    double2* Y = (double2*)ws;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FFT], s));
    if ((rc = forward_passes(h, P, h_mesh, zeta, ldz, np, Y, (long long)ngh, 1, geo))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FFT], s));
    h->ev_used[ISDF_STAGE_FFT] = true;
    // W = (vol/ng^2) Y Y^T over the weighted half spectrum (upper triangle, then mirrored)
    GemmParams p;
    p.M = np; p.N = np; p.K = (int)(2 * ngh);
    p.A = (const double*)Y; p.lda = 2 * (long long)ngh; p.a_kmajor = 0;
    p.B = (const double*)Y; p.ldb = 2 * (long long)ngh; p.b_kmajor = 0;
    p.C = W; p.ldc = np; p.Cin = nullptr; p.ldcin = 0; p.E = nullptr; p.lde = 0;
    p.alpha = vol / ((double)ng * (double)ng); p.neg = 0; p.epi = EPI_STORE; p.upper_only = 1;
    if ((rc = gemm_launch(h, p))) return rc;
    dim3 blk(32, 8), grd((np + 31) / 32, (np + 7) / 8);
    symmetrize_upper_kernel<<<grd, blk, 0, s>>>(W, np, np);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_COULOMB], s));
    h->ev_used[ISDF_STAGE_COULOMB] = true;
    CUDA_TRY(cudaFreeAsync(P.d_tw, s));
    return ISDF_OK;
}

// ---- slab-decomposed 3-D transform (grid sharded over ranks by x-planes) --------------------
// Each rank holds x-planes [x_off[r], x_off[r+1]) of every vector.  z and y passes are local;
// one all-to-all (done by the caller, NCCL) re-shards the half spectra from x-planes to
// (y,z)-column slabs; the x pass with the fused Coulomb weight then runs on whole x lines.
namespace {

__global__ void fft_pack_kernel(const double2* __restrict__ spec, long long lds, int nxl, int ncols,
                                int c0, int ncl, double2* __restrict__ send)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncl) return;
    const int x = blockIdx.y, v = blockIdx.z;
    send[((long long)v * nxl + x) * ncl + c] = spec[(long long)v * lds + (long long)x * ncols + c0 + c];
}

__global__ void fft_unpack_kernel(const double2* __restrict__ recv, int nxl, int x0, int ncl,
                                  double2* __restrict__ Y, long long ldy)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncl) return;
    const int x = blockIdx.y, v = blockIdx.z;
    Y[(long long)v * ldy + (long long)(x0 + x) * ncl + c] = recv[((long long)v * nxl + x) * ncl + c];
}

}  // namespace

// local z (R2C) and y passes: in[nvec, ldin] real, x-planes of this rank -> spec[nvec][nxl][ny*nzh] complex
extern "C" int isdf_fft_zy(isdf_handle* h, const double* in, long long ldin, int nvec, int nxl,
                           const int* h_mesh, void* spec, long long lds)
{
    ARG_CHECK(h && in && h_mesh && spec && nvec > 0 && nxl > 0 && nxl <= h_mesh[0]);
    const int ny = h_mesh[1], nz = h_mesh[2], nzh = nz / 2 + 1;
    ARG_CHECK(ldin >= (long long)nxl * ny * nz && lds >= (long long)nxl * ny * nzh);
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    CoulGeom geo; memset(&geo, 0, sizeof(geo));
    {
        const int LB = pick_lb(P.ax[2].lalloc(), fft_smem_budget(true));
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        if ((rc = set_smem(fft_z_r2c_kernel, smem))) return rc;
        dim3 grid((nxl * ny + 2 * LB - 1) / (2 * LB), nvec);
        fft_z_r2c_kernel<<<grid, FFT_THREADS, smem, s>>>(in, ldin, (double2*)spec, lds, P.ax[2], nxl * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        if ((rc = set_smem(fft_cols_kernel<0>, smem))) return rc;
        dim3 grid((nzh + CH - 1) / CH, nxl, nvec);
        fft_cols_kernel<0><<<grid, FFT_THREADS, smem, s>>>((double2*)spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 2;
    CUDA_TRY(cudaFreeAsync(P.d_tw, s));
    return ISDF_OK;
}

// send buffer for the all-to-all: block of rank r = [nvec][nxl][h_col_off[r+1]-h_col_off[r]], blocks in rank order
extern "C" int isdf_fft_pack(isdf_handle* h, const void* spec, long long lds, int nvec, int nxl, int ncols,
                             const int* h_col_off, int nranks, void* send)
{
    ARG_CHECK(h && spec && h_col_off && send && nranks >= 1 && h_col_off[0] == 0 && h_col_off[nranks] == ncols);
    for (int r = 0; r < nranks; ++r) {
        const int c0 = h_col_off[r], ncl = h_col_off[r + 1] - c0;
        if (ncl <= 0) continue;
        double2* dst = (double2*)send + (long long)nvec * nxl * c0;
        dim3 grid((ncl + 127) / 128, nxl, nvec);
        fft_pack_kernel<<<grid, 128, 0, h->stream>>>((const double2*)spec, lds, nxl, ncols, c0, ncl, dst);
        h->other_launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

// receive buffer (block of rank r = [nvec][nx_r][ncl]) -> Y[nvec][nx][ncl] (row stride ldy, complex units)
extern "C" int isdf_fft_unpack(isdf_handle* h, const void* recv, int nvec, const int* h_x_off, int nranks,
                               int ncl, void* Y, long long ldy)
{
    ARG_CHECK(h && recv && h_x_off && Y && nranks >= 1 && ncl > 0 && h_x_off[0] == 0);
    ARG_CHECK(ldy >= (long long)h_x_off[nranks] * ncl);
    for (int r = 0; r < nranks; ++r) {
        const int x0 = h_x_off[r], nxl = h_x_off[r + 1] - x0;
        if (nxl <= 0) continue;
        const double2* src = (const double2*)recv + (long long)nvec * ncl * x0;
        dim3 grid((ncl + 127) / 128, nxl, nvec);
        fft_unpack_kernel<<<grid, 128, 0, h->stream>>>(src, nxl, x0, ncl, (double2*)Y, ldy);
        h->other_launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

// x pass over the column slab [col0, col0+ncl) of the (y,z) plane, fused with the
// sqrt(weight * coulG) factor of the W build: Y[nvec][nx][ncl] in place.
extern "C" int isdf_fft_x_coulomb(isdf_handle* h, void* Y, long long ldy, int nvec,
                                  const int* h_mesh, const double* h_a, int col0, int ncl)
{
    ARG_CHECK(h && Y && h_mesh && h_a && nvec > 0 && ncl > 0 && col0 >= 0);
    const int nx = h_mesh[0];
    ARG_CHECK(col0 + ncl <= h_mesh[1] * (h_mesh[2] / 2 + 1) && ldy >= (long long)nx * ncl);
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0);
    geo.col0 = col0;
    const size_t budget = fft_smem_budget();
    const int CH = std::min(pick_lb(P.ax[0].lalloc(), budget), ncl);
    const size_t smem = 2 * sizeof(double2) * CH * P.ax[0].lalloc();
    if ((rc = set_smem(fft_cols_kernel<1>, smem))) return rc;
    dim3 grid((ncl + CH - 1) / CH, 1, nvec);
    fft_cols_kernel<1><<<grid, FFT_THREADS, smem, h->stream>>>((double2*)Y, ldy, 0, P.ax[0], ncl, CH, geo);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    CUDA_TRY(cudaFreeAsync(P.d_tw, h->stream));
    return ISDF_OK;
}

extern "C" int isdf_symmetrize_upper(isdf_handle* h, double* W, long long ld, int n)
{
    ARG_CHECK(h && W && n > 0 && ld >= n);
    dim3 blk(32, 8), grd((n + 31) / 32, (n + 7) / 8);
    symmetrize_upper_kernel<<<grd, blk, 0, h->stream>>>(W, ld, n);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}
