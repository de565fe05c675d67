// Residue arithmetic for the second int8 form of the metric product W = alpha A A^T
// (isdf_syrk_i8 with ISDF_SYRK_CRT=1; "Ozaki scheme II": one int8 GEMM per modulus instead of one per
// pair of digit planes).  Per K chunk every row is scaled by its power of two and rounded to a B-bit
// integer; the integer matrix is reduced modulo NM pairwise coprime moduli <= 256 (centred residues fit
// int8), each residue matrix is multiplied with itself exactly on the int8 tensor cores (int32
// accumulation, (128^2) * 32768 < 2^31), the int32 results are reduced modulo the same moduli, and the
// integer dot products -- bounded by kc * 2^(2B) < P/2, P the product of the moduli -- are rebuilt from
// their residues by Garner's mixed-radix algorithm with centred digits and a Horner sum in FP64 (the
// high digits of a small value are zero, so the Horner sum is accurate relative to the value itself).
// 16 moduli carry B = 53 bits: full FP64 inputs for 16 tensor-core products instead of 36.
//
// Like fft_core.cuh this header also compiles with g++ (tests/emul/crt_emul.cpp), where the per-element
// routines are checked against exact big-integer arithmetic.
#pragma once
#include <cstdint>
#include <cmath>

namespace isdf_crt {

constexpr int NM_MAX = 16;
// pairwise coprime: 2^8, 3*5*17, 11*23, 251, 13*19, 241, 239, 233, 229, 227, 223, 7*31, 211, 199, 197, 193
__host__ __device__ constexpr int modulus(int k)
{
    constexpr int p[NM_MAX] = {256, 255, 253, 251, 247, 241, 239, 233, 229, 227, 223, 217, 211, 199, 197, 193};
    return p[k];
}

// x^-1 mod m for small coprime x, m (compile-time loop; the tables below are built from it)
__host__ __device__ constexpr int inv_mod(int x, int m)
{
    x %= m;
    for (int y = 1; y < m; ++y) if ((x * y) % m == 1) return y;
    return 0;
}

// Garner tables: M[k][j] = (p_0 ... p_{j-1}) mod p_k for j < k, c[k] = (p_0 ... p_{k-1})^-1 mod p_k
struct GarnerTable { int M[NM_MAX][NM_MAX]; int c[NM_MAX]; };
__host__ __device__ constexpr GarnerTable make_garner_table()
{
    GarnerTable t{};
    for (int k = 0; k < NM_MAX; ++k) {
        int prod = 1 % modulus(k);
        for (int j = 0; j < NM_MAX; ++j) {
            t.M[k][j] = j < k ? prod : 0;
            if (j < k) prod = prod * (modulus(j) % modulus(k)) % modulus(k);
        }
        t.c[k] = k == 0 ? 1 : inv_mod(prod, modulus(k));
    }
    return t;
}

// Input bits B for nm moduli and chunks of at most 2^15 columns: the largest B <= 53 with
// 2^15 * 2^(2B) <= P/2 - 1 (host side, called once per product)
inline int input_bits(int nm)
{
    double lg = 0.0;
    for (int k = 0; k < nm; ++k) lg += std::log2((double)modulus(k));
    int b = (int)std::floor((lg - 1.0 - 15.0 - 1e-6) / 2.0);
    return b > 53 ? 53 : b;
}

// centred residue of a 64-bit integer: in [-(p/2), (p-1)/2] ... for p = 256 the range is [-128, 127]
template <int P>
__host__ __device__ inline int centred_mod(long long x)
{
    int r = (int)(x % P);                 // C++: sign of the dividend, |r| < P
    if (r > (P - 1) / 2) r -= P;
    else if (r < -(P / 2)) r += P;
    return r;
}

// centred residue of an integer-valued double |y| <= 2^53: q = rint(y / P) in FP64 is off by at most one
// (|y / P| 2^-52 < 0.05), y - q P is a small integer and therefore exact in one fma, one correction step.
// Same value as centred_mod<P>((long long)y); ~8 FP64 / integer instructions instead of a 64-bit
// constant division (ncu: the split kernel was issue-bound at 28 instructions per residue).
template <int P>
__host__ __device__ inline int centred_mod_f64(double y)
{
    const double q = rint(y * (1.0 / (double)P));
    int r = (int)fma(-q, (double)P, y);
    if (r > (P - 1) / 2) r -= P;
    else if (r < -(P / 2)) r += P;
    return r;
}

// centred residue of an int32 accumulator modulo a run-time modulus (epilogue of the product kernel):
// q = rint(v / p) in FP64 is off by at most one for |v| < 2^31, one correction step
__host__ __device__ inline int centred_mod_rt(int v, int p, double inv_p)
{
    const int q = (int)rint((double)v * inv_p);
    int r = v - q * p;
    if (r > (p - 1) / 2) r -= p;
    else if (r < -(p / 2)) r += p;
    return r;
}

// all residues of one scaled integer
template <int K>
struct ResidueLoop {
    __host__ __device__ static inline void run(long long x, int8_t* out, int nm)
    {
        ResidueLoop<K - 1>::run(x, out, nm);
        if (K - 1 < nm) out[K - 1] = (int8_t)centred_mod<modulus(K - 1)>(x);
    }
};
template <>
struct ResidueLoop<0> {
    __host__ __device__ static inline void run(long long, int8_t*, int) {}
};
__host__ __device__ inline void residues(long long x, int8_t* out, int nm) { ResidueLoop<NM_MAX>::run(x, out, nm); }

// Garner: centred mixed-radix digits d[k] with x = d0 + p0 (d1 + p1 (d2 + ...)), then the Horner sum in
// FP64.  r[k] = x mod modulus(k), any representative with |r[k]| <= 2^15.  With the prefix products reduced
// modulo p_k in a table, digit k is one short dot product (|d_j M_kj| < 2^15, at most 15 terms) and two
// reductions by a compile-time modulus:  d_k = ((r_k - sum_j d_j M_kj) c_k) mod p_k, centred.
// All loops have compile-time bounds so that they unroll and the tables become immediates.
__host__ __device__ inline double reconstruct(const int* r, int nm)
{
    constexpr GarnerTable G = make_garner_table();
    int d[NM_MAX];
#pragma unroll
    for (int k = 0; k < NM_MAX; ++k) {
        const int p = modulus(k);
        int acc = 0;
#pragma unroll
        for (int j = 0; j < k; ++j) acc += d[j] * G.M[k][j];
        int t = ((r[k] - acc) % p) * G.c[k] % p;
        if (t > (p - 1) / 2) t -= p;
        else if (t < -(p / 2)) t += p;
        d[k] = k < nm ? t : 0;
    }
    double s = 0.0;
#pragma unroll
    for (int k = NM_MAX - 1; k >= 0; --k)
        if (k < nm) s = fma(s, (double)modulus(k), (double)d[k]);      // explicit: the same bits on host and device
    return s;
}

// ---- the two element-wise kernels of the CRT form, as per-item routines (kernels in syrk_i8.cu) ----
// Operand blocks are stored pre-tiled in the SWIZZLE_128B order the tensor-core descriptors expect:
// block (i / 128, k / 128) is 16 KiB contiguous, row r of a block at r * 128 bytes, its 16-byte chunk c
// at position c ^ (r & 7)  (same layout as the digit planes of i8_split_kernel).
constexpr int BLOCK_BYTES = 128 * 128;

// residue plane K-1 of 16 scaled integers: four packed words, one 16-byte store (planes one after the
// other, so that only the 16 integers stay live)
template <int K>
struct PlaneLoop {
    __host__ __device__ static inline void run(const double* x, int8_t* dst, long long plane_stride, int nm)
    {
        PlaneLoop<K - 1>::run(x, dst, plane_stride, nm);
        if (K - 1 < nm) {
            uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
            for (int j = 0; j < 16; ++j)
                w[j >> 2] |= ((uint32_t)centred_mod_f64<modulus(K - 1)>(x[j]) & 0xffu) << (8 * (j & 3));
            uint32_t* o = reinterpret_cast<uint32_t*>(dst + (long long)(K - 1) * plane_stride);     // 16-byte aligned
            o[0] = w[0]; o[1] = w[1]; o[2] = w[2]; o[3] = w[3];
        }
    }
};
template <>
struct PlaneLoop<0> {
    __host__ __device__ static inline void run(const double*, int8_t*, long long, int) {}
};

// item idx = (row i, group of 16 consecutive k): residue plane m of rint(A[i, k0+k] / sc[i] * 2^bits)
// modulo modulus(m); rows >= n and columns >= kc are written as zeros
__host__ __device__ inline void split_item(long long idx, const double* A, long long lda, int n, int nrows_pad,
                                           long long k0, int kc, int kcpad, const double* sc, int8_t* planes,
                                           long long rb_stride, long long plane_stride, int nm, int bits)
{
    const int groups = kcpad / 16;
    if (idx >= (long long)nrows_pad * groups) return;
    const int i = (int)(idx / groups), kg = (int)(idx - (long long)i * groups) * 16;
    double x[16];                                           // integer-valued, |x| <= 2^bits <= 2^53
    if (i < n) {
        const double mult = ldexp(1.0, bits) / sc[i];       // exact: powers of two
        const double* src = A + (long long)i * lda + k0 + kg;
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = (kg + j < kc) ? rint(src[j] * mult) : 0.0;
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = 0.0;
    }
    const int r = i & 127, c = (kg & 127) >> 4;
    int8_t* dst = planes + (long long)(i >> 7) * rb_stride + (long long)(kg >> 7) * BLOCK_BYTES + r * 128 + ((c ^ (r & 7)) << 4);
    PlaneLoop<NM_MAX>::run(x, dst, plane_stride, nm);
}

// item (row, col), col >= row: C[row, col] += f sc[row] sc[col] x, x rebuilt from R[m][row][col]
__host__ __device__ inline void reconstruct_item(int row, int col, const int8_t* R, long long ldr, long long res_stride,
                                                 int nm, int n, const double* sc, double f, double* C, long long ldc)
{
    if (col >= n || col < row) return;
    int r[NM_MAX];
    const int8_t* src = R + (long long)row * ldr + col;
#pragma unroll
    for (int m = 0; m < NM_MAX; ++m) r[m] = m < nm ? (int)src[m * res_stride] : 0;
    const double x = reconstruct(r, nm);
    C[(long long)row * ldc + col] += x * (f * sc[row] * sc[col]);
}

}  // namespace isdf_crt
