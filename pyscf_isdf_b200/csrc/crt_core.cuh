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
        if (k < nm) s = s * (double)modulus(k) + (double)d[k];
    return s;
}

}  // namespace isdf_crt
