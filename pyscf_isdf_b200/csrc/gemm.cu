// FP64 GEMM for the ISDF hot path (sm_100a).
//
// One kernel family covers every dense contraction of stages 2-5:
//   * both operand layouts (k-major rows, or k-contiguous rows),
//   * accumulator initialised from a matrix (C = Cin -/+ A^T B),
//   * fused epilogues: scale, Hadamard square ((Phi_IP^T Phi)^2 pair densities),
//     Hadamard multiply (M o W of the exchange build).
//
// Arithmetic contract: every C element is one FMA chain over ascending k
// (no split-K, no re-association), which the C oracle reproduces bit for bit.
//
// Tiling: BM x BN CTA tile, 256 threads as a 16 x 16 grid, TM x TN register
// micro-tile per thread (8x8 -> 128x128 tile, 4x4 -> 64x64 tile), BK = 16,
// double-buffered shared memory, register-staged global prefetch.
#include "common.cuh"

namespace {

constexpr int BK = 16;
constexpr int NTHREADS = 256;

template <int BX>
struct Tile {
    static constexpr int LD = BX + 2;          // 2-way-conflict-free transposed stores, 16B-aligned rows
    static constexpr int NJ = BX / 32;         // double2 loads per thread per k-tile
};

// ---- global -> registers -------------------------------------------------
// k-major operand: elem(k, x) = p[k*ld + x].  thread t: k = t/16, x = (t%16)*2 + 32*j
// k-contig operand: elem(k, x) = p[x*ld + k]. thread t: k = 2*(t%8) (+1), x = t/8 + 32*j
template <int BX, bool KMAJOR>
__device__ __forceinline__ void load_tile(const double* __restrict__ p, long long ld,
                                          int dimX, int K, int x0, int k0,
                                          double2 (&r)[Tile<BX>::NJ], bool fast, int tid)
{
    constexpr int NJ = Tile<BX>::NJ;
    if (KMAJOR) {
        const int k = k0 + (tid >> 4);
        const int xb = x0 + ((tid & 15) << 1);
        if (fast) {
            const double* q = p + (long long)k * ld + xb;
#pragma unroll
            for (int j = 0; j < NJ; ++j) r[j] = *reinterpret_cast<const double2*>(q + 32 * j);
        } else {
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int x = xb + 32 * j;
                double2 v = make_double2(0.0, 0.0);
                if (k < K) {
                    const double* q = p + (long long)k * ld + x;
                    if (x < dimX) v.x = q[0];
                    if (x + 1 < dimX) v.y = q[1];
                }
                r[j] = v;
            }
        }
    } else {
        const int k = k0 + ((tid & 7) << 1);
        const int xb = x0 + (tid >> 3);
        if (fast) {
#pragma unroll
            for (int j = 0; j < NJ; ++j)
                r[j] = *reinterpret_cast<const double2*>(p + (long long)(xb + 32 * j) * ld + k);
        } else {
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const int x = xb + 32 * j;
                double2 v = make_double2(0.0, 0.0);
                if (x < dimX) {
                    const double* q = p + (long long)x * ld + k;
                    if (k < K) v.x = q[0];
                    if (k + 1 < K) v.y = q[1];
                }
                r[j] = v;
            }
        }
    }
}

// ---- registers -> shared (always stored as S[k][x]) ------------------------
template <int BX, bool KMAJOR, bool NEG>
__device__ __forceinline__ void store_tile(double* __restrict__ S,
                                           const double2 (&r)[Tile<BX>::NJ], int tid)
{
    constexpr int NJ = Tile<BX>::NJ;
    constexpr int LD = Tile<BX>::LD;
    if (KMAJOR) {
        double* s = S + (tid >> 4) * LD + ((tid & 15) << 1);
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            double2 v = r[j];
            if (NEG) { v.x = -v.x; v.y = -v.y; }
            *reinterpret_cast<double2*>(s + 32 * j) = v;
        }
    } else {
        double* s = S + ((tid & 7) << 1) * LD + (tid >> 3);
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            double2 v = r[j];
            if (NEG) { v.x = -v.x; v.y = -v.y; }
            s[32 * j] = v.x;
            s[LD + 32 * j] = v.y;
        }
    }
}

template <int TM, int TN, bool AK, bool BKM, bool NEG>
__global__ void __launch_bounds__(NTHREADS, 1)
dgemm_kernel(GemmParams p)
{
    constexpr int BM = 16 * TM, BN = 16 * TN;
    constexpr int LDA = Tile<BM>::LD, LDB = Tile<BN>::LD;
    extern __shared__ __align__(16) double smem[];
    double* As = smem;                       // [2][BK][LDA]
    double* Bs = smem + 2 * BK * LDA;        // [2][BK][LDB]

    const int tid = threadIdx.x;
    const int m0 = blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;
    if (p.upper_only && m0 >= n0 + BN) return;
    const int M = p.M, N = p.N, K = p.K;

    const int tx = tid & 15, ty = tid >> 4;

    double acc[TM][TN];
    // accumulator init (the FMA chain starts here)
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
            double v = 0.0;
            if (p.Cin != nullptr && m < M && n < N) v = p.Cin[(long long)m * p.ldcin + n];
            acc[i][j] = v;
        }
    }

    const bool alignedA = ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) && ((p.lda & 1) == 0);
    const bool alignedB = ((reinterpret_cast<uintptr_t>(p.B) & 15) == 0) && ((p.ldb & 1) == 0);
    const bool fullM = alignedA && (m0 + BM <= M);
    const bool fullN = alignedB && (n0 + BN <= N);

    double2 ra[Tile<BM>::NJ], rb[Tile<BN>::NJ];
    const int nk = (K + BK - 1) / BK;
    if (nk > 0) {
        const bool fk = (BK <= K);
        load_tile<BM, AK>(p.A, p.lda, M, K, m0, 0, ra, fullM && fk, tid);
        load_tile<BN, BKM>(p.B, p.ldb, N, K, n0, 0, rb, fullN && fk, tid);
        store_tile<BM, AK, NEG>(As, ra, tid);
        store_tile<BN, BKM, false>(Bs, rb, tid);
    }
    __syncthreads();

    for (int kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            const int k0 = (kt + 1) * BK;
            const bool fk = (k0 + BK <= K);
            load_tile<BM, AK>(p.A, p.lda, M, K, m0, k0, ra, fullM && fk, tid);
            load_tile<BN, BKM>(p.B, p.ldb, N, K, n0, k0, rb, fullN && fk, tid);
        }
        const double* as = As + cur * BK * LDA + ty * 4;
        const double* bs = Bs + cur * BK * LDB + tx * 4;
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            double a[TM], b[TN];
#pragma unroll
            for (int g = 0; g < TM / 4; ++g) {
                const double2 v0 = *reinterpret_cast<const double2*>(as + k * LDA + g * 64);
                const double2 v1 = *reinterpret_cast<const double2*>(as + k * LDA + g * 64 + 2);
                a[g * 4 + 0] = v0.x; a[g * 4 + 1] = v0.y; a[g * 4 + 2] = v1.x; a[g * 4 + 3] = v1.y;
            }
#pragma unroll
            for (int g = 0; g < TN / 4; ++g) {
                const double2 v0 = *reinterpret_cast<const double2*>(bs + k * LDB + g * 64);
                const double2 v1 = *reinterpret_cast<const double2*>(bs + k * LDB + g * 64 + 2);
                b[g * 4 + 0] = v0.x; b[g * 4 + 1] = v0.y; b[g * 4 + 2] = v1.x; b[g * 4 + 3] = v1.y;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j)
                    acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            store_tile<BM, AK, NEG>(As + (cur ^ 1) * BK * LDA, ra, tid);
            store_tile<BN, BKM, false>(Bs + (cur ^ 1) * BK * LDB, rb, tid);
        }
        __syncthreads();
    }

    // epilogue
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
        if (m >= M) continue;
#pragma unroll
        for (int jg = 0; jg < TN / 4; ++jg) {
            const int n = n0 + jg * 64 + tx * 4;
            double v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                double x = acc[i][jg * 4 + q];
                if (p.epi == EPI_SQUARE) x = x * x;
                else if (p.epi == EPI_MUL) {
                    if (n + q < N) x = x * p.E[(long long)m * p.lde + n + q];
                } else x = p.alpha * x;
                v[q] = x;
            }
            double* c = p.C + (long long)m * p.ldc + n;
            if (n + 3 < N && ((reinterpret_cast<uintptr_t>(c) & 15) == 0)) {
                *reinterpret_cast<double2*>(c) = make_double2(v[0], v[1]);
                *reinterpret_cast<double2*>(c + 2) = make_double2(v[2], v[3]);
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) if (n + q < N) c[q] = v[q];
            }
        }
    }
}

template <int TM, int TN, bool AK, bool BKM, bool NEG>
int launch_cfg(isdf_handle* h, const GemmParams& p)
{
    constexpr int BM = 16 * TM, BN = 16 * TN;
    const size_t smem = sizeof(double) * 2 * BK * (Tile<BM>::LD + Tile<BN>::LD);
    auto kern = dgemm_kernel<TM, TN, AK, BKM, NEG>;
    static bool configured = false;
    if (!configured) {
        CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    dim3 grid((p.M + BM - 1) / BM, (p.N + BN - 1) / BN);
    const bool prof = h->prof_on && h->prof_n < h->prof_cap;
    if (prof) CUDA_TRY(cudaEventRecord(h->prof_ev[2 * h->prof_n], h->stream));
    kern<<<grid, NTHREADS, smem, h->stream>>>(p);
    CUDA_TRY(cudaGetLastError());
    if (prof) {
        CUDA_TRY(cudaEventRecord(h->prof_ev[2 * h->prof_n + 1], h->stream));
        h->prof_n++;
        // algorithmic flops of this launch: 2*K per computed output element
        double elems = (double)p.M * (double)p.N;
        if (p.upper_only) {
            long long tiles = 0;
            for (unsigned bx = 0; bx < grid.x; ++bx)
                for (unsigned by = 0; by < grid.y; ++by)
                    if (!((long long)bx * BM >= (long long)by * BN + BN)) tiles++;
            elems = (double)tiles * BM * BN;
        }
        h->prof_flops += 2.0 * elems * (double)p.K;
    }
    h->gemm_launches++;
    return ISDF_OK;
}

template <int TM, int TN>
int dispatch_layout(isdf_handle* h, const GemmParams& p)
{
    const int key = (p.a_kmajor ? 4 : 0) | (p.b_kmajor ? 2 : 0) | (p.neg ? 1 : 0);
    switch (key) {
    case 0: return launch_cfg<TM, TN, false, false, false>(h, p);
    case 1: return launch_cfg<TM, TN, false, false, true>(h, p);
    case 2: return launch_cfg<TM, TN, false, true, false>(h, p);
    case 3: return launch_cfg<TM, TN, false, true, true>(h, p);
    case 4: return launch_cfg<TM, TN, true, false, false>(h, p);
    case 5: return launch_cfg<TM, TN, true, false, true>(h, p);
    case 6: return launch_cfg<TM, TN, true, true, false>(h, p);
    default: return launch_cfg<TM, TN, true, true, true>(h, p);
    }
}

}  // namespace

int gemm_launch(isdf_handle* h, const GemmParams& p)
{
    if (p.M <= 0 || p.N <= 0) return ISDF_OK;
    ARG_CHECK(p.A && p.B && p.C && p.K >= 0);
    ARG_CHECK(p.epi != EPI_MUL || p.E != nullptr);
    // 128x128 tiles when they fill the machine, else 64x64
    const long long big_tiles = (long long)((p.M + 127) / 128) * ((p.N + 127) / 128);
    if (big_tiles >= h->num_sms) return dispatch_layout<8, 8>(h, p);
    return dispatch_layout<4, 4>(h, p);
}
