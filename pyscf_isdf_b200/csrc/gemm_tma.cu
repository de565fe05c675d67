// FP64 GEMM, warp-specialised form: operand tiles arrive by 2-D tensor-map TMA (cp.async.bulk.tensor.2d, SASS
// UTMALDG) from ONE producer lane into an mbarrier ring; the DMMA consumer warps never meet a CTA-wide barrier
// in the K loop (they release a stage with fence.proxy.async + mbarrier.arrive, see the loop).  Same arithmetic contract as gemm.cu (every C element is one ascending-k FMA chain from Cin,
// K tail and ragged edges zero-filled -- here by the TMA unit's out-of-bounds fill), so results are
// bit-identical to dgemm_kernel; tests/test_g1_stages_gpu.py compares the two forms bit for bit.
//
// Shared-memory image of one operand tile (BX x 16 k), SWIZZLE_128B (the 16-byte chunk index of every
// 128-byte row is XORed with the row index mod 8):
//   k-major operand  (memory rows = k, x contiguous): BX/16 boxes of {16 x, 16 k}: [x >> 4][k][16 x]
//   k-contiguous operand (memory rows = x, k contiguous): one box of {16 k, BX x}:  [x][16 k]
// A TMA box is dense, so the padding / XOR-on-k layout of gemm.cu is not available; instead the rows of
// the 8x8x4 DMMA tiles are a permutation of the tile's x indices chosen so that the fragment loads of a
// (half-)warp hit distinct banks under the hardware swizzle:
//   tile i of a 16-wide x group holds x = 2 S(g) + (i & 1), g = lane / 4,
//   S(g) = (g >> 1) + 4 (g & 1) for a k-major operand (lane's two tiles = one 16-byte chunk: LDS.128),
//   S(g) = g                    for a k-contiguous one (LDS.64).
// The epilogue applies the same permutation to C's rows (A's x) and columns (B's x).
#include "common.cuh"
#include <cuda.h>
#include <cstdlib>

namespace {

constexpr int BK = 16;
constexpr long long WAIT_LIMIT = 4000000000LL;      // ~2 s of SM clocks

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1])
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    long long t0 = 0;
    for (int spin = 0;; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) break;
        if (spin == 64) t0 = clock64();
        if (spin > 64 && clock64() - t0 > WAIT_LIMIT) __trap();      // never hang the device on a protocol bug
    }
}
// one box of the tensor map -> shared memory, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
                 "[%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

template <bool KMAJOR>
__device__ __forceinline__ int perm(int g) { return KMAJOR ? (g >> 1) + 4 * (g & 1) : g; }

// Fragment loads of one k-step (4 k values) for four DMMA tiles = two 16-wide x groups starting at the warp's
// x offset `wx` (a multiple of 32).  v[2 grp + ii] belongs to tile i = 2 grp + ii.
template <bool KMAJOR>
__device__ __forceinline__ void load_frag(const double* __restrict__ S, int wx, int kk, int g, int t, double (&v)[4])
{
    if (KMAJOR) {
        const int k = kk + t;
        const int sw = (perm<true>(g) ^ (k & 7)) << 1;
#pragma unroll
        for (int grp = 0; grp < 2; ++grp) {
            const double2 x = *reinterpret_cast<const double2*>(S + ((wx >> 4) + grp) * 256 + k * 16 + sw);
            v[2 * grp] = x.x; v[2 * grp + 1] = x.y;
        }
    } else {
        const int k = kk + t;
#pragma unroll
        for (int grp = 0; grp < 2; ++grp)
#pragma unroll
            for (int ii = 0; ii < 2; ++ii) {
                const int x = wx + 16 * grp + 2 * g + ii;
                v[2 * grp + ii] = S[x * 16 + ((((k >> 1) ^ (x & 7)) << 1) | (k & 1))];
            }
    }
}

template <int WM, int WN, int NSTAGE, bool AK, bool BKM, bool NEG>
__global__ void __launch_bounds__(32 * (WM * WN + 1), (WM * WN <= 4 ? 3 : (WM * WN <= 8 ? 2 : 1)))
dgemm_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, GemmParams p)
{
    constexpr int BM = 32 * WM, BN = 32 * WN, NCW = WM * WN;
    constexpr int A_SZ = BM * BK, B_SZ = BN * BK, STAGE = A_SZ + B_SZ;      // doubles
    extern __shared__ unsigned char smem_raw[];

    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    if (p.upper_only && m0 >= n0 + BN) return;
    const int M = p.M, N = p.N, K = p.K;
    const int nk = (K + BK - 1) / BK;

    // tiles at the first 1024-byte boundary (SWIZZLE_128B), the barriers behind them
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t tiles_u32 = (raw + 1023u) & ~1023u;
    double* tiles = reinterpret_cast<double*>(smem_raw + (tiles_u32 - raw));
    const uint32_t bar_full = tiles_u32 + NSTAGE * STAGE * 8, bar_empty = bar_full + NSTAGE * 8;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, NCW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp == NCW) {
        // ---- producer: one lane feeds the ring -----------------------------------------------------------
        if (lane == 0) {
            for (int kt = 0; kt < nk; ++kt) {
                const int s = kt % NSTAGE;
                if (kt >= NSTAGE) mbar_wait(bar_empty + 8 * s, ((kt / NSTAGE) - 1) & 1);
                const uint32_t full = bar_full + 8 * s;
                mbar_expect_tx(full, STAGE * 8);
                const uint32_t dA = tiles_u32 + s * STAGE * 8, dB = dA + A_SZ * 8;
                if (AK) {
#pragma unroll
                    for (int b = 0; b < BM / 16; ++b) tma_load_2d(dA + b * 2048, &mapA, m0 + 16 * b, kt * BK, full);
                } else {
                    tma_load_2d(dA, &mapA, kt * BK, m0, full);
                }
                if (BKM) {
#pragma unroll
                    for (int b = 0; b < BN / 16; ++b) tma_load_2d(dB + b * 2048, &mapB, n0 + 16 * b, kt * BK, full);
                } else {
                    tma_load_2d(dB, &mapB, kt * BK, n0, full);
                }
            }
        }
        return;
    }

    // ---- consumers: 32x32 warp tiles ------------------------------------------------------------------------
    const int g = lane >> 2, t = lane & 3;
    const int wm = (warp / WN) * 32, wn = (warp % WN) * 32;
    // C coordinates of this lane's accumulators: acc[i][j][e], rows from A's permutation, columns from B's
    auto row_of = [&](int i) { return m0 + wm + 16 * (i >> 1) + 2 * perm<AK>(g) + (i & 1); };
    auto col_of = [&](int jp, int e) { return n0 + wn + 16 * jp + 2 * perm<BKM>(2 * t + e); };   // tiles 2jp, 2jp+1: +0, +1

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = row_of(i);
#pragma unroll
        for (int jp = 0; jp < 2; ++jp)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = col_of(jp, e);
                double v0 = 0.0, v1 = 0.0;
                if (p.Cin != nullptr && m < M) {
                    const double* c = p.Cin + (long long)m * p.ldcin + n;
                    if (n < N) v0 = c[0];
                    if (n + 1 < N) v1 = c[1];
                }
                acc[i][2 * jp][e] = v0; acc[i][2 * jp + 1][e] = v1;
            }
    }

    // Fragments are single-buffered (64 accumulator + 16 fragment registers); the LDS latency of one warp is
    // covered by the other warps of its sub-core, which own the DMMA pipe meanwhile -- no CTA-wide barrier ever
    // lines them up.
    //
    // Handing a stage back: the tiles are written by the async proxy (TMA) and read by the generic proxy (LDS).
    // The mbarrier alone orders the two only in the RAW direction (the completion of a bulk copy carries an
    // implicit proxy fence); for WAR -- our reads, then the next TMA write into the same stage -- every lane
    // issues fence.proxy.async after its last read and before the warp's arrive on the empty barrier.  Without
    // it co-resident CTAs (>= 2 per SM) read stale rows of the previous tile of a stage: 199 of 200 products
    // of a flush shape differed from the LDGSTS kernel; with it 0 of 1800 (tools/gpu_tma_variants.sh,
    // profiles/r02_gemm_tma.md).  A fence on the acquire side does not help, nor does arriving after the
    // DMMAs that consumed the fragments.
    for (int kt = 0; kt < nk; ++kt) {
        const int s = kt % NSTAGE;
        const double* As = tiles + s * STAGE;
        mbar_wait(bar_full + 8 * s, (kt / NSTAGE) & 1);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            double a[4], b[4];
            load_frag<AK>(As, wm, 4 * ks, g, t, a);
            load_frag<BKM>(As + A_SZ, wn, 4 * ks, g, t, b);
            if (ks == 3) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_empty + 8 * s);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j], NEG ? -a[i] : a[i], b[j]);
        }
    }

    // ---- epilogue -------------------------------------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = row_of(i);
        if (m >= M) continue;
#pragma unroll
        for (int jp = 0; jp < 2; ++jp)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int n = col_of(jp, e);
                if (n >= N) continue;
                double v0 = acc[i][2 * jp][e], v1 = acc[i][2 * jp + 1][e];
                if (p.epi == EPI_SQUARE) { v0 = v0 * v0; v1 = v1 * v1; }
                else if (p.epi == EPI_MUL) {
                    const double* ep = p.E + (long long)m * p.lde + n;
                    v0 = v0 * ep[0];
                    if (n + 1 < N) v1 = v1 * ep[1];
                } else { v0 = p.alpha * v0; v1 = p.alpha * v1; }
                double* c = p.C + (long long)m * p.ldc + n;
                if (n + 1 < N && ((reinterpret_cast<uintptr_t>(c) & 15) == 0)) {
                    *reinterpret_cast<double2*>(c) = make_double2(v0, v1);
                } else {
                    c[0] = v0;
                    if (n + 1 < N) c[1] = v1;
                }
            }
    }
}

// ---- host: tensor maps + launch --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    // the driver entry point, looked up once (the library does not link libcuda)
    static EncodeTiledFn fn = [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            ptr = nullptr;
        return (EncodeTiledFn)ptr;
    }();
    return fn;
}

// operand [rows of k | rows of x] -> 2-D map with 128-byte inner boxes, zero fill outside (K tail, ragged x edge)
int make_map(CUtensorMap* map, const double* ptr, long long ld, int dimX, int K, bool kmajor, int bx)
{
    // returns 1 (not an error code) when the map cannot be made: the caller then takes the LDGSTS kernel
    EncodeTiledFn enc = encode_fn();
    if (!enc) return 1;
    cuuint64_t dims[2] = {(cuuint64_t)(kmajor ? dimX : K), (cuuint64_t)(kmajor ? K : dimX)};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
    cuuint32_t box[2] = {16, (cuuint32_t)(kmajor ? BK : bx)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)ptr, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? ISDF_OK : 1;
}

template <int WM, int WN, int NSTAGE, bool AK, bool BKM, bool NEG>
int launch_tma_cfg(isdf_handle* h, const GemmParams& p)
{
    constexpr int BM = 32 * WM, BN = 32 * WN, NT = 32 * (WM * WN + 1);
    static_assert(BM <= 256 && BN <= 256, "one TMA box per k-contiguous operand tile");
    const size_t smem = (size_t)NSTAGE * (BM + BN) * BK * 8 + 1024 + 2 * NSTAGE * 8;
    auto kern = dgemm_tma_kernel<WM, WN, NSTAGE, AK, BKM, NEG>;
    { int rc_; if ((rc_ = isdf_func_max_smem((const void*)kern, smem))) return rc_; }
    CUtensorMap mapA, mapB;
    { int rc_; if ((rc_ = make_map(&mapA, p.A, p.lda, p.M, p.K, AK, BM))) return rc_; }
    { int rc_; if ((rc_ = make_map(&mapB, p.B, p.ldb, p.N, p.K, BKM, BN))) return rc_; }
    dim3 grid((p.M + BM - 1) / BM, (p.N + BN - 1) / BN);
    const bool prof = h->prof_on && h->prof_n < h->prof_cap;
    if (prof) CUDA_TRY(cudaEventRecord(h->prof_ev[2 * h->prof_n], h->stream));
    kern<<<grid, NT, smem, h->stream>>>(mapA, mapB, p);
    CUDA_TRY(cudaGetLastError());
    if (prof) {
        CUDA_TRY(cudaEventRecord(h->prof_ev[2 * h->prof_n + 1], h->stream));
        h->prof_n++;
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

template <int WM, int WN, int NSTAGE>
int dispatch_tma(isdf_handle* h, const GemmParams& p)
{
    const int key = (p.a_kmajor ? 4 : 0) | (p.b_kmajor ? 2 : 0) | (p.neg ? 1 : 0);
    switch (key) {
    case 0: return launch_tma_cfg<WM, WN, NSTAGE, false, false, false>(h, p);
    case 1: return launch_tma_cfg<WM, WN, NSTAGE, false, false, true>(h, p);
    case 2: return launch_tma_cfg<WM, WN, NSTAGE, false, true, false>(h, p);
    case 3: return launch_tma_cfg<WM, WN, NSTAGE, false, true, true>(h, p);
    case 4: return launch_tma_cfg<WM, WN, NSTAGE, true, false, false>(h, p);
    case 5: return launch_tma_cfg<WM, WN, NSTAGE, true, false, true>(h, p);
    case 6: return launch_tma_cfg<WM, WN, NSTAGE, true, true, false>(h, p);
    default: return launch_tma_cfg<WM, WN, NSTAGE, true, true, true>(h, p);
    }
}

}  // namespace

// TMA needs 16-byte aligned bases and row strides; everything else (ragged edges, K tail) is the unit's zero fill
bool gemm_tma_applicable(const GemmParams& p)
{
    return p.K > 0 && ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) && ((p.lda & 1) == 0) &&
           ((reinterpret_cast<uintptr_t>(p.B) & 15) == 0) && ((p.ldb & 1) == 0) &&
           p.lda * 8 < (1LL << 40) && p.ldb * 8 < (1LL << 40);
}

// returns 1 when a tensor map could not be encoded (nothing launched): gemm_launch falls back to dgemm_kernel.
// form: 7 = 64x128 CTA tile (2x4 consumer warps, 2 CTAs per SM), 8 = 128x64 (4x2), 9 = 128x128 (4x4, one CTA per SM),
// 10 = 64x64 (2x2, 3 CTAs per SM), 11 = 32x128 (1x4, 3 stages, 3 CTAs per SM: products with M <= 32, e.g. the
// row blocks of the panel sweep)
int gemm_tma_launch(isdf_handle* h, const GemmParams& p, int form)
{
    switch (form) {
    case 8: return dispatch_tma<4, 2, 4>(h, p);
    case 9: return dispatch_tma<4, 4, 4>(h, p);
    case 10: return dispatch_tma<2, 2, 4>(h, p);
    case 11: return dispatch_tma<1, 4, 3>(h, p);
    default: return dispatch_tma<2, 4, 4>(h, p);
    }
}
