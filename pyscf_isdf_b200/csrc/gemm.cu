// FP64 GEMM for the ISDF hot path on the B200 FP64 tensor pipe (sm_100a).
//
// One kernel family covers every dense contraction of stages 2-5:
//   * both operand layouts (k-major rows, or k-contiguous rows),
//   * accumulator initialised from a matrix (C = Cin -/+ A^T B),
//   * fused epilogues: scale, Hadamard square ((Phi_IP^T Phi)^2 pair densities),
//     Hadamard multiply (M o W of the exchange build).
//
// Inner product: mma.sync.aligned.m8n8k4.f64 (SASS DMMA).  Measured on B200
// (tools/fp64_probe.cu, profiles/r01_fp64_probe.txt): the instruction
// accumulates c + a0 b0 + a1 b1 + a2 b2 + a3 b3 as ONE ascending-k FMA chain
// (bit-identical to fma() over 1.28 M random cases), and its peak equals the
// DFMA peak (37.0 vs 36.7 TFLOP/s).  So the arithmetic contract of this
// library -- every C element is one FMA chain over ascending k starting from
// Cin, no split-K, no re-association, which the C oracle reproduces bit for
// bit -- holds on the tensor pipe, with 1/32 of the issue slots of DFMA.
//
// Operand feed: a 3-stage shared-memory ring filled with 16-byte cp.async
// (SASS LDGSTS) copies whose src-size operand zero-fills everything out of
// range (ragged M/N edges and the K tail); padded row strides (== 4 mod 16
// doubles) make the DMMA fragment loads bank-conflict free (ncu: 0 conflicts).
// Operands that are not 16-byte aligned go through a synchronous loader.
// A TMA variant (per-row cp.async.bulk + mbarrier complete_tx, SASS UBLKCP)
// was built and measured first (profiles/r01_gemm_loaders.md): equal speed
// for k-major operands (31.4 vs 30.9 TFLOP/s on the 128x128 tile) but 2-4x
// slower for k-contiguous ones (a 16-double row is a 128-byte bulk copy, 256
// of them per stage), and ncu shows the kernel is DMMA-pipe bound either way
// (tensor pipe 84 % active, no operand-feed stalls) -- so LDGSTS is the loader.
//
// Tiling: warp tile (8 MI) x 32 with MI = 4 (32x32, 32 FP64 accumulators per lane; default) or
// MI = 8 (64x32); default CTA tile 64x64 = four 32x32 warps, 4 CTAs per SM (128 registers).
// A sub-core retires one DMMA per 16 clocks, so small grids are bound by the K loop of one
// warp: the 32x32 warp tile halves that path (profiles/r01_gemm_loaders.md, last table).
#include "common.cuh"
#include <cstdlib>

namespace {

constexpr int BK = 16;
constexpr int WTN = 32;                    // warp tile: (8 MI) x 32, MI = 8 (throughput) or 4 (latency)
constexpr int MINB = 3;                    // CTAs per SM the 64x128 tile is built for (registers <= 168)

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1])
                 : "d"(a), "d"(b));
}

// Shared-memory image of one operand tile (BX rows/cols x BK), unpadded, XOR-swizzled so that
// the DMMA fragment loads (lane = 4*g + t reads element (x = x0 + g, k = k0 + t)) hit 16
// distinct 8-byte words per half-warp:
//   k-major     S[k][x ^ ((k & 3) << 2)]      row stride BX
//   k-contiguous S[x][k ^ ((x & 3) << 2)]      row stride BK
// The swizzle permutes 4-double groups only, so 16-byte cp.async chunks stay intact.
// No padding: 3 stages of a 64x128 tile are 72 KiB and three CTAs fit one SM.
template <int BX, bool KMAJOR>
struct OpTile {
    static constexpr int LD = KMAJOR ? BX : BK;
    static constexpr int ROWS = KMAJOR ? BK : BX;
    static constexpr int SIZE = ROWS * LD;
    // offset of element (x, k)
    __device__ static __forceinline__ int at(int x, int k)
    {
        return KMAJOR ? k * BX + (x ^ ((k & 3) << 2)) : x * BK + (k ^ ((x & 3) << 2));
    }
};

// ---- synchronous zero-filling loader (operands that are not 16-byte aligned) -------------------------
template <int BX, bool KMAJOR, int NT>
__device__ __forceinline__ void load_slow(double* S, const double* __restrict__ g, long long ld,
                                          int dimX, int K, int x0, int k0, int tid)
{
    using T = OpTile<BX, KMAJOR>;
    for (int idx = tid; idx < BX * BK; idx += NT) {
        int x, k;
        if (KMAJOR) { k = idx / BX; x = idx - k * BX; }
        else { x = idx / BK; k = idx - x * BK; }
        double v = 0.0;
        if (x0 + x < dimX && k0 + k < K)
            v = KMAJOR ? g[(long long)(k0 + k) * ld + x0 + x] : g[(long long)(x0 + x) * ld + k0 + k];
        S[T::at(x, k)] = v;
    }
}

// ---- LDGSTS producer (cp.async, 16-byte chunks, zero-fill of the out-of-range part) --------
__device__ __forceinline__ void cp_async16(void* dst, const void* src, int src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

template <int BX, bool KMAJOR, int NT>
__device__ __forceinline__ void load_async(double* S, const double* __restrict__ g, long long ld,
                                           int dimX, int K, int x0, int k0, int tid)
{
    using T = OpTile<BX, KMAJOR>;
    if (KMAJOR) {
        constexpr int CPR = BX / 2;                     // 16-byte chunks per k row
#pragma unroll
        for (int idx = tid; idx < BK * CPR; idx += NT) {
            const int r = idx / CPR, c = (idx - r * CPR) * 2;
            const int x = x0 + c, k = k0 + r;
            int bytes = 0;
            if (k < K) bytes = (x + 1 < dimX) ? 16 : (x < dimX ? 8 : 0);
            const double* src = bytes ? g + (long long)k * ld + x : g;
            cp_async16(S + T::at(c, r), src, bytes);
        }
    } else {
        constexpr int CPR = BK / 2;                     // 16-byte chunks per x row
#pragma unroll
        for (int idx = tid; idx < BX * CPR; idx += NT) {
            const int r = idx / CPR, c = (idx - r * CPR) * 2;
            const int x = x0 + r, k = k0 + c;
            int bytes = 0;
            if (x < dimX) bytes = (k + 1 < K) ? 16 : (k < K ? 8 : 0);
            const double* src = bytes ? g + (long long)x * ld + k : g;
            cp_async16(S + T::at(r, c), src, bytes);
        }
    }
}

// ---- one k-tile of DMMAs on a resident stage ------------------------------------------------
template <typename TA, typename TB, bool NEG, int MI>
__device__ __forceinline__ void compute_stage(const double* __restrict__ As, const double* __restrict__ Bs,
                                              double (&acc)[MI][4][2], int wm, int wn, int g, int t)
{
#pragma unroll
    for (int kk = 0; kk < BK; kk += 4) {
        double a[MI], b[4];
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            const double v = As[TA::at(wm + 8 * i + g, kk + t)];
            a[i] = NEG ? -v : v;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            b[j] = Bs[TB::at(wn + 8 * j + g, kk + t)];
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j], a[i], b[j]);
    }
}

template <int WM, int WN, int NSTAGE, bool AK, bool BKM, bool NEG, int MI>
__global__ void __launch_bounds__(32 * WM * WN, (MI == 8 && WM * WN == 4) ? MINB : (WM * WN <= 4 ? 4 : (WM * WN <= 8 ? 2 : 1)))
dgemm_kernel(GemmParams p)
{
    constexpr int WTM = 8 * MI;
    constexpr int BM = WTM * WM, BN = WTN * WN, NT = 32 * WM * WN;
    using TA = OpTile<BM, AK>;
    using TB = OpTile<BN, BKM>;
    constexpr int STAGE = TA::SIZE + TB::SIZE;
    extern __shared__ __align__(128) double smem[];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wm = (warp / WN) * WTM, wn = (warp % WN) * WTN;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    if (p.upper_only && m0 >= n0 + BN) return;
    const int M = p.M, N = p.N, K = p.K;

    // fast path needs 16-byte aligned rows in both operands
    const bool fast = ((reinterpret_cast<uintptr_t>(p.A) & 15) == 0) && ((p.lda & 1) == 0) &&
                      ((reinterpret_cast<uintptr_t>(p.B) & 15) == 0) && ((p.ldb & 1) == 0);
    const int nfull = fast ? (K + BK - 1) / BK : 0;
    auto issue = [&](int kt) {
        double* As = smem + (kt % NSTAGE) * STAGE;
        load_async<BM, AK, NT>(As, p.A, p.lda, M, K, m0, kt * BK, tid);
        load_async<BN, BKM, NT>(As + TA::SIZE, p.B, p.ldb, N, K, n0, kt * BK, tid);
    };
    for (int kt = 0; kt < NSTAGE - 1; ++kt) {
        if (kt < nfull) issue(kt);
        cp_async_commit();
    }

    // accumulator init (the FMA chain starts here)
    double acc[MI][4][2];
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const int m = m0 + wm + 8 * i + g;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + wn + 8 * j + 2 * t;
            double v0 = 0.0, v1 = 0.0;
            if (p.Cin != nullptr && m < M) {
                const double* c = p.Cin + (long long)m * p.ldcin + n;
                if (n < N) v0 = c[0];
                if (n + 1 < N) v1 = c[1];
            }
            acc[i][j][0] = v0; acc[i][j][1] = v1;
        }
    }

    for (int kt = 0; kt < nfull; ++kt) {
        const int s = kt % NSTAGE;
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();                       // tile kt landed; every warp is done with the stage refilled below
        if (kt + NSTAGE - 1 < nfull) issue(kt + NSTAGE - 1);
        cp_async_commit();
        const double* As = smem + s * STAGE;
        compute_stage<TA, TB, NEG, MI>(As, As + TA::SIZE, acc, wm, wn, g, t);
    }
    // unaligned operands: synchronous zero-filling loads
    for (int k0 = nfull * BK; k0 < K; k0 += BK) {
        __syncthreads();
        double* As = smem;
        double* Bs = As + TA::SIZE;
        load_slow<BM, AK, NT>(As, p.A, p.lda, M, K, m0, k0, tid);
        load_slow<BN, BKM, NT>(Bs, p.B, p.ldb, N, K, n0, k0, tid);
        __syncthreads();
        compute_stage<TA, TB, NEG, MI>(As, Bs, acc, wm, wn, g, t);
    }

    // epilogue: lane holds C[m][n], C[m][n+1] of every 8x8 tile
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        const int m = m0 + wm + 8 * i + g;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + wn + 8 * j + 2 * t;
            if (n >= N) continue;
            double v0 = acc[i][j][0], v1 = acc[i][j][1];
            if (p.epi == EPI_SQUARE) { v0 = v0 * v0; v1 = v1 * v1; }
            else if (p.epi == EPI_MUL) {
                const double* e = p.E + (long long)m * p.lde + n;
                v0 = v0 * e[0];
                if (n + 1 < N) v1 = v1 * e[1];
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

template <int WM, int WN, int NSTAGE, bool AK, bool BKM, bool NEG, int MI>
int launch_cfg(isdf_handle* h, const GemmParams& p)
{
    constexpr int BM = 8 * MI * WM, BN = WTN * WN, NT = 32 * WM * WN;
    const size_t smem = sizeof(double) * NSTAGE * (OpTile<BM, AK>::SIZE + OpTile<BN, BKM>::SIZE);
    auto kern = dgemm_kernel<WM, WN, NSTAGE, AK, BKM, NEG, MI>;
    { int rc_; if ((rc_ = isdf_func_max_smem((const void*)kern, smem))) return rc_; }
    dim3 grid((p.M + BM - 1) / BM, (p.N + BN - 1) / BN);
    const bool prof = h->prof_on && h->prof_n < h->prof_cap;
    if (prof) CUDA_TRY(cudaEventRecord(h->prof_ev[2 * h->prof_n], h->stream));
    kern<<<grid, NT, smem, h->stream>>>(p);
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

template <int WM, int WN, int NSTAGE, int MI>
int dispatch_layout(isdf_handle* h, const GemmParams& p)
{
    const int key = (p.a_kmajor ? 4 : 0) | (p.b_kmajor ? 2 : 0) | (p.neg ? 1 : 0);
    switch (key) {
    case 0: return launch_cfg<WM, WN, NSTAGE, false, false, false, MI>(h, p);
    case 1: return launch_cfg<WM, WN, NSTAGE, false, false, true, MI>(h, p);
    case 2: return launch_cfg<WM, WN, NSTAGE, false, true, false, MI>(h, p);
    case 3: return launch_cfg<WM, WN, NSTAGE, false, true, true, MI>(h, p);
    case 4: return launch_cfg<WM, WN, NSTAGE, true, false, false, MI>(h, p);
    case 5: return launch_cfg<WM, WN, NSTAGE, true, false, true, MI>(h, p);
    case 6: return launch_cfg<WM, WN, NSTAGE, true, true, false, MI>(h, p);
    default: return launch_cfg<WM, WN, NSTAGE, true, true, true, MI>(h, p);
    }
}

}  // namespace

// ---- debugging aid (ISDF_GEMM_VERIFY=1): every TMA-form product is repeated by the LDGSTS kernel into a
// scratch copy of C and compared bit for bit; the first differing call is reported with its parameters.
// Allocates and synchronises -- never on in a measured or shipped run.
__global__ void gemm_verify_kernel(const double* __restrict__ C, long long ldc, const double* __restrict__ R, long long ldr,
                                   int M, int N, int upper, unsigned long long* out)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)M * N) return;
    const int m = (int)(idx / N), n = (int)(idx - (long long)m * N);
    if (upper && n < m) return;
    const unsigned long long a = __double_as_longlong(C[(long long)m * ldc + n]);
    const unsigned long long b = __double_as_longlong(R[(long long)m * ldr + n]);
    if (a != b) {
        atomicAdd(out, 1ULL);
        atomicMin(out + 1, (unsigned long long)idx);
    }
}

static bool verify_on()
{
    static const bool on = [] { const char* e = getenv("ISDF_GEMM_VERIFY"); return e && e[0] == '1'; }();
    return on;
}

static int gemm_verified(isdf_handle* h, const GemmParams& p)
{
    static int call = 0;
    const int me = call++;
    double* R = nullptr;
    unsigned long long* out = nullptr;
    const size_t bytes = (size_t)p.M * p.ldc * sizeof(double);
    CUDA_TRY(cudaMalloc(&R, bytes));
    CUDA_TRY(cudaMalloc(&out, 16));
    CUDA_TRY(cudaMemcpyAsync(R, p.C, bytes, cudaMemcpyDeviceToDevice, h->stream));
    const unsigned long long init[2] = {0ULL, ~0ULL};
    CUDA_TRY(cudaMemcpyAsync(out, init, 16, cudaMemcpyHostToDevice, h->stream));
    GemmParams q = p;
    q.C = R;
    if (p.Cin == p.C) q.Cin = R;
    int rc = dispatch_layout<2, 2, 3, 4>(h, q);
    if (!rc) rc = gemm_tma_launch(h, p, 10);
    if (rc) return rc;
    const long long tot = (long long)p.M * p.N;
    gemm_verify_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(p.C, p.ldc, R, p.ldc, p.M, p.N, p.upper_only, out);
    unsigned long long res[2];
    CUDA_TRY(cudaMemcpyAsync(res, out, 16, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (res[0])
        fprintf(stderr, "GEMM_VERIFY call %d: %llu of %lld differ, first at (m %lld, n %lld): M %d N %d K %d ak %d bk %d neg %d epi %d "
                "upper %d lda %lld ldb %lld ldc %lld ldcin %lld A %p B %p C %p Cin %p\n", me, res[0], tot, (long long)(res[1] / p.N),
                (long long)(res[1] % p.N), p.M, p.N, p.K, p.a_kmajor, p.b_kmajor, p.neg, p.epi, p.upper_only, p.lda, p.ldb, p.ldc,
                p.ldcin, (const void*)p.A, (const void*)p.B, (void*)p.C, (const void*)p.Cin);
    cudaFree(R);
    cudaFree(out);
    return ISDF_OK;
}

static bool tma_default()
{
    static const bool on = [] { const char* e = getenv("ISDF_GEMM_TMA"); return !(e && e[0] == '0'); }();
    return on;
}

int gemm_launch(isdf_handle* h, const GemmParams& p)
{
    if (p.M <= 0 || p.N <= 0) return ISDF_OK;
    ARG_CHECK(p.A && p.B && p.C && p.K >= 0);
    ARG_CHECK(p.epi != EPI_MUL || p.E != nullptr);
    // Default since round 2: the warp-specialised kernel of gemm_tma.cu (2-D tensor-map TMA producer lane, mbarrier
    // ring, four 32x32 DMMA consumer warps on a 64x64 CTA tile, 3 CTAs per SM) for every product whose operands
    // TMA can address (16-byte aligned bases and row strides) -- measured on the ten shapes of tools/gemm_bench.py
    // it is faster than the LDGSTS kernel below on each (flush 36.3 vs 33.1 TFLOP/s = cuBLAS's 36.3; W 33.3 vs
    // 28.9; narrow solve 9.2 vs 5.8; profiles/r02_gemm_tma.md) with bit-identical results.  The LDGSTS kernel stays
    // for operands with odd leading dimensions or 8-byte aligned bases, and as tile 4 / ISDF_GEMM_TMA=0.
    // Among its tiles, 64x64 CTAs of four 32x32 warps (128 registers, 4 CTAs per SM) win: against
    // 64x128 CTAs of four 64x32 warps it is 1-5 % faster on the large streaming products and
    // 1.5x faster on small grids, which are bound by the K loop of a single warp (one DMMA per 16
    // clocks and sub-core): half the DMMAs per warp and K step, all four sub-cores busy
    // (profiles/r01_gemm_loaders.md).  Tuning knob: 2 = 64x128 / 64x32 warps, 3 = 64x64 / two 64x32
    // warps, 4 = the LDGSTS default, 5 = 64x128 / eight 32x32 warps, 6 = 128x128 / sixteen 32x32 warps;
    // 7-11: TMA form with 64x128, 128x64, 128x128, 64x64, 32x128 CTA tiles (default: 64x64, 32x128 when M <= 32).
    if (((h->gemm_force >= 7 && h->gemm_force <= 11) || (h->gemm_force == 0 && tma_default())) && gemm_tma_applicable(p)) {
        if (h->gemm_force == 0 && verify_on()) return gemm_verified(h, p);
        const int rc = gemm_tma_launch(h, p, h->gemm_force ? h->gemm_force : (p.M <= 32 ? 11 : 10));
        if (rc != 1) return rc;                 // 1: no tensor map for these operands, nothing launched
    }
    if (h->gemm_force == 2) return dispatch_layout<1, 4, 3, 8>(h, p);
    if (h->gemm_force == 3) return dispatch_layout<1, 2, 3, 8>(h, p);
    if (h->gemm_force == 5) return dispatch_layout<2, 4, 3, 4>(h, p);
    if (h->gemm_force == 6) return dispatch_layout<4, 4, 3, 4>(h, p);
    return dispatch_layout<2, 2, 3, 4>(h, p);
}
