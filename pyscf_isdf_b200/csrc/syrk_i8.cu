// FP64-accurate symmetric product C = alpha * A A^T on the int8 tensor cores (tcgen05,
// sm_100a) by error-free splitting (the Ozaki scheme).
//
// B200 has no FP64 tcgen05 kind: the FP64 tensor pipe (DMMA, gemm.cu) peaks at 37 TFLOP/s
// while kind::i8 runs at 4.5 POP/s with exact int32 accumulation.  Every row of a K-chunk
// of A is scaled by a power of two and cut into S signed 7-bit digit planes
//     A[i,k] = 2^e_i * sum_s q_s[i,k] * 2^(-6-7s),   |q_s| <= 64,
// the digit planes are multiplied exactly on the tensor cores,
//     P_g = sum_{s+t=g} q_s q_t^T          (int32, exact for (g+1)*Kc*64^2 < 2^31),
// and C += alpha * 2^(e_i+e_j) * sum_{g<S} 2^(-12-7g) P_g is accumulated in FP64.  With
// S = 8 the result is as accurate as an FP64 GEMM (profiles/r01_syrk_i8.md); the cost is
// S(S+1)/2 = 36 int8 products instead of one FP64 product at 1/120 of the rate.
//
// It is used for the Coulomb metric W = (vol/Ng^2) Y Y^T, whose entries
// only need the 1e-8 J/K tolerance; the interpolation-point selection stays on the
// bit-reproducible FP64 chain of gemm.cu.
//
// Two kernels.  syrk_i8_kernel2 (default, further down): 256x256 tile per CTA, 2x2 clusters.
// syrk_i8_kernel<CL> (first version, ISDF_SYRK_KERNEL=1), one CTA per SM, persistent over
// upper-triangle 128x256 tiles:
//   warp 0 (one lane)  bulk-copy producer (cp.async.bulk, the TMA unit): the digit planes are
//                      stored pre-tiled and pre-swizzled, [plane][row block][K block] of
//                      128 rows x 128 bytes in the SWIZZLE_128B order, so one operand block is
//                      one contiguous 16 KiB copy (full DRAM pages; a strided 128-row box was
//                      latency bound, profiles/r01_syrk_i8.md) into a 4-stage shared-memory ring
//   warp 1 (one lane)  tcgen05.mma kind::i8, M=128 N=256 K=32, accumulators in TMEM
//                      (2 x 256 columns: the epilogue of unit u overlaps the MMAs of u+1)
//   warps 2-5          epilogue: tcgen05.ld -> int32 -> FP64, scale, C += in global memory
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include "crt_core.cuh"
#include <cmath>
#include <vector>
#include <algorithm>
#include <cstdlib>

namespace {

constexpr int TM = 128;            // tile rows   (UMMA M)
constexpr int TN = 256;            // tile cols   (UMMA N)
constexpr int BKB = 128;           // K bytes per stage = one 128-byte swizzle atom
constexpr int UK = 32;             // K per tcgen05.mma (kind::i8)
constexpr int NSTAGE = 4;
constexpr int A_BYTES = TM * BKB;  // 16 KiB
constexpr int B_BYTES = TN * BKB;  // 32 KiB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SMEM_BYTES = NSTAGE * STAGE_BYTES + 1024;    // + alignment slack
constexpr int NTHREADS = 192;
constexpr int KC_MAX = 32768;      // int32 exactness: 12 * 32768 * 64^2 < 2^31
constexpr int S_MAX = 12;
constexpr int STRIP = 8;           // tile columns per strip of the persistent schedule
constexpr long long WAIT_LIMIT = 4000000000LL;             // clocks before a stuck barrier traps

// ---------------------------------------------------------------- PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok = 0;
    const long long t0 = clock64();
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) break;
        if (clock64() - t0 > WAIT_LIMIT) __trap();          // never hang the device on a protocol bug
    }
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_load_mc(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint16_t mask)
{
    // the copy lands at the same shared-memory offset (and signals the same barrier offset) in every CTA of `mask`
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void tc_commit_mc(uint32_t bar, uint16_t mask)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 :: "r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, int* v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tc_ld16(uint32_t taddr, int* v)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor of a K-major tile whose rows are 128 bytes (one
// SWIZZLE_128B atom): 8-row groups are 1024 bytes apart (stride byte offset), the leading
// byte offset is unused, descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t saddr)
{
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// Instruction descriptor, kind::i8: D = S32 (bits 4-5 = 2), A and B signed 8 bit (bits 7-9,
// 10-12 = 1), both K-major (bits 15, 16 = 0), N >> 3 in bits 17-22, M >> 4 in bits 24-28.
constexpr uint32_t IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

struct SyrkArgs {
    const int8_t* planes;              // [S][nrb][nkb_full] blocks of 128 rows x 128 bytes (swizzled)
    long long plane_stride, rb_stride; // bytes
    const int2* tiles; int ntiles;     // (gi, tj): a cluster of CL CTAs takes row tiles CL*gi + rank, cols 256*tj..
    int n, nkb, S;                     // matrix order, 128-byte K blocks in this chunk, digit planes
    double* C; long long ldc;
    const double* sc;                  // row scales 2^e_i of this chunk
    double alpha;
    // CRT form only (syrk_i8_kernel2<true>): S = number of moduli, unit u multiplies residue plane u with
    // itself and stores the centred residues of the int32 products as int8, R[u][row][col]
    int8_t* R; long long ldr, res_stride;
};

// moduli and their reciprocals for the CRT form (crt_core.cuh)
__constant__ int c_mod[isdf_crt::NM_MAX];
__constant__ double c_invp[isdf_crt::NM_MAX];


// CL = cluster size: the CL CTAs of a cluster own CL vertically adjacent 128x256 tiles; each
// loads its own A block and 1/CL of the common B block, which is multicast to the whole cluster
// (L2 -> SM traffic per tile 16 + 32/CL KiB per stage instead of 48 KiB, and the CTAs of a
// cluster advance in lockstep).  A stage is refilled only after all CL CTAs released it.
template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(NTHREADS, 1)
syrk_i8_kernel(const SyrkArgs a)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t bar_full[NSTAGE], bar_empty[NSTAGE], bar_tfull[2], bar_tempty[2];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;     // SWIZZLE_128B tiles: 1024-byte aligned

    if (threadIdx.x == 0) {
        for (int i = 0; i < NSTAGE; ++i) { mbar_init(smem_u32(&bar_full[i]), 1); mbar_init(smem_u32(&bar_empty[i]), CL); }
        for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&bar_tfull[i]), 1); mbar_init(smem_u32(&bar_tempty[i]), 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {                                                   // whole TMEM: two 256-column accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();                                    // peers' barriers are initialised
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const int rank = CL > 1 ? (int)cluster_rank() : 0;
    const int cid = blockIdx.x / CL, ncl = gridDim.x / CL;
    constexpr uint16_t CMASK = (uint16_t)((1u << CL) - 1);

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int st = 0; uint32_t ph = 0;
            for (int t = cid; t < a.ntiles; t += ncl) {
                const int2 tl = a.tiles[t];
                const int8_t* arow = a.planes + (long long)(tl.x * CL + rank) * a.rb_stride;
                const int8_t* brow = a.planes + (long long)(2 * tl.y) * a.rb_stride;
                for (int g = a.S - 1; g >= 0; --g)
                    for (int s = 0; s <= g; ++s) {
                        const int8_t* pa = arow + (long long)s * a.plane_stride;
                        const int8_t* pb = brow + (long long)(g - s) * a.plane_stride;
                        for (int kb = 0; kb < a.nkb; ++kb) {
                            mbar_wait(smem_u32(&bar_empty[st]), ph ^ 1);    // all CL CTAs are done with this stage
                            const uint32_t full = smem_u32(&bar_full[st]);
                            const uint32_t sa = smem0 + st * STAGE_BYTES, sb = sa + A_BYTES;
                            mbar_expect_tx(full, STAGE_BYTES);
                            bulk_load(sa, pa + (long long)kb * A_BYTES, A_BYTES, full);
                            if (CL == 1) {
                                bulk_load(sb, pb + (long long)kb * A_BYTES, A_BYTES, full);
                                bulk_load(sb + A_BYTES, pb + a.rb_stride + (long long)kb * A_BYTES, A_BYTES, full);
                            } else {
                                constexpr int PB = B_BYTES / CL;            // this CTA's share of the B tile
                                const int off = rank * PB;
                                bulk_load_mc(sb + off, pb + (long long)(off / A_BYTES) * a.rb_stride +
                                             (long long)kb * A_BYTES + (off % A_BYTES), PB, full, CMASK);
                            }
                            if (++st == NSTAGE) { st = 0; ph ^= 1; }
                        }
                    }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            int st = 0; uint32_t ph = 0;
            uint32_t unit = 0;
            for (int t = cid; t < a.ntiles; t += ncl)
                for (int g = a.S - 1; g >= 0; --g, ++unit) {
                    const uint32_t buf = unit & 1, bph = (unit >> 1) & 1;
                    mbar_wait(smem_u32(&bar_tempty[buf]), bph ^ 1);     // epilogue has drained this accumulator
                    tc_fence_after();
                    const uint32_t tmem_d = tmem_base + buf * TN;
                    const int niter = (g + 1) * a.nkb;
                    for (int it = 0; it < niter; ++it) {
                        mbar_wait(smem_u32(&bar_full[st]), ph);
                        tc_fence_after();
                        const uint32_t sa = smem0 + st * STAGE_BYTES, sb = sa + A_BYTES;
#pragma unroll
                        for (int k = 0; k < BKB / UK; ++k)
                            tc_mma_i8(tmem_d, umma_desc_k128(sa + k * UK), umma_desc_k128(sb + k * UK), IDESC,
                                      (it | k) != 0);
                        // frees the stage (in every CTA of the cluster) when these MMAs retire
                        if (CL == 1) tc_commit(smem_u32(&bar_empty[st]));
                        else tc_commit_mc(smem_u32(&bar_empty[st]), CMASK);
                        if (++st == NSTAGE) { st = 0; ph ^= 1; }
                    }
                    tc_commit(smem_u32(&bar_tfull[buf]));               // accumulator complete
                }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
        const int q = warp & 3;
        const bool vec2 = (a.ldc % 2 == 0) && ((uintptr_t)a.C % 16 == 0);
        uint32_t unit = 0;
        for (int t = cid; t < a.ntiles; t += ncl) {
            const int2 tl = a.tiles[t];
            const int row = (tl.x * CL + rank) * TM + q * 32 + lane;
            const double srow = row < a.n ? a.sc[row] : 0.0;
            double* crow = a.C + (long long)row * a.ldc;
            for (int g = a.S - 1; g >= 0; --g, ++unit) {
                const uint32_t buf = unit & 1, bph = (unit >> 1) & 1;
                const double f = a.alpha * srow * exp2((double)(-12 - 7 * g));
                mbar_wait(smem_u32(&bar_tfull[buf]), bph);
                tc_fence_after();
#pragma unroll 1
                for (int cc = 0; cc < TN / 32; ++cc) {
                    int v[32];
                    tc_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * TN + cc * 32, v);
                    if (cc == TN / 32 - 1) {                            // all TMEM reads of this unit are done
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(&bar_tempty[buf]));
                    }
                    const int col0 = tl.y * TN + cc * 32;
                    if (row < a.n && col0 + 31 >= row && col0 < a.n) {
                        if (col0 >= row && col0 + 32 <= a.n) {
                            // interior chunk: all 32 loads in flight before the first use
                            double c[32];
                            if (vec2) {
                                const double2* p2 = reinterpret_cast<const double2*>(crow + col0);
#pragma unroll
                                for (int j = 0; j < 16; ++j) { const double2 t2 = p2[j]; c[2 * j] = t2.x; c[2 * j + 1] = t2.y; }
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j) c[j] = crow[col0 + j];
                            }
#pragma unroll
                            for (int j = 0; j < 32; ++j) c[j] += (double)v[j] * (f * a.sc[col0 + j]);
                            if (vec2) {
                                double2* p2 = reinterpret_cast<double2*>(crow + col0);
#pragma unroll
                                for (int j = 0; j < 16; ++j) p2[j] = make_double2(c[2 * j], c[2 * j + 1]);
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j) crow[col0 + j] = c[j];
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const int col = col0 + j;
                                if (col < a.n && col >= row) crow[col] += (double)v[j] * (f * a.sc[col]);
                            }
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();            // no CTA exits while a peer may still signal its barriers
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(512) : "memory");
    }
}

// ---- second kernel: 256x256 tile per CTA, clusters of 2x2 CTAs (512x512 per cluster) ----------
// The 128x256-tile kernel above is bound by what one SM can take in from L2 (48 KiB per
// 128x256x128 MMA block, ~55 B/clk/SM; profiles/r01_syrk_i8.md).  Here a CTA owns two row tiles
// that share one B block (two MMAs per K step into the two halves of TMEM), 64 KiB per stage for
// twice the work, and the 2x2 cluster multicasts A along rows and B along columns: every CTA
// issues 32 KiB of copies per stage and receives 64 KiB.  TMEM is one 512-column buffer, so the
// epilogue (8 warps) is not overlapped with the next unit; it is ~3 % of a unit.
constexpr int NST2 = 3;
constexpr int STAGE2 = 4 * A_BYTES;                       // A0 | A1 | B0 | B1, 16 KiB each
constexpr int SMEM2 = NST2 * STAGE2 + 1024;
constexpr int NTHREADS2 = 320;                            // producer, MMA, 8 epilogue warps
constexpr int CT = 512;                                   // cluster tile

// CRT = true: the residue form (crt_core.cuh): unit g multiplies residue plane g with itself over the
// chunk (one product per modulus instead of g+1 per digit weight) and the epilogue stores the int32
// accumulators reduced modulo the g-th modulus; everything else -- schedule, copies, barriers -- is shared.
template <bool CRT>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(NTHREADS2, 1)
syrk_i8_kernel2(const SyrkArgs a)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t bar_full[NST2], bar_empty[NST2], bar_tfull, bar_tempty;
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NST2; ++i) { mbar_init(smem_u32(&bar_full[i]), 1); mbar_init(smem_u32(&bar_empty[i]), 4); }
        mbar_init(smem_u32(&bar_tfull), 1);
        mbar_init(smem_u32(&bar_tempty), 8);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const int rank = (int)cluster_rank();
    const int r = rank >> 1, c = rank & 1;                 // position in the 2x2 cluster
    const int cid = blockIdx.x / 4, ncl = gridDim.x / 4;
    const uint16_t mask_row = (uint16_t)(3u << (2 * r));   // the two CTAs of my cluster row
    const uint16_t mask_col = (uint16_t)((1u << c) | (1u << (c + 2)));

    if (warp == 0) {
        if (lane == 0) {
            // ===== bulk-copy producer: A_c of my row block -> my cluster row, B_r of my column tile -> my cluster column
            int st = 0; uint32_t ph = 0;
            for (int t = cid; t < a.ntiles; t += ncl) {
                const int2 tl = a.tiles[t];
                const int8_t* arow = a.planes + (long long)(4 * tl.x + 2 * r + c) * a.rb_stride;
                const int8_t* brow = a.planes + (long long)(4 * tl.y + 2 * c + r) * a.rb_stride;
                for (int g = a.S - 1; g >= 0; --g)
                    for (int s = CRT ? g : 0; s <= g; ++s) {
                        const int8_t* pa = arow + (long long)s * a.plane_stride;
                        const int8_t* pb = brow + (long long)(CRT ? s : g - s) * a.plane_stride;
                        for (int kb = 0; kb < a.nkb; ++kb) {
                            mbar_wait(smem_u32(&bar_empty[st]), ph ^ 1);
                            const uint32_t full = smem_u32(&bar_full[st]);
                            const uint32_t sa = smem0 + st * STAGE2, sb = sa + 2 * A_BYTES;
                            mbar_expect_tx(full, STAGE2);
                            bulk_load_mc(sa + c * A_BYTES, pa + (long long)kb * A_BYTES, A_BYTES, full, mask_row);
                            bulk_load_mc(sb + r * A_BYTES, pb + (long long)kb * A_BYTES, A_BYTES, full, mask_col);
                            if (++st == NST2) { st = 0; ph ^= 1; }
                        }
                    }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer: two M=128 MMAs per K step share the B block =====
            int st = 0; uint32_t ph = 0;
            uint32_t unit = 0;
            for (int t = cid; t < a.ntiles; t += ncl) {
                const int2 tl = a.tiles[t];
                const int row0 = tl.x * CT + r * 256, col0 = tl.y * CT + c * 256;
                const bool idle = row0 > col0 + 255 || row0 >= a.n || col0 >= a.n;    // nothing of this tile is stored
                for (int g = a.S - 1; g >= 0; --g, ++unit) {
                    mbar_wait(smem_u32(&bar_tempty), (unit & 1) ^ 1);
                    tc_fence_after();
                    const int niter = (CRT ? 1 : g + 1) * a.nkb;
                    for (int it = 0; it < niter; ++it) {
                        mbar_wait(smem_u32(&bar_full[st]), ph);
                        tc_fence_after();
                        if (!idle) {
                            const uint32_t sa = smem0 + st * STAGE2, sb = sa + 2 * A_BYTES;
#pragma unroll
                            for (int k = 0; k < BKB / UK; ++k) {
                                const uint64_t db = umma_desc_k128(sb + k * UK);
                                tc_mma_i8(tmem_base, umma_desc_k128(sa + k * UK), db, IDESC, (it | k) != 0);
                                tc_mma_i8(tmem_base + TN, umma_desc_k128(sa + A_BYTES + k * UK), db, IDESC, (it | k) != 0);
                            }
                        }
                        tc_commit_mc(smem_u32(&bar_empty[st]), 0xF);
                        if (++st == NST2) { st = 0; ph ^= 1; }
                    }
                    tc_commit(smem_u32(&bar_tfull));
                }
            }
        }
    } else {
        // ===== epilogue: warps 2..9; lane quarter = warp % 4, accumulator half = (warp - 2) / 4 =====
        const int q = warp & 3, hh = (warp - 2) >> 2;
        const bool vec2 = (a.ldc % 2 == 0) && ((uintptr_t)a.C % 16 == 0);
        uint32_t unit = 0;
        for (int t = cid; t < a.ntiles; t += ncl) {
            const int2 tl = a.tiles[t];
            const int row0 = tl.x * CT + r * 256, colt = tl.y * CT + c * 256;
            const bool idle = row0 > colt + 255 || row0 >= a.n || colt >= a.n;
            const int row = row0 + hh * 128 + q * 32 + lane;
            const double srow = row < a.n ? a.sc[row] : 0.0;
            double* crow = a.C + (long long)row * a.ldc;
            for (int g = a.S - 1; g >= 0; --g, ++unit) {
                const double f = a.alpha * srow * exp2((double)(-12 - 7 * g));
                mbar_wait(smem_u32(&bar_tfull), unit & 1);
                tc_fence_after();
                if (!idle) {
#pragma unroll 1
                    for (int cc = 0; cc < TN / 16; ++cc) {
                        int v[16];
                        tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + hh * TN + cc * 16, v);
                        if (cc == TN / 16 - 1) {                        // last TMEM read: the MMA warp may go on
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(smem_u32(&bar_tempty));
                        }
                        const int col0 = colt + cc * 16;
                        if constexpr (CRT) {
                            // residues of the 16 accumulators modulo the unit's modulus, one 16-byte store
                            // (ldr is a multiple of 16 >= n, so a chunk that starts inside the matrix fits the row)
                            if (row < a.n && col0 + 15 >= row && col0 < a.n) {
                                const int pm = c_mod[g];
                                const double ipm = c_invp[g];
                                uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    w[j >> 2] |= ((uint32_t)isdf_crt::centred_mod_rt(v[j], pm, ipm) & 0xffu) << (8 * (j & 3));
                                *reinterpret_cast<uint4*>(a.R + (long long)g * a.res_stride + (long long)row * a.ldr + col0) =
                                    make_uint4(w[0], w[1], w[2], w[3]);
                            }
                        } else
                        if (row < a.n && col0 + 15 >= row && col0 < a.n) {
                            if (col0 >= row && col0 + 16 <= a.n) {
                                double cv[16];
                                if (vec2) {
                                    const double2* p2 = reinterpret_cast<const double2*>(crow + col0);
#pragma unroll
                                    for (int j = 0; j < 8; ++j) { const double2 t2 = p2[j]; cv[2 * j] = t2.x; cv[2 * j + 1] = t2.y; }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) cv[j] = crow[col0 + j];
                                }
#pragma unroll
                                for (int j = 0; j < 16; ++j) cv[j] += (double)v[j] * (f * a.sc[col0 + j]);
                                if (vec2) {
                                    double2* p2 = reinterpret_cast<double2*>(crow + col0);
#pragma unroll
                                    for (int j = 0; j < 8; ++j) p2[j] = make_double2(cv[2 * j], cv[2 * j + 1]);
                                } else {
#pragma unroll
                                    for (int j = 0; j < 16; ++j) crow[col0 + j] = cv[j];
                                }
                            } else {
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    const int col = col0 + j;
                                    if (col < a.n && col >= row) crow[col] += (double)v[j] * (f * a.sc[col]);
                                }
                            }
                        }
                    }
                }
                else {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(&bar_tempty));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(512) : "memory");
    }
}

// ---- digit planes -----------------------------------------------------------------
// Row statistics, one CTA per row: sc[c][i] = 2^e with max_k |A[i, chunk c]| = m * 2^e,
// m in [0.5, 1), and nrm[i] = ||A[i, :]||_2.
__global__ void i8_rowstats_kernel(const double* __restrict__ A, long long lda, long long K, int kc_full, int nchunks,
                                   int n, double* __restrict__ sc, double* __restrict__ nrm)
{
    const int i = blockIdx.x;
    const double* row = A + (long long)i * lda;
    __shared__ double red[32];
    double ssq = 0.0;
    for (int c = 0; c < nchunks; ++c) {
        const long long k0 = (long long)c * kc_full;
        const int kc = (int)min((long long)kc_full, K - k0);
        double m = 0.0;
        for (int k = threadIdx.x; k < kc; k += blockDim.x) {
            const double v = row[k0 + k];
            m = fmax(m, fabs(v));
            ssq = fma(v, v, ssq);
        }
        for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        __syncthreads();
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x < 32) {
            m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
            for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            if (threadIdx.x == 0) {
                int e = 0;
                if (m > 0.0) frexp(m, &e);
                sc[(long long)c * n + i] = ldexp(1.0, e);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) ssq += __shfl_xor_sync(0xffffffffu, ssq, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ssq;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        nrm[i] = sqrt(t);
    }
}

// rho[c] = max_i sc[c][i] / nrm[i]: how large chunk c is relative to the whole rows (one CTA per chunk)
__global__ void i8_rho_kernel(const double* __restrict__ sc, const double* __restrict__ nrm, int n, double* __restrict__ rho)
{
    const double* s = sc + (long long)blockIdx.x * n;
    double m = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        if (nrm[i] > 0.0) m = fmax(m, s[i] / nrm[i]);
    __shared__ double red[32];
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = fmax(m, red[w]);
        rho[blockIdx.x] = fmax(m, red[0]);
    }
}

// s-th signed 7-bit digit of A[i, k0+k] / sc[i], 16 consecutive k per thread, stored in the
// pre-tiled SWIZZLE_128B order: block (i / 128, k / 128) is 16 KiB contiguous, row r of a block
// at r * 128 bytes, its 16-byte chunk c at position c ^ (r & 7).  Rows >= n (padding up to a
// multiple of 256) and columns >= kc are written as zeros.
__global__ void i8_split_kernel(const double* __restrict__ A, long long lda, int n, int nrows_pad, long long k0,
                                int kc, int kcpad, const double* __restrict__ sc, int8_t* __restrict__ planes,
                                long long rb_stride, long long plane_stride, int S)
{
    const int groups = kcpad / 16;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)nrows_pad * groups) return;
    const int i = (int)(idx / groups), kg = (int)(idx - (long long)i * groups) * 16;
    double x[16];
    if (i < n) {
        const double inv = 1.0 / sc[i];                 // exact: a power of two
        const double* src = A + (long long)i * lda + k0 + kg;
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = (kg + j < kc) ? src[j] * inv : 0.0;
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = 0.0;
    }
    const int r = i & 127, c = (kg & 127) >> 4;
    int8_t* dst = planes + (long long)(i >> 7) * rb_stride + (long long)(kg >> 7) * A_BYTES + r * 128 + ((c ^ (r & 7)) << 4);
    for (int s = 0; s < S; ++s) {
        const double mult = s == 0 ? 64.0 : 128.0;
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const double y = x[j] * mult;
            const double qd = rint(y);
            x[j] = y - qd;                              // exact remainder, |.| <= 1/2
            w[j >> 2] |= ((uint32_t)(__double2int_rn(qd)) & 0xffu) << (8 * (j & 3));
        }
        *reinterpret_cast<uint4*>(dst + s * plane_stride) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---- CRT form: residue planes and reconstruction (crt_core.cuh) --------------------------------
// Residue plane m of rint(A[i, k0+k] / sc[i] * 2^bits) modulo modulus(m), same pre-tiled SWIZZLE_128B
// layout as the digit planes (16 consecutive k per thread, one 16-byte store per plane).
__global__ void __launch_bounds__(256, 3) i8_split_crt_kernel(const double* __restrict__ A, long long lda, int n, int nrows_pad, long long k0,
                                    int kc, int kcpad, const double* __restrict__ sc, int8_t* __restrict__ planes,
                                    long long rb_stride, long long plane_stride, int nm, int bits)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    isdf_crt::split_item(idx, A, lda, n, nrows_pad, k0, kc, kcpad, sc, planes, rb_stride, plane_stride, nm, bits);
}

// C[row, col] += alpha sc[row] sc[col] 2^(-2 bits) x(row, col) for col >= row, x rebuilt from its nm residues
__global__ void crt_reconstruct_kernel(const int8_t* __restrict__ R, long long ldr, long long res_stride, int nm,
                                       int n, const double* __restrict__ sc, double f, double* __restrict__ C, long long ldc)
{
    isdf_crt::reconstruct_item((int)blockIdx.y, (int)(blockIdx.x * blockDim.x + threadIdx.x), R, ldr, res_stride, nm, n, sc, f, C, ldc);
}

struct Chunks { int nchunks, kc_full; };

Chunks plan_chunks(long long K)
{
    Chunks c;
    c.nchunks = (int)((K + KC_MAX - 1) / KC_MAX);
    const long long per = (K + c.nchunks - 1) / c.nchunks;
    c.kc_full = (int)align_up((size_t)per, BKB);
    return c;
}


thread_local double g_last_ops = 0.0;      // int8 operations (2 per multiply-add) issued by the last isdf_syrk_i8 call

int upload_crt_tables(void* arg)
{
    isdf_handle* h = (isdf_handle*)arg;
    int hm[isdf_crt::NM_MAX]; double hi[isdf_crt::NM_MAX];
    for (int k = 0; k < isdf_crt::NM_MAX; ++k) { hm[k] = isdf_crt::modulus(k); hi[k] = 1.0 / (double)hm[k]; }
    CUDA_TRY(cudaMemcpyToSymbolAsync(c_mod, hm, sizeof(hm), 0, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpyToSymbolAsync(c_invp, hi, sizeof(hi), 0, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));    // the host copies are locals
    return ISDF_OK;
}

int cluster_size()
{
    static const int cl = [] {                                // read once, thread-safe
        const char* e = getenv("ISDF_SYRK_CLUSTER");          // tuning knob: 1, 2, 4 or 8 CTAs per cluster
        const int c = e ? atoi(e) : 4;
        return (c != 1 && c != 2 && c != 4 && c != 8) ? 4 : c;
    }();
    return cl;
}

// 2 (default): 256x256 tiles, 2x2 clusters (syrk_i8_kernel2); 1: 128x256 tiles, CL x 1 clusters
int kernel_variant()
{
    static const int v = [] {
        const char* e = getenv("ISDF_SYRK_KERNEL");
        return (e && atoi(e) == 1) ? 1 : 2;
    }();
    return v;
}

// The residue form of the product (crt_core.cuh) with 16 moduli is the default: on the C64 workload it gives the
// same W as the digit-plane form to 2e-12 (J/K to 2e-13) in 462 ms instead of 737 ms (profiles/r01_optin_forms.md).
// ISDF_SYRK_CRT=<moduli> (8..16) changes the ceiling, ISDF_SYRK_CRT=0 selects the digit planes.
// Needs the 256x256 kernel (variant 2).
int crt_moduli()
{
    const char* e = getenv("ISDF_SYRK_CRT");
    int nm = e ? atoi(e) : 16;
    if (nm <= 0 || kernel_variant() != 2) return 0;
    return std::max(8, std::min(nm, (int)isdf_crt::NM_MAX));
}

// ISDF_SYRK_OVERLAP (default 1, residue form only): the element-wise kernels of chunk c+1 / c-1 run on a second
// stream beside the tensor-core products of chunk c (double-buffered residue planes and residue output).
// The second set of buffers is only taken while it stays small: n <= 16384 rows (18 GB in all for C64's
// n = 12480; Si216's n = 42120 would need 100 GB and runs the one-stream form).
constexpr int OVERLAP_MAX_ROWS = 16384;
bool overlap_requested(int n)
{
    const char* e = getenv("ISDF_SYRK_OVERLAP");      // default on since round 2 (bit-identical to the one-stream form); 0 = off
    return (e ? atoi(e) > 0 : true) && crt_moduli() > 0 && n <= OVERLAP_MAX_ROWS;
}

int rows_padded(int n, int cl)
{
    if (kernel_variant() == 2) return (int)align_up((size_t)n, (size_t)CT);
    return (int)align_up((size_t)n, (size_t)std::max(TN, TM * cl));
}

template <bool CRT>
int launch_syrk2(isdf_handle* h, const SyrkArgs& a, int& nclusters, cudaStream_t stream = nullptr)
{
    if (!stream) stream = h->stream;
    if (nclusters == 0) {
        { int rc_; if ((rc_ = isdf_func_max_smem((const void*)syrk_i8_kernel2<CRT>, SMEM2))) return rc_; }
        int fit = h->num_sms / 4;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(h->num_sms / 4 * 4); cfg.blockDim = dim3(NTHREADS2); cfg.dynamicSmemBytes = SMEM2;
        cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension;
        at.val.clusterDim.x = 4; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at; cfg.numAttrs = 1;
        int nmax = 0;
        if (cudaOccupancyMaxActiveClusters(&nmax, syrk_i8_kernel2<CRT>, &cfg) == cudaSuccess && nmax > 0) fit = std::min(fit, nmax);
        else (void)cudaGetLastError();
        nclusters = std::max(1, std::min(fit, a.ntiles));
    }
    syrk_i8_kernel2<CRT><<<nclusters * 4, NTHREADS2, SMEM2, stream>>>(a);
    return ISDF_OK;
}

template <int CL>
int launch_syrk(isdf_handle* h, const SyrkArgs& a, int& nclusters)
{
    auto kern = syrk_i8_kernel<CL>;
    if (nclusters == 0) {
        { int rc_; if ((rc_ = isdf_func_max_smem((const void*)kern, SMEM_BYTES))) return rc_; }
        int fit = h->num_sms / CL;
        if (CL > 1) {                                          // how many clusters are co-resident on this GPU
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(h->num_sms / CL * CL); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = SMEM_BYTES;
            cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = CL; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
            cfg.attrs = &at; cfg.numAttrs = 1;
            int nmax = 0;
            if (cudaOccupancyMaxActiveClusters(&nmax, kern, &cfg) == cudaSuccess && nmax > 0) fit = std::min(fit, nmax);
            else (void)cudaGetLastError();
        }
        nclusters = std::max(1, std::min(fit, a.ntiles));
    }
    kern<<<nclusters * CL, NTHREADS, SMEM_BYTES, h->stream>>>(a);
    return ISDF_OK;
}

}  // namespace

extern "C" double isdf_syrk_i8_last_ops(void) { return g_last_ops; }

extern "C" size_t isdf_syrk_i8_ws_bytes(int n, long long K, int nslices)
{
    if (n <= 0 || K <= 0 || nslices <= 0) return 0;
    const Chunks c = plan_chunks(K);
    const int cl = cluster_size();
    const int tr = (n + TM - 1) / TM, tc = (n + TN - 1) / TN;
    const int nm = crt_moduli();
    const size_t nbuf = overlap_requested(n) ? 2 : 1;
    const size_t crt_extra = nm ? nbuf * align_up((size_t)nm * (size_t)n * align_up((size_t)n, 16), 1024) : 0;    // residues of the products
    return crt_extra + nbuf * align_up((size_t)std::max(nslices, nm) * (size_t)rows_padded(n, cl) * ((c.kc_full / BKB | 1) * BKB), 1024) +
           align_up((size_t)(c.nchunks + 1) * n * sizeof(double), 256) + align_up((size_t)c.nchunks * sizeof(double), 256) +
           align_up((size_t)tr * tc * sizeof(int2), 256) + 2048;
}

extern "C" int isdf_syrk_i8(isdf_handle* h, int n, long long K, const double* A, long long lda,
                            double* C, long long ldc, double alpha, int nslices, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && A && C && ws);
    ARG_CHECK(n > 0 && K > 0 && lda >= K && ldc >= n);
    ARG_CHECK(nslices >= 2 && nslices <= S_MAX);
    if (ws_bytes < isdf_syrk_i8_ws_bytes(n, K, nslices)) {
        isdf_set_error("isdf_syrk_i8: workspace too small");
        return ISDF_ERR_WS;
    }
    const Chunks ch = plan_chunks(K);
    const int cl = cluster_size();
    const int tr = (n + TM - 1) / TM, tc = (n + TN - 1) / TN;
    const int nrows_pad = rows_padded(n, cl);              // whole 256-row B tiles and whole clusters of row tiles
    // an odd number of 16 KiB blocks per row block: a power-of-two stride would map the ~30 row and
    // column blocks that are live at the same K offset onto the same L2 sets
    const long long rb_stride = (long long)(ch.kc_full / BKB | 1) * A_BYTES;
    const long long plane_stride = rb_stride * (nrows_pad / TM);
    char* base = (char*)align_up((size_t)ws, 1024);
    const int nm_max = crt_moduli();
    const bool overlap = overlap_requested(n);
    int8_t* planes = (int8_t*)base;
    base += align_up((size_t)std::max(nslices, nm_max) * plane_stride, 1024);
    int8_t* planes2 = planes;                              // second set of residue planes (overlap form)
    if (overlap) { planes2 = (int8_t*)base; base += align_up((size_t)std::max(nslices, nm_max) * plane_stride, 1024); }
    const long long ldr = (long long)align_up((size_t)n, 16);
    const long long res_stride = (long long)n * ldr;
    int8_t* resid = (int8_t*)base;                         // [nm][n][ldr] residues of one chunk's products (CRT form)
    if (nm_max) base += align_up((size_t)nm_max * res_stride, 1024);
    int8_t* resid2 = resid;
    if (overlap) { resid2 = (int8_t*)base; base += align_up((size_t)nm_max * res_stride, 1024); }
    double* sc_all = (double*)base;                        // [nchunks][n] row scales, then [n] row norms
    double* nrm = sc_all + (size_t)ch.nchunks * n;
    base += align_up((size_t)(ch.nchunks + 1) * n * sizeof(double), 256);
    double* d_rho = (double*)base;
    base += align_up((size_t)ch.nchunks * sizeof(double), 256);
    int2* d_tiles = (int2*)base;

    // persistent schedule: strips of STRIP tile columns, row groups ascending inside a strip, so that
    // the tiles in flight share few distinct row/column blocks (L2 reuse of the digit planes)
    std::vector<int2> tiles;
    const int variant = kernel_variant();
    if (variant == 2) {
        const int nb = nrows_pad / CT;                     // 512x512 cluster tiles of the upper triangle
        for (int bj = 0; bj < nb; ++bj)
            for (int bi = 0; bi <= bj; ++bi) tiles.push_back(make_int2(bi, bj));
    } else {
        const int ngr = (tr + cl - 1) / cl;
        for (int s0 = 0; s0 < tc; s0 += STRIP)
            for (int gi = 0; gi < ngr; ++gi)
                for (int tj = s0; tj < std::min(tc, s0 + STRIP); ++tj)
                    if (gi * cl <= 2 * tj + 1) tiles.push_back(make_int2(gi, tj));
    }
    if (getenv("ISDF_SYRK_DEBUG_SAMETILE"))               // traffic experiment only (results are garbage):
        for (auto& t : tiles) t = make_int2(0, 0);         // every cluster streams the same blocks
    const int ntiles = (int)tiles.size();
    CUDA_TRY(cudaMemcpyAsync(d_tiles, tiles.data(), ntiles * sizeof(int2), cudaMemcpyHostToDevice, h->stream));

    // Digit planes per chunk.  A chunk whose entries are small against the whole rows needs fewer
    // planes for the same normwise accuracy: the error of chunk c in C[i,j] is about
    // sqrt(kc) (S+3) 2^(-7S) sc_i sc_j, to be kept below 2^-T |a_i| |a_j| with T = 7 nslices - 6
    // (nslices = 8: T = 50, the accuracy of an FP64 GEMM over this K).  The weighted spectra of
    // the W build decay with |G|, so the high-frequency chunks take 2-4 planes less.
    ARG_CHECK(ch.nchunks * sizeof(double) <= 65536);
    i8_rowstats_kernel<<<n, 256, 0, h->stream>>>(A, lda, K, ch.kc_full, ch.nchunks, n, sc_all, nrm);
    i8_rho_kernel<<<ch.nchunks, 256, 0, h->stream>>>(sc_all, nrm, n, d_rho);
    double* h_rho = (double*)h->h_pinned;
    CUDA_TRY(cudaMemcpyAsync(h_rho, d_rho, ch.nchunks * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));          // also covers `tiles` (pageable host memory)
    h->other_launches += 2;
    std::vector<int> planes_of(ch.nchunks, 0);
    const double T = 7.0 * nslices - 6.0;
    for (int c = 0; c < ch.nchunks; ++c) {
        const long long k0 = (long long)c * ch.kc_full;
        if (k0 >= K || !(h_rho[c] > 0.0)) continue;
        const double kc = (double)std::min<long long>(ch.kc_full, K - k0);
        int S = 0;
        while (S < nslices && 7.0 * S < T + 0.5 * std::log2(kc) + std::log2(S + 3.0) + 2.0 * std::log2(h_rho[c])) ++S;
        planes_of[c] = S;
    }

    CUDA_TRY(cudaMemset2DAsync(C, ldc * sizeof(double), 0, (size_t)n * sizeof(double), n, h->stream));

    if (getenv("ISDF_SYRK_VERBOSE")) {
        fprintf(stderr, "isdf_syrk_i8: n=%d K=%lld chunks=%d x %d, planes per chunk:", n, K, ch.nchunks, ch.kc_full);
        for (int c = 0; c < ch.nchunks; ++c) fprintf(stderr, " %d", planes_of[c]);
        fprintf(stderr, "\n");
    }
    g_last_ops = 0.0;
    int nclusters = 0;
    if (nm_max) {
        // the moduli tables live in constant memory: one upload per device, under the set-up lock
        static const char tables_key = 0;
        int rc_;
        if ((rc_ = isdf_once_per_device(&tables_key, upload_crt_tables, h))) return rc_;
    }
    if (overlap) {
        // Same kernels, same order per buffer and the same order of the updates of C as the loop below; only
        // the streams differ.  s1 carries split(0), split(1), then recon(i), split(i+2) for i = 0, 1, ...;
        // h->stream carries the products.  planes/resid buffer of chunk i: i & 1.
        std::vector<int> act;
        for (int c = 0; c < ch.nchunks; ++c)
            if ((long long)c * ch.kc_full < K && planes_of[c] > 0) act.push_back(c);
        const int m = (int)act.size();
        cudaStream_t s1 = nullptr;
        CUDA_TRY(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
        std::vector<cudaEvent_t> ev(3 * (size_t)m + 1);
        for (auto& e : ev) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        cudaEvent_t* ev_split = ev.data();
        cudaEvent_t* ev_syrk = ev.data() + m;
        cudaEvent_t* ev_recon = ev.data() + 2 * m;
        CUDA_TRY(cudaEventRecord(ev[3 * m], h->stream));          // s1 starts after the row statistics
        CUDA_TRY(cudaStreamWaitEvent(s1, ev[3 * m], 0));
        struct Geo { long long k0; int kc, kcpad, nm, bits; const double* sc; };
        auto geo = [&](int i) {
            Geo g;
            const int c = act[i];
            g.k0 = (long long)c * ch.kc_full;
            g.kc = (int)std::min<long long>(ch.kc_full, K - g.k0);
            g.kcpad = (int)align_up((size_t)g.kc, BKB);
            g.nm = 8;
            while (g.nm < nm_max && isdf_crt::input_bits(g.nm) < 7 * planes_of[c] - 5) ++g.nm;
            g.bits = isdf_crt::input_bits(g.nm);
            g.sc = sc_all + (size_t)c * n;
            return g;
        };
        auto split = [&](int i) -> int {
            const Geo g = geo(i);
            if (i >= 2) CUDA_TRY(cudaStreamWaitEvent(s1, ev_syrk[i - 2], 0));     // the products that read this buffer
            const long long items = (long long)nrows_pad * (g.kcpad / 16);
            i8_split_crt_kernel<<<(unsigned)((items + 255) / 256), 256, 0, s1>>>(
                A, lda, n, nrows_pad, g.k0, g.kc, g.kcpad, g.sc, (i & 1) ? planes2 : planes, rb_stride, plane_stride, g.nm, g.bits);
            CUDA_TRY(cudaEventRecord(ev_split[i], s1));
            return ISDF_OK;
        };
        int rc;
        if (m > 0 && (rc = split(0)) != ISDF_OK) return rc;
        if (m > 1 && (rc = split(1)) != ISDF_OK) return rc;
        for (int i = 0; i < m; ++i) {
            const Geo g = geo(i);
            CUDA_TRY(cudaStreamWaitEvent(h->stream, ev_split[i], 0));
            if (i >= 2) CUDA_TRY(cudaStreamWaitEvent(h->stream, ev_recon[i - 2], 0));    // the reader of this residue buffer
            SyrkArgs a;
            a.planes = (i & 1) ? planes2 : planes; a.plane_stride = plane_stride; a.rb_stride = rb_stride;
            a.tiles = d_tiles; a.ntiles = ntiles; a.n = n; a.nkb = g.kcpad / BKB; a.S = g.nm;
            a.C = C; a.ldc = ldc; a.sc = g.sc; a.alpha = alpha;
            a.R = (i & 1) ? resid2 : resid; a.ldr = ldr; a.res_stride = res_stride;
            if ((rc = launch_syrk2<true>(h, a, nclusters)) != ISDF_OK) return rc;
            CUDA_TRY(cudaEventRecord(ev_syrk[i], h->stream));
            CUDA_TRY(cudaStreamWaitEvent(s1, ev_syrk[i], 0));
            dim3 rg((unsigned)((n + 127) / 128), (unsigned)n);
            crt_reconstruct_kernel<<<rg, 128, 0, s1>>>((i & 1) ? resid2 : resid, ldr, res_stride, g.nm, n, g.sc,
                                                        alpha * ldexp(1.0, -2 * g.bits), C, ldc);
            CUDA_TRY(cudaEventRecord(ev_recon[i], s1));
            if (i + 2 < m && (rc = split(i + 2)) != ISDF_OK) return rc;
            g_last_ops += 2.0 * g.nm * (double)ntiles * (double)CT * CT * (double)g.kcpad;
            h->other_launches += 3;
        }
        if (m > 0) CUDA_TRY(cudaStreamWaitEvent(h->stream, ev_recon[m - 1], 0));     // C is complete for whatever follows
        CUDA_TRY(cudaGetLastError());
        for (auto& e : ev) CUDA_TRY(cudaEventDestroy(e));          // released when the recorded work has finished
        CUDA_TRY(cudaStreamDestroy(s1));
        return ISDF_OK;
    }
    for (int c = 0; c < ch.nchunks; ++c) {
        const long long k0 = (long long)c * ch.kc_full;
        if (k0 >= K) break;
        const int kc = (int)std::min<long long>(ch.kc_full, K - k0);
        const int kcpad = (int)align_up((size_t)kc, BKB);
        const int S = planes_of[c];
        if (S == 0) continue;                              // below the accuracy target altogether
        const double* sc = sc_all + (size_t)c * n;
        const long long items = (long long)nrows_pad * (kcpad / 16);
        if (nm_max) {
            // the chunk's digit-plane count S stands for 7 S - 6 accurate bits; take the fewest moduli that carry them
            int nm = 8;
            while (nm < nm_max && isdf_crt::input_bits(nm) < 7 * S - 5) ++nm;
            const int bits = isdf_crt::input_bits(nm);
            i8_split_crt_kernel<<<(unsigned)((items + 255) / 256), 256, 0, h->stream>>>(
                A, lda, n, nrows_pad, k0, kc, kcpad, sc, planes, rb_stride, plane_stride, nm, bits);
            SyrkArgs a;
            a.planes = planes; a.plane_stride = plane_stride; a.rb_stride = rb_stride;
            a.tiles = d_tiles; a.ntiles = ntiles; a.n = n; a.nkb = kcpad / BKB; a.S = nm;
            a.C = C; a.ldc = ldc; a.sc = sc; a.alpha = alpha;
            a.R = resid; a.ldr = ldr; a.res_stride = res_stride;
            int rc;
            if ((rc = launch_syrk2<true>(h, a, nclusters)) != ISDF_OK) return rc;
            dim3 rg((unsigned)((n + 127) / 128), (unsigned)n);
            crt_reconstruct_kernel<<<rg, 128, 0, h->stream>>>(resid, ldr, res_stride, nm, n, sc,
                                                              alpha * ldexp(1.0, -2 * bits), C, ldc);
            g_last_ops += 2.0 * nm * (double)ntiles * (double)CT * CT * (double)kcpad;
            h->other_launches += 3;
            continue;
        }
        i8_split_kernel<<<(unsigned)((items + 255) / 256), 256, 0, h->stream>>>(
            A, lda, n, nrows_pad, k0, kc, kcpad, sc, planes, rb_stride, plane_stride, S);
        SyrkArgs a;
        a.planes = planes; a.plane_stride = plane_stride; a.rb_stride = rb_stride;
        a.tiles = d_tiles; a.ntiles = ntiles; a.n = n; a.nkb = kcpad / BKB; a.S = S;
        a.C = C; a.ldc = ldc; a.sc = sc; a.alpha = alpha;
        a.R = nullptr; a.ldr = 0; a.res_stride = 0;
        int rc;
        if (variant == 2) rc = launch_syrk2<false>(h, a, nclusters);
        else switch (cl) {
        case 1: rc = launch_syrk<1>(h, a, nclusters); break;
        case 2: rc = launch_syrk<2>(h, a, nclusters); break;
        case 8: rc = launch_syrk<8>(h, a, nclusters); break;
        default: rc = launch_syrk<4>(h, a, nclusters); break;
        }
        if (rc != ISDF_OK) return rc;
        g_last_ops += 2.0 * (S * (S + 1) / 2) * (double)ntiles * (variant == 2 ? (double)CT * CT : (double)cl * TM * TN) * (double)kcpad;
        h->other_launches += 2;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}
