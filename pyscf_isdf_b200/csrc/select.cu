// Stage 2: interpolation-point selection by greedy diagonally-pivoted Cholesky
// of the pair-density Gram matrix G[g,g'] = (sum_mu phi[mu,g] phi[mu,g'])^2.
//
// The textbook loop (pivot = first maximum of the residual diagonal d; new
// column = (G[:,p] - L[:, :k] L[p, :k]) / sqrt(d_p); d -= column^2 -- the rule
// of pyscf/lib/scipy_helper.py:84-105) re-reads the whole factor every step
// and is 25x off the roofline.  This file runs the SAME arithmetic, blocked:
//
//   * per panel, a candidate set C (the ~1-2k largest residuals) is cut out by
//     a radix threshold search; theta = largest residual outside C;
//   * the panel factorisation runs on C only (T = G[C,C] - Lc^T Lc by GEMM,
//     then a persistent multi-CTA kernel picks pivots one by one).  Residuals
//     only decrease, so while max_C d > theta the candidate argmax IS the
//     global argmax: pivots are exact, not approximate.  The panel ends when
//     that certificate fails or NB pivots are found;
//   * the NB new columns for all grid points are then two GEMMs (pair density
//     with fused square, and the L L^T update streaming the factor once) and a
//     per-point triangular sweep.
//
// Every L entry is one FMA chain over ascending column index starting from
// G[g,p], exactly as in the unblocked loop, so pivots and factor are
// bit-identical to the C oracle (oracle/c/isdf_oracle.c).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <vector>
#include <algorithm>
#include <cmath>

namespace {

constexpr int NB = 128;          // max pivots per panel
constexpr int MCAP = 4096;       // candidate capacity
constexpr int PANEL_CTAS = 16;
constexpr int PANEL_THREADS = 256;

struct SelState {
    unsigned long long prefix;      // radix search: fixed high bits
    long long remaining;            // candidate quota left for the current prefix
    unsigned long long tie_key;     // candidates = {key > tie_key} + first tie_take of {key == tie_key}
    long long tie_take;
    unsigned long long theta_key;   // largest key outside the candidate set
    int all_fit;
    int m;                          // number of candidates
    int nb_eff;                     // pivots accepted by the last panel
    int stop;                       // 1: residual <= tol reached
    double tol;
    double dmax0;
    unsigned int barrier;           // grid barrier counter for the panel kernel
    int pad;
};

__device__ __forceinline__ unsigned long long dkey(double d)
{
    return d > 0.0 ? (unsigned long long)__double_as_longlong(d) : 0ull;
}

// d[g] = (sum_mu phi[mu,g]^2)^2, FMA chain over ascending mu starting from 0
__global__ void init_diag_kernel(const double* __restrict__ phi, long long ldphi, int nao,
                                 long long ng, double* __restrict__ d)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ng) return;
    double s = 0.0;
    for (int mu = 0; mu < nao; ++mu) {
        const double v = phi[(long long)mu * ldphi + g];
        s = fma(v, v, s);
    }
    d[g] = s * s;
}

__global__ void state_init_kernel(const double* __restrict__ d, long long ng, double tol, SelState* st)
{
    __shared__ double red[32];
    double mx = 0.0;
    for (long long g = threadIdx.x; g < ng; g += blockDim.x) mx = fmax(mx, d[g]);
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmax(mx, red[w]);
        st->dmax0 = mx;
        st->tol = tol < 0.0 ? (double)ng * 2.220446049250313e-16 * mx : tol;
        st->stop = 0;
        st->nb_eff = 0;
        st->m = 0;
        st->barrier = 0;
    }
}

// ---- radix threshold search ------------------------------------------------
__global__ void radix_begin_kernel(SelState* st, long long quota)
{
    st->prefix = 0; st->remaining = quota; st->all_fit = 0;
    st->tie_key = 0; st->tie_take = 0;
}

__global__ void radix_hist_kernel(const double* __restrict__ d, long long ng, const SelState* st,
                                  int shift, int bits, unsigned int* __restrict__ hist)
{
    __shared__ unsigned int sh[2048];
    if (st->all_fit) return;
    const int nbins = 1 << bits;
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const unsigned long long prefix = st->prefix;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += stride) {
        const unsigned long long k = dkey(d[g]);
        if (k != 0 && (k >> (shift + bits)) == prefix) atomicAdd(&sh[(k >> shift) & (nbins - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nbins; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

__global__ void radix_scan_kernel(SelState* st, int shift, int bits, unsigned int* hist, int last)
{
    if (threadIdx.x != 0) return;
    const int nbins = 1 << bits;
    if (!st->all_fit) {
        long long cum = 0;
        int ob = -1;
        for (int b = nbins - 1; b >= 0; --b) {
            if (cum + (long long)hist[b] > st->remaining) { ob = b; break; }
            cum += hist[b];
        }
        if (ob < 0) {
            // everything under this prefix fits (can only happen on the first pass)
            st->all_fit = 1;
            st->tie_key = 0; st->tie_take = 0;
        } else {
            st->remaining -= cum;
            st->prefix = (st->prefix << bits) | (unsigned long long)ob;
            if (last) { st->tie_key = st->prefix; st->tie_take = st->remaining; }
        }
    }
    for (int b = 0; b < nbins; ++b) hist[b] = 0;
}

// ---- ordered compaction of the candidate set (single CTA, ascending g) --------
__global__ void __launch_bounds__(1024) compact_kernel(const double* __restrict__ d, long long ng,
                                                       SelState* st, int* __restrict__ cand,
                                                       double* __restrict__ dC)
{
    __shared__ int s_w[32][2];
    __shared__ long long s_base[2];
    __shared__ unsigned long long s_theta[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned long long tie_key = st->tie_key;
    const long long tie_take = st->tie_take;
    if (tid == 0) { s_base[0] = 0; s_base[1] = 0; }
    __syncthreads();
    unsigned long long theta = 0;
    for (long long g0 = 0; g0 < ng; g0 += blockDim.x) {
        const long long g = g0 + tid;
        unsigned long long k = 0;
        if (g < ng) k = dkey(d[g]);
        const bool gt = k > tie_key;
        const bool eq = (k == tie_key) && (k != 0);
        const unsigned bgt = __ballot_sync(0xffffffffu, gt), beq = __ballot_sync(0xffffffffu, eq);
        if (lane == 0) { s_w[wid][0] = __popc(bgt); s_w[wid][1] = __popc(beq); }
        __syncthreads();
        int ogt = 0, oeq = 0, tgt = 0, teq = 0;
        for (int w = 0; w < 32; ++w) {
            if (w < wid) { ogt += s_w[w][0]; oeq += s_w[w][1]; }
            tgt += s_w[w][0]; teq += s_w[w][1];
        }
        const long long eq_rank = s_base[1] + oeq + __popc(beq & ((1u << lane) - 1u));
        const bool take_eq = eq && (eq_rank < tie_take);
        // position: candidates are emitted in ascending g; ties taken and greater keys interleave,
        // so count both kinds that precede this thread.
        const unsigned btk = __ballot_sync(0xffffffffu, take_eq);
        // number of taken ties in earlier warps of this chunk
        long long tk_before = 0;
        {
            // ties in earlier warps that are within quota: ranks are contiguous, so
            // taken = clamp(tie_take - base_eq - 0, 0, oeq)
            long long room = tie_take - s_base[1];
            if (room < 0) room = 0;
            tk_before = oeq < room ? oeq : room;
        }
        const long long pos = s_base[0] + ogt + tk_before + __popc(bgt & ((1u << lane) - 1u)) +
                              __popc(btk & ((1u << lane) - 1u));
        if (gt || take_eq) {
            if (pos < MCAP) { cand[pos] = (int)g; dC[pos] = d[g]; }
        } else if (k != 0) {
            theta = k > theta ? k : theta;
        }
        __syncthreads();
        if (tid == 0) {
            long long room = tie_take - s_base[1];
            if (room < 0) room = 0;
            const long long tk = teq < room ? teq : room;
            s_base[0] += tgt + tk;
            s_base[1] += teq;
        }
        __syncthreads();
    }
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, theta, o);
        theta = t > theta ? t : theta;
    }
    if (lane == 0) s_theta[wid] = theta;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 32; ++w) theta = s_theta[w] > theta ? s_theta[w] : theta;
        st->theta_key = theta;
        st->m = (int)(s_base[0] < MCAP ? s_base[0] : MCAP);
    }
}

// out[r, j] = in[r, idx[j]]  (idx on device, int32)
__global__ void gather_cols_kernel(const double* __restrict__ in, long long ldin, int nrows,
                                   const int* __restrict__ idx, int ncols,
                                   double* __restrict__ out, long long ldout)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncols) return;
    const long long src = idx[j];
    for (int r = blockIdx.y; r < nrows; r += gridDim.y)
        out[(long long)r * ldout + j] = in[(long long)r * ldin + src];
}

__global__ void gather_cols_i64_kernel(const double* __restrict__ in, long long ldin, int nrows,
                                       const long long* __restrict__ idx, int ncols,
                                       double* __restrict__ out, long long ldout)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncols) return;
    const long long src = idx[j];
    for (int r = blockIdx.y; r < nrows; r += gridDim.y)
        out[(long long)r * ldout + j] = in[(long long)r * ldin + src];
}


struct Partial { double v; int gidx; int cidx; };

__device__ __forceinline__ bool better(double v, int g, double bv, int bg)
{
    return (v > bv) || (v == bv && g < bg);
}

__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int nblocks, unsigned int& epoch)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch += 1;
        __threadfence();
        atomicAdd(counter, 1u);
        const unsigned int target = epoch * nblocks;
        while (*((volatile unsigned int*)counter) < target) { }
        __threadfence();
    }
    __syncthreads();
}

// Panel factorisation on the candidate set.  Launched cooperatively
// (PANEL_CTAS co-resident CTAs); one grid barrier per pivot.
__global__ void __launch_bounds__(PANEL_THREADS)
panel_kernel(SelState* st, const double* __restrict__ T, long long ldT,
             const int* __restrict__ cand, double* __restrict__ dC,
             double* __restrict__ LppT, int* __restrict__ piv_local,
             double* __restrict__ resid_out, Partial* __restrict__ partials,
             int k0, int nbmax)
{
    __shared__ double s_row[NB];
    __shared__ Partial s_part[PANEL_THREADS / 32];
    __shared__ Partial s_best;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nthr = gridDim.x * blockDim.x;
    const int gtid = blockIdx.x * blockDim.x + tid;
    const int m = st->m;
    const double tol = st->tol;
    const double theta = st->theta_key ? __longlong_as_double((long long)st->theta_key) : 0.0;
    unsigned int epoch = 0;
    int kk = 0;
    int stop = 0;
    for (; kk < nbmax; ++kk) {
        // ---- argmax over owned candidates (ties -> lowest grid index) ----
        double bv = -1e300; int bg = 0x7fffffff, bc = -1;
        for (int c = gtid; c < m; c += nthr) {
            const double v = dC[c]; const int g = cand[c];
            if (better(v, g, bv, bg)) { bv = v; bg = g; bc = c; }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int og = __shfl_xor_sync(0xffffffffu, bg, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (better(ov, og, bv, bg)) { bv = ov; bg = og; bc = oc; }
        }
        if (lane == 0) { s_part[wid].v = bv; s_part[wid].gidx = bg; s_part[wid].cidx = bc; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < PANEL_THREADS / 32; ++w)
                if (better(s_part[w].v, s_part[w].gidx, bv, bg)) { bv = s_part[w].v; bg = s_part[w].gidx; bc = s_part[w].cidx; }
            Partial* out = partials + (kk & 1) * PANEL_CTAS + blockIdx.x;
            out->v = bv; out->gidx = bg; out->cidx = bc;
        }
        grid_barrier(&st->barrier, gridDim.x, epoch);
        if (tid == 0) {
            const Partial* in = partials + (kk & 1) * PANEL_CTAS;
            double v = in[0].v; int g = in[0].gidx, c = in[0].cidx;
            for (int b = 1; b < (int)gridDim.x; ++b)
                if (better(in[b].v, in[b].gidx, v, g)) { v = in[b].v; g = in[b].gidx; c = in[b].cidx; }
            s_best.v = v; s_best.gidx = g; s_best.cidx = c;
        }
        __syncthreads();
        const double sv = s_best.v; const int sc = s_best.cidx;
        if (sc < 0 || !(sv > tol)) { stop = 1; break; }
        if (kk > 0 && !(sv > theta)) break;       // certificate failed: end the panel
        const double piv = sqrt(sv);
        for (int j = tid; j < kk; j += blockDim.x) s_row[j] = LppT[(long long)j * MCAP + sc];
        __syncthreads();
        const double* Trow = T + (long long)sc * ldT;
        for (int c = gtid; c < m; c += nthr) {
            double acc = Trow[c];
            for (int j = 0; j < kk; ++j) acc = fma(-LppT[(long long)j * MCAP + c], s_row[j], acc);
            const double L = acc / piv;
            LppT[(long long)kk * MCAP + c] = L;
            double dn = fma(-L, L, dC[c]);
            if (c == sc) dn = -1.0;
            dC[c] = dn;
        }
        if (gtid == 0) { piv_local[kk] = sc; resid_out[k0 + kk] = sv; }
        __syncthreads();
    }
    if (gtid == 0) { st->nb_eff = kk; st->stop = stop; }
}

// LppP[k][j] = LppT[j][piv_local[k]]  (within-panel factor rows of the chosen pivots)
__global__ void gather_lpp_kernel(const double* __restrict__ LppT, const int* __restrict__ piv_local,
                                  int nb, double* __restrict__ LppP)
{
    const int k = blockIdx.x, j = threadIdx.x;
    if (k < nb && j < nb) LppP[k * NB + j] = (j < k) ? LppT[(long long)j * MCAP + piv_local[k]] : 0.0;
}

__global__ void record_pivots_kernel(const int* __restrict__ cand, const int* __restrict__ piv_local,
                                     int nb, int k0, long long* __restrict__ ip_dev)
{
    const int k = threadIdx.x;
    if (k < nb) ip_dev[k0 + k] = cand[piv_local[k]];
}

__global__ void mark_selected_kernel(const long long* __restrict__ ip_dev, int k0, int nb, double* __restrict__ d)
{
    const int k = threadIdx.x;
    if (k < nb) d[ip_dev[k0 + k]] = -1.0;
}

// Per-grid-point triangular sweep over the panel columns + residual update.
// rows: Lt_panel[k * ldl + g], k < nb.  Divisor of column k is sqrt(resid[k]).
__global__ void __launch_bounds__(128)
panel_sweep_kernel(double* __restrict__ Ltp, long long ldl, long long ng, int nb,
                   const double* __restrict__ LppP, const double* __restrict__ resid,
                   double* __restrict__ d)
{
    extern __shared__ double s_l[];            // packed lower triangle, then nb pivots
    double* s_piv = s_l + NB * (NB + 1) / 2;
    for (int i = threadIdx.x; i < nb * nb; i += blockDim.x) {
        const int k = i / nb, j = i - k * nb;
        if (j < k) s_l[k * (k + 1) / 2 + j] = LppP[k * NB + j];
    }
    for (int i = threadIdx.x; i < nb; i += blockDim.x) s_piv[i] = sqrt(resid[i]);
    __syncthreads();
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ng) return;
    double dd = d[g];
    double* col = Ltp + g;
    for (int kb = 0; kb < nb; kb += 8) {
        double acc[8], res[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = (kb + i < nb) ? col[(long long)(kb + i) * ldl] : 0.0;
        for (int j = 0; j < kb; ++j) {
            const double x = -col[(long long)j * ldl];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (kb + i < nb) acc[i] = fma(x, s_l[(kb + i) * (kb + i + 1) / 2 + j], acc[i]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (kb + i < nb) {
#pragma unroll
                for (int jj = 0; jj < i; ++jj)
                    acc[i] = fma(-res[jj], s_l[(kb + i) * (kb + i + 1) / 2 + kb + jj], acc[i]);
                res[i] = acc[i] / s_piv[kb + i];
                col[(long long)(kb + i) * ldl] = res[i];
                dd = fma(-res[i], res[i], dd);
            }
        }
    }
    d[g] = dd;
}

struct SelWs {
    double* d; unsigned int* hist; int* cand; double* dC; double* PhiC; double* LcOld; double* T;
    double* LppT; int* piv_local; double* PhiP; double* LtP; double* LppP; double* resid_dev;
    long long* ip_dev; Partial* partials; SelState* st;
};

size_t carve(SelWs& w, void* ws, size_t bytes, int nao, long long ng, int npts)
{
    WsAlloc a(ws, bytes);
    w.d = a.get<double>(ng);
    w.hist = a.get<unsigned int>(2048);
    w.cand = a.get<int>(MCAP);
    w.dC = a.get<double>(MCAP);
    w.PhiC = a.get<double>((size_t)nao * MCAP);
    w.LcOld = a.get<double>((size_t)npts * MCAP);
    w.T = a.get<double>((size_t)MCAP * MCAP);
    w.LppT = a.get<double>((size_t)NB * MCAP);
    w.piv_local = a.get<int>(NB);
    w.PhiP = a.get<double>((size_t)nao * NB);
    w.LtP = a.get<double>((size_t)npts * NB);
    w.LppP = a.get<double>((size_t)NB * NB);
    w.resid_dev = a.get<double>(npts);
    w.ip_dev = a.get<long long>(npts);
    w.partials = a.get<Partial>(2 * PANEL_CTAS);
    w.st = a.get<SelState>(1);
    return a.off + 256;
}

}  // namespace

static int gather_dev(isdf_handle* h, const double* in, long long ldin, int nrows,
                      const int* idx, int ncols, double* out, long long ldout)
{
    if (nrows <= 0 || ncols <= 0) return ISDF_OK;
    dim3 grid((ncols + 127) / 128, std::min(nrows, 4096));
    gather_cols_kernel<<<grid, 128, 0, h->stream>>>(in, ldin, nrows, idx, ncols, out, ldout);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}

extern "C" int isdf_gather_cols(isdf_handle* h, const double* in, long long ldin, int nrows,
                                const long long* h_idx, int ncols, double* out, long long ldout)
{
    ARG_CHECK(h && in && h_idx && out && nrows >= 0 && ncols >= 0);
    if (nrows == 0 || ncols == 0) return ISDF_OK;
    long long* d_idx = nullptr;
    CUDA_TRY(cudaMallocAsync((void**)&d_idx, sizeof(long long) * ncols, h->stream));
    CUDA_TRY(cudaMemcpyAsync(d_idx, h_idx, sizeof(long long) * ncols, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));   // h_idx may be pageable
    dim3 grid((ncols + 127) / 128, std::min(nrows, 4096));
    gather_cols_i64_kernel<<<grid, 128, 0, h->stream>>>(in, ldin, nrows, d_idx, ncols, out, ldout);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    CUDA_TRY(cudaFreeAsync(d_idx, h->stream));
    return ISDF_OK;
}

extern "C" size_t isdf_select_ip_ws_bytes(int nao, long long ng, int npts)
{
    SelWs w;
    return carve(w, nullptr, 0, nao, ng, npts);
}

static int gemm_simple(isdf_handle* h, int M, int N, int K, const double* A, long long lda,
                       const double* B, long long ldb, double* C, long long ldc,
                       const double* Cin, int neg, int epi)
{
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.A = A; p.lda = lda; p.a_kmajor = 1;
    p.B = B; p.ldb = ldb; p.b_kmajor = 1;
    p.C = C; p.ldc = ldc; p.Cin = Cin; p.ldcin = ldc;
    p.E = nullptr; p.lde = 0; p.alpha = 1.0; p.neg = neg; p.epi = epi; p.upper_only = 0;
    return gemm_launch(h, p);
}

extern "C" int isdf_select_ip(isdf_handle* h, const double* phi, int nao, long long ng, long long ldphi,
                              int npts, double tol, double* Lt, long long ldl,
                              long long* h_ip, double* h_resid, int* h_rank,
                              void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi && Lt && h_ip && h_rank && ws);
    ARG_CHECK(nao > 0 && ng > 0 && ng < 0x7fffffffLL && npts > 0 && npts <= ng && ldphi >= ng && ldl >= ng);
    SelWs w;
    const size_t need = carve(w, ws, ws_bytes, nao, ng, npts);
    if (need > ws_bytes) { isdf_set_error("select_ip workspace too small: need %zu have %zu", need, ws_bytes); return ISDF_ERR_WS; }
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_SELECT], s));

    init_diag_kernel<<<(unsigned)((ng + 255) / 256), 256, 0, s>>>(phi, ldphi, nao, ng, w.d);
    state_init_kernel<<<1, 1024, 0, s>>>(w.d, ng, tol, w.st);
    CUDA_TRY(cudaMemsetAsync(w.hist, 0, 2048 * sizeof(unsigned int), s));
    CUDA_TRY(cudaGetLastError());
    h->other_launches += 2;

    static bool sweep_cfg = false;
    const size_t sweep_smem = sizeof(double) * (NB * (NB + 1) / 2 + NB);
    if (!sweep_cfg) {
        CUDA_TRY(cudaFuncSetAttribute(panel_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_smem));
        sweep_cfg = true;
    }

    SelState* hst = (SelState*)h->h_pinned;
    const int shifts[6] = {52, 41, 30, 20, 10, 0};
    const int bitsv[6] = {11, 11, 11, 10, 10, 10};
    long long quota = 1024;
    int k0 = 0;
    const int hist_blocks = std::min<long long>((ng + 255) / 256, 2 * h->num_sms);
    while (k0 < npts) {
        const int nbmax = std::min(NB, npts - k0);
        // ---- candidate set ----
        radix_begin_kernel<<<1, 1, 0, s>>>(w.st, quota);
        for (int pass = 0; pass < 6; ++pass) {
            radix_hist_kernel<<<hist_blocks, 256, 0, s>>>(w.d, ng, w.st, shifts[pass], bitsv[pass], w.hist);
            radix_scan_kernel<<<1, 32, 0, s>>>(w.st, shifts[pass], bitsv[pass], w.hist, pass == 5);
        }
        compact_kernel<<<1, 1024, 0, s>>>(w.d, ng, w.st, w.cand, w.dC);
        CUDA_TRY(cudaGetLastError());
        h->other_launches += 14;
        CUDA_TRY(cudaMemcpyAsync(hst, w.st, sizeof(SelState), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        const int m = hst->m;
        if (m <= 0) break;                       // no positive residual left
        // ---- T = G[C,C] - Lc^T Lc ----
        int rc;
        if ((rc = gather_dev(h, phi, ldphi, nao, w.cand, m, w.PhiC, MCAP))) return rc;
        if ((rc = gather_dev(h, Lt, ldl, k0, w.cand, m, w.LcOld, MCAP))) return rc;
        if ((rc = gemm_simple(h, m, m, nao, w.PhiC, MCAP, w.PhiC, MCAP, w.T, MCAP, nullptr, 0, EPI_SQUARE))) return rc;
        if (k0 > 0)
            if ((rc = gemm_simple(h, m, m, k0, w.LcOld, MCAP, w.LcOld, MCAP, w.T, MCAP, w.T, 1, EPI_STORE))) return rc;
        // ---- panel factorisation on the candidates ----
        {
            SelState* st = w.st; const double* T = w.T; long long ldT = MCAP; const int* cand = w.cand;
            double* dC = w.dC; double* LppT = w.LppT; int* pl = w.piv_local; double* rs = w.resid_dev;
            Partial* pa = w.partials; int k0a = k0, nbm = nbmax;
            void* args[] = {&st, &T, &ldT, &cand, &dC, &LppT, &pl, &rs, &pa, &k0a, &nbm};
            CUDA_TRY(cudaMemsetAsync(&w.st->barrier, 0, sizeof(unsigned int), s));
            CUDA_TRY(cudaLaunchCooperativeKernel((void*)panel_kernel, dim3(PANEL_CTAS), dim3(PANEL_THREADS), args, 0, s));
            h->other_launches++;
        }
        CUDA_TRY(cudaMemcpyAsync(hst, w.st, sizeof(SelState), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        const int nb = hst->nb_eff;
        const int stop = hst->stop;
        if (nb > 0) {
            // ---- flush: new factor rows for every grid point ----
            record_pivots_kernel<<<1, NB, 0, s>>>(w.cand, w.piv_local, nb, k0, w.ip_dev);
            gather_lpp_kernel<<<nb, NB, 0, s>>>(w.LppT, w.piv_local, nb, w.LppP);
            h->other_launches += 2;
            if ((rc = gather_dev(h, w.PhiC, MCAP, nao, w.piv_local, nb, w.PhiP, NB))) return rc;
            if ((rc = gather_dev(h, w.LcOld, MCAP, k0, w.piv_local, nb, w.LtP, NB))) return rc;
            double* Ltp = Lt + (long long)k0 * ldl;
            if ((rc = gemm_simple(h, nb, (int)ng, nao, w.PhiP, NB, phi, ldphi, Ltp, ldl, nullptr, 0, EPI_SQUARE))) return rc;
            if (k0 > 0)
                if ((rc = gemm_simple(h, nb, (int)ng, k0, w.LtP, NB, Lt, ldl, Ltp, ldl, Ltp, 1, EPI_STORE))) return rc;
            panel_sweep_kernel<<<(unsigned)((ng + 127) / 128), 128, sweep_smem, s>>>(
                Ltp, ldl, ng, nb, w.LppP, w.resid_dev + k0, w.d);
            mark_selected_kernel<<<1, NB, 0, s>>>(w.ip_dev, k0, nb, w.d);
            CUDA_TRY(cudaGetLastError());
            h->other_launches += 2;
            k0 += nb;
        }
        if (stop) break;
        // adapt the candidate quota: grow when the certificate ends panels early
        if (nb < nbmax / 2 && quota < MCAP) quota *= 2;
        else if (nb == nbmax && quota > 1024 && m > 8 * nbmax) quota = quota;   // keep
    }
    CUDA_TRY(cudaMemcpyAsync(h_ip, w.ip_dev, sizeof(long long) * k0, cudaMemcpyDeviceToHost, s));
    if (h_resid) CUDA_TRY(cudaMemcpyAsync(h_resid, w.resid_dev, sizeof(double) * k0, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_SELECT], s));
    h->ev_used[ISDF_STAGE_SELECT] = true;
    CUDA_TRY(cudaStreamSynchronize(s));
    *h_rank = k0;
    return ISDF_OK;
}
