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
#include <cstdlib>
#include <cooperative_groups.h>

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

// one CTA of 1024 threads: suffix sums over <= 2048 bins, then the owning bin is located in parallel
__global__ void __launch_bounds__(1024) radix_scan_kernel(SelState* st, int shift, int bits, unsigned int* hist, int last)
{
    __shared__ long long suf[2049];      // suf[b] = sum_{b' >= b} hist[b']
    __shared__ int s_ob;
    const int nbins = 1 << bits;
    const int tid = threadIdx.x;
    if (st->all_fit) {
        for (int b = tid; b < nbins; b += blockDim.x) hist[b] = 0;
        return;
    }
    for (int b = tid; b < 2049; b += blockDim.x) suf[b] = (b < nbins) ? (long long)hist[b] : 0;
    if (tid == 0) s_ob = -1;
    __syncthreads();
    for (int off = 1; off < nbins; off <<= 1) {
        long long v[2];
        for (int i = 0; i < 2; ++i) {
            const int b = tid + i * 1024;
            v[i] = (b < nbins) ? suf[b] + ((b + off < nbins) ? suf[b + off] : 0) : 0;
        }
        __syncthreads();
        for (int i = 0; i < 2; ++i) { const int b = tid + i * 1024; if (b < nbins) suf[b] = v[i]; }
        __syncthreads();
    }
    const long long remaining = st->remaining;
    // ob = largest b with suf[b] > remaining (the bin where the quota is crossed)
    for (int b = tid; b < nbins; b += blockDim.x)
        if (suf[b] > remaining && suf[b + 1] <= remaining) s_ob = b;
    __syncthreads();
    if (tid == 0) {
        const int ob = s_ob;
        if (ob < 0) {
            // everything under this prefix fits (can only happen on the first pass)
            st->all_fit = 1;
            st->tie_key = 0; st->tie_take = 0;
        } else {
            st->remaining -= suf[ob + 1];
            st->prefix = (st->prefix << bits) | (unsigned long long)ob;
            if (last) { st->tie_key = st->prefix; st->tie_take = st->remaining; }
        }
    }
    __syncthreads();
    for (int b = tid; b < nbins; b += blockDim.x) hist[b] = 0;
}

// ---- ordered compaction of the candidate set --------------------------------------------
// candidates = {key > tie_key} + the first tie_take of {key == tie_key} in ascending g.
// Three launches: per-chunk counts, a scan over the chunks, per-chunk scatter.
constexpr int COMPACT_CHUNK = 8192;

template <bool SCATTER>
__global__ void __launch_bounds__(1024)
compact_chunk_kernel(const double* __restrict__ d, long long ng, const SelState* __restrict__ st,
                     long long* __restrict__ chunk_cnt, const long long* __restrict__ chunk_base,
                     unsigned long long* __restrict__ chunk_theta,
                     int* __restrict__ cand, double* __restrict__ dC)
{
    __shared__ int s_w[32][2];
    __shared__ long long s_base[2];
    __shared__ unsigned long long s_theta[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned long long tie_key = st->tie_key;
    const long long tie_take = st->tie_take;
    const long long g_begin = (long long)blockIdx.x * COMPACT_CHUNK;
    const long long g_end = min(ng, g_begin + COMPACT_CHUNK);
    if (tid == 0) {
        s_base[0] = SCATTER ? chunk_base[2 * blockIdx.x] : 0;
        s_base[1] = SCATTER ? chunk_base[2 * blockIdx.x + 1] : 0;
    }
    __syncthreads();
    unsigned long long theta = 0;       // largest key below tie_key in this chunk (count pass)
    for (long long g0 = g_begin; g0 < g_end; g0 += blockDim.x) {
        const long long g = g0 + tid;
        unsigned long long k = 0;
        if (g < g_end) k = dkey(d[g]);
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
        if (SCATTER) {
            const long long eq_rank = s_base[1] + oeq + __popc(beq & ((1u << lane) - 1u));
            const bool take_eq = eq && (eq_rank < tie_take);
            const unsigned btk = __ballot_sync(0xffffffffu, take_eq);
            // ties of earlier warps within the quota: ranks are contiguous
            long long room = tie_take - s_base[1];
            if (room < 0) room = 0;
            const long long tk_before = oeq < room ? oeq : room;
            const long long pos = s_base[0] + ogt + tk_before + __popc(bgt & ((1u << lane) - 1u)) +
                                  __popc(btk & ((1u << lane) - 1u));
            if ((gt || take_eq) && pos < MCAP) { cand[pos] = (int)g; dC[pos] = d[g]; }
        } else if (!gt && !eq && k != 0) {
            theta = k > theta ? k : theta;
        }
        __syncthreads();
        if (tid == 0) {
            long long room = tie_take - s_base[1];
            if (room < 0) room = 0;
            const long long tk = teq < room ? teq : room;
            s_base[0] += tgt + (SCATTER ? tk : 0);
            s_base[1] += teq;
        }
        __syncthreads();
    }
    if (!SCATTER) {
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long t = __shfl_xor_sync(0xffffffffu, theta, o);
            theta = t > theta ? t : theta;
        }
        if (lane == 0) s_theta[wid] = theta;
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 32; ++w) theta = s_theta[w] > theta ? s_theta[w] : theta;
            chunk_cnt[2 * blockIdx.x] = s_base[0];          // keys above the threshold
            chunk_cnt[2 * blockIdx.x + 1] = s_base[1];      // ties
            chunk_theta[blockIdx.x] = theta;
        }
    }
}

// exclusive scan over the chunks: output position base and tie rank base of every chunk
__global__ void compact_scan_kernel(SelState* st, int nchunks, const long long* __restrict__ chunk_cnt,
                                    long long* __restrict__ chunk_base,
                                    const unsigned long long* __restrict__ chunk_theta)
{
    if (threadIdx.x != 0) return;
    const long long tie_take = st->tie_take;
    long long pos = 0, eqs = 0;
    unsigned long long theta = 0;
    for (int c = 0; c < nchunks; ++c) {
        chunk_base[2 * c] = pos;
        chunk_base[2 * c + 1] = eqs;
        long long room = tie_take - eqs;
        if (room < 0) room = 0;
        const long long teq = chunk_cnt[2 * c + 1];
        pos += chunk_cnt[2 * c] + (teq < room ? teq : room);
        eqs += teq;
        theta = chunk_theta[c] > theta ? chunk_theta[c] : theta;
    }
    if (eqs > tie_take) theta = st->tie_key;        // an untaken tie is the largest residual outside the set
    st->theta_key = theta;
    st->m = (int)(pos < MCAP ? pos : MCAP);
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
             int k0, int nbmax, int use_smem)
{
    extern __shared__ double s_lpp[];          // [NB][cpb]: this CTA's columns of the panel factor (use_smem)
    __shared__ double s_row[NB];
    __shared__ Partial s_part[PANEL_THREADS / 32];
    __shared__ Partial s_best;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nthr = gridDim.x * blockDim.x;
    const int gtid = blockIdx.x * blockDim.x + tid;
    const int m = st->m;
    // candidates are dealt evenly over the CTAs (a contiguous block of cpb per CTA, one per thread):
    // with m = 1024 a linear thread index would leave 12 of the 16 CTAs without work
    const int cpb = (((m + (int)gridDim.x - 1) / (int)gridDim.x) + 31) & ~31;
    const int myc = (tid < cpb && blockIdx.x * cpb + tid < m) ? blockIdx.x * cpb + tid : -1;
    const double tol = st->tol;
    const double theta = st->theta_key ? __longlong_as_double((long long)st->theta_key) : 0.0;
    unsigned int epoch = 0;
    int kk = 0;
    int stop = 0;
    for (; kk < nbmax; ++kk) {
        // ---- argmax over owned candidates (ties -> lowest grid index) ----
        double bv = -1e300; int bg = 0x7fffffff, bc = -1;
        if (myc >= 0) {
            const int c = myc;
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
        if (myc >= 0) {
            const int c = myc;
            double acc = Trow[c];
            if (use_smem)
                for (int j = 0; j < kk; ++j) acc = fma(-s_lpp[j * cpb + tid], s_row[j], acc);
            else
                for (int j = 0; j < kk; ++j) acc = fma(-LppT[(long long)j * MCAP + c], s_row[j], acc);
            const double L = acc / piv;
            LppT[(long long)kk * MCAP + c] = L;
            if (use_smem) s_lpp[kk * cpb + tid] = L;
            double dn = fma(-L, L, dC[c]);
            if (c == sc) dn = -1.0;
            dC[c] = dn;
        }
        if (gtid == 0) { piv_local[kk] = sc; resid_out[k0 + kk] = sv; }
        __syncthreads();
    }
    if (gtid == 0) { st->nb_eff = kk; st->stop = stop; }
}

// Same panel factorisation with the PANEL_CTAS CTAs launched as ONE thread-block cluster
// (opt-in cluster size 16): the per-pivot exchange goes through distributed shared memory
// and the hardware cluster barrier instead of global memory and an atomic spin barrier.
//   * every CTA pushes its local best into slot [parity][rank] of EVERY CTA's table, so after
//     the barrier the reduction over CTAs is a local shared-memory loop that all threads run
//     redundantly (same data, same order -> same winner in every CTA);
//   * the winner's earlier panel columns (what the FMA chain multiplies by) are read from the
//     owning CTA's shared-memory copy of its factor columns (DSMEM) -- needs the shared-memory
//     layout, i.e. cpb <= 192 candidates per CTA (the host falls back to panel_kernel otherwise);
//   * residuals of the owned candidate live in a register.
// The arithmetic (argmax rule, tie-break, FMA chain order, division) is that of panel_kernel,
// statement for statement, so the pivots and LppT are bit-identical.  Opt-in: ISDF_PANEL_CLUSTER=1.
__global__ void __launch_bounds__(PANEL_THREADS)
panel_cluster_kernel(SelState* st, const double* __restrict__ T, long long ldT,
                     const int* __restrict__ cand, double* __restrict__ dC,
                     double* __restrict__ LppT, int* __restrict__ piv_local,
                     double* __restrict__ resid_out, int k0, int nbmax)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ double s_lpp[];          // [NB][cpb]: this CTA's columns of the panel factor
    __shared__ double s_row[NB];
    __shared__ Partial s_part[PANEL_THREADS / 32];
    __shared__ Partial s_all[2][PANEL_CTAS];   // [parity][cluster rank]: filled remotely by every CTA
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned int crank = cluster.block_rank();          // == blockIdx.x (one cluster spans the grid)
    const int m = st->m;
    const int cpb = (((m + PANEL_CTAS - 1) / PANEL_CTAS) + 31) & ~31;
    const int myc = (tid < cpb && (int)crank * cpb + tid < m) ? (int)crank * cpb + tid : -1;
    const double tol = st->tol;
    const double theta = st->theta_key ? __longlong_as_double((long long)st->theta_key) : 0.0;
    double myd = myc >= 0 ? dC[myc] : 0.0;
    const int myg = myc >= 0 ? cand[myc] : 0x7fffffff;
    int kk = 0;
    int stop = 0;
    cluster.sync();                              // every CTA of the cluster is resident before the first remote store
    for (; kk < nbmax; ++kk) {
        // ---- argmax over owned candidates (ties -> lowest grid index) ----
        double bv = -1e300; int bg = 0x7fffffff, bc = -1;
        if (myc >= 0 && better(myd, myg, bv, bg)) { bv = myd; bg = myg; bc = myc; }
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int og = __shfl_xor_sync(0xffffffffu, bg, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (better(ov, og, bv, bg)) { bv = ov; bg = og; bc = oc; }
        }
        if (lane == 0) { s_part[wid].v = bv; s_part[wid].gidx = bg; s_part[wid].cidx = bc; }
        __syncthreads();
        if (wid == 0) {
            // every lane folds the warp partials in the same order, lane r < PANEL_CTAS pushes the
            // CTA's best into CTA r's table
            bv = s_part[0].v; bg = s_part[0].gidx; bc = s_part[0].cidx;
            for (int w = 1; w < PANEL_THREADS / 32; ++w)
                if (better(s_part[w].v, s_part[w].gidx, bv, bg)) { bv = s_part[w].v; bg = s_part[w].gidx; bc = s_part[w].cidx; }
            if (lane < PANEL_CTAS) {
                Partial* remote = cluster.map_shared_rank(&s_all[kk & 1][crank], lane);
                Partial p; p.v = bv; p.gidx = bg; p.cidx = bc;
                *remote = p;
            }
        }
        cluster.sync();                          // release/acquire: the remote stores are visible
        double sv = s_all[kk & 1][0].v; int sg = s_all[kk & 1][0].gidx, sc = s_all[kk & 1][0].cidx;
        for (int b = 1; b < PANEL_CTAS; ++b) {
            const Partial q = s_all[kk & 1][b];
            if (better(q.v, q.gidx, sv, sg)) { sv = q.v; sg = q.gidx; sc = q.cidx; }
        }
        if (sc < 0 || !(sv > tol)) { stop = 1; break; }
        if (kk > 0 && !(sv > theta)) break;       // certificate failed: end the panel
        const double piv = sqrt(sv);
        // the pivot's earlier columns of this panel, from the owning CTA's shared memory
        const int owner = sc / cpb, oc = sc - owner * cpb;
        const double* rl = cluster.map_shared_rank(s_lpp, owner);
        double acc = 0.0;
        if (myc >= 0) acc = T[(long long)sc * ldT + myc];
        for (int j = tid; j < kk; j += blockDim.x) s_row[j] = rl[j * cpb + oc];
        __syncthreads();
        if (myc >= 0) {
            for (int j = 0; j < kk; ++j) acc = fma(-s_lpp[j * cpb + tid], s_row[j], acc);
            const double L = acc / piv;
            LppT[(long long)kk * MCAP + myc] = L;
            s_lpp[kk * cpb + tid] = L;
            double dn = fma(-L, L, myd);
            if (myc == sc) dn = -1.0;
            myd = dn;
        }
        if (crank == 0 && tid == 0) { piv_local[kk] = sc; resid_out[k0 + kk] = sv; }
        // s_row / s_part are rewritten only after the next iteration's __syncthreads / cluster.sync
    }
    if (myc >= 0) dC[myc] = myd;
    if (crank == 0 && tid == 0) { st->nb_eff = kk; st->stop = stop; }
    cluster.sync();                              // no CTA leaves while its shared memory may still be read
}

// ---- third form of the panel kernel: cluster + one-way pushes (ISDF_PANEL_CLUSTER=2) -----------
// cluster.sync() costs a GPU-scope memory barrier per pivot (MEMBAR.ALL.GPU + L1 invalidate in SASS).
// Here nothing is pulled and nothing is fenced: before the exchange a CTA does not know the winner, so
// it pushes its local best (16 bytes) AND that candidate's earlier panel columns (8*kk bytes) into every
// CTA's shared memory with st.async, each store completing bytes on the receiver's mbarrier.  A receiver
// arms the barrier with the byte count it expects (16 senders), waits for the phase, scans the 16
// entries and reads the winner's columns from its own shared memory.  Two buffers / two barriers
// alternate by pivot parity; a sender can run at most one pivot ahead of the slowest CTA (it needs that
// CTA's push of pivot kk+1, issued after the CTA's end-of-pivot __syncthreads), so a buffer is never
// overwritten while it is still being read.  Same arithmetic as panel_kernel, statement for statement.
constexpr int PUSH_CPB_MAX = 160;                 // factor columns [NB][cpb] + push buffers must fit 227 KB

__device__ __forceinline__ unsigned int sel_smem_u32(const void* p) { return (unsigned int)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned int sel_mapa(unsigned int addr, unsigned int rank)
{
    unsigned int r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void sel_st_async_f64(unsigned int raddr, double v, unsigned int rbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];"
                 :: "r"(raddr), "d"(v), "r"(rbar) : "memory");
}
__device__ __forceinline__ void sel_st_async_2xb64(unsigned int raddr, unsigned long long a, unsigned long long b, unsigned int rbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b64 [%0], {%1, %2}, [%3];"
                 :: "r"(raddr), "l"(a), "l"(b), "r"(rbar) : "memory");
}
__device__ __forceinline__ void sel_mbar_wait(unsigned int bar, unsigned int parity)
{
    unsigned int ok = 0;
    const long long t0 = clock64();
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) break;
        if (clock64() - t0 > 4000000000LL) __trap();       // never hang the device on a protocol bug
    }
}

__global__ void __launch_bounds__(PANEL_THREADS)
panel_push_kernel(SelState* st, const double* __restrict__ T, long long ldT,
                  const int* __restrict__ cand, double* __restrict__ dC,
                  double* __restrict__ LppT, int* __restrict__ piv_local,
                  double* __restrict__ resid_out, int k0, int nbmax)
{
    extern __shared__ double s_lpp[];                        // [NB][cpb]: this CTA's columns of the panel factor
    __shared__ __align__(16) Partial s_tab[2][PANEL_CTAS];   // [parity][sender]
    __shared__ __align__(16) double s_rows[2][PANEL_CTAS][NB];   // [parity][sender][earlier panel columns of its best]
    __shared__ __align__(8) unsigned long long s_bar[2];
    __shared__ Partial s_part[PANEL_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned int crank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    const int m = st->m;
    const int cpb = (((m + PANEL_CTAS - 1) / PANEL_CTAS) + 31) & ~31;
    const int myc = (tid < cpb && (int)crank * cpb + tid < m) ? (int)crank * cpb + tid : -1;
    const double tol = st->tol;
    const double theta = st->theta_key ? __longlong_as_double((long long)st->theta_key) : 0.0;
    double myd = myc >= 0 ? dC[myc] : 0.0;
    const int myg = myc >= 0 ? cand[myc] : 0x7fffffff;
    if (tid == 0) {
        for (int i = 0; i < 2; ++i)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(sel_smem_u32(&s_bar[i])), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");     // peers' barriers are initialised
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    int kk = 0;
    int stop = 0;
    for (; kk < nbmax; ++kk) {
        const int par = kk & 1;
        // ---- argmax over owned candidates (ties -> lowest grid index) ----
        double bv = -1e300; int bg = 0x7fffffff, bc = -1;
        if (myc >= 0 && better(myd, myg, bv, bg)) { bv = myd; bg = myg; bc = myc; }
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int og = __shfl_xor_sync(0xffffffffu, bg, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (better(ov, og, bv, bg)) { bv = ov; bg = og; bc = oc; }
        }
        if (lane == 0) { s_part[wid].v = bv; s_part[wid].gidx = bg; s_part[wid].cidx = bc; }
        __syncthreads();
        // every thread folds the warp partials in the same order: the CTA's best
        bv = s_part[0].v; bg = s_part[0].gidx; bc = s_part[0].cidx;
        for (int w = 1; w < PANEL_THREADS / 32; ++w)
            if (better(s_part[w].v, s_part[w].gidx, bv, bg)) { bv = s_part[w].v; bg = s_part[w].gidx; bc = s_part[w].cidx; }
        // ---- push the CTA's best and its earlier panel columns to all CTAs ----
        const unsigned int bar_local = sel_smem_u32(&s_bar[par]);
        if (tid == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                         :: "r"(bar_local), "r"((unsigned int)(PANEL_CTAS * (16 + 8 * kk))) : "memory");
        if (tid < PANEL_CTAS) {
            const unsigned int dst = sel_mapa(sel_smem_u32(&s_tab[par][crank]), (unsigned int)tid);
            const unsigned int rbar = sel_mapa(bar_local, (unsigned int)tid);
            const unsigned long long w0 = (unsigned long long)__double_as_longlong(bv);
            const unsigned long long w1 = ((unsigned long long)(unsigned int)bc << 32) | (unsigned long long)(unsigned int)bg;   // {gidx, cidx}
            sel_st_async_2xb64(dst, w0, w1, rbar);
        }
        {
            const int tb = bc >= 0 ? bc - (int)crank * cpb : 0;
            for (int idx = tid; idx < PANEL_CTAS * kk; idx += PANEL_THREADS) {
                const int r = idx & (PANEL_CTAS - 1), j = idx / PANEL_CTAS;
                const double v = bc >= 0 ? s_lpp[j * cpb + tb] : 0.0;
                sel_st_async_f64(sel_mapa(sel_smem_u32(&s_rows[par][crank][j]), (unsigned int)r), v,
                                 sel_mapa(bar_local, (unsigned int)r));
            }
        }
        sel_mbar_wait(bar_local, (unsigned int)((kk >> 1) & 1));
        // ---- all 16 entries are here: the winner, identically in every CTA ----
        double sv = s_tab[par][0].v; int sg = s_tab[par][0].gidx, sc = s_tab[par][0].cidx, sb = 0;
        for (int b = 1; b < PANEL_CTAS; ++b) {
            const Partial q = s_tab[par][b];
            if (better(q.v, q.gidx, sv, sg)) { sv = q.v; sg = q.gidx; sc = q.cidx; sb = b; }
        }
        if (sc < 0 || !(sv > tol)) { stop = 1; break; }
        if (kk > 0 && !(sv > theta)) break;       // certificate failed: end the panel
        const double piv = sqrt(sv);
        if (myc >= 0) {
            double acc = T[(long long)sc * ldT + myc];
            const double* row = s_rows[par][sb];
            for (int j = 0; j < kk; ++j) acc = fma(-s_lpp[j * cpb + tid], row[j], acc);
            const double L = acc / piv;
            LppT[(long long)kk * MCAP + myc] = L;
            s_lpp[kk * cpb + tid] = L;
            double dn = fma(-L, L, myd);
            if (myc == sc) dn = -1.0;
            myd = dn;
        }
        if (crank == 0 && tid == 0) { piv_local[kk] = sc; resid_out[k0 + kk] = sv; }
        __syncthreads();          // the CTA is done with this pivot's buffers before any of its threads pushes the next
    }
    if (myc >= 0) dC[myc] = myd;
    if (crank == 0 && tid == 0) { st->nb_eff = kk; st->stop = stop; }
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");     // no CTA leaves while peers may still push to it
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// LppP[k][j] = LppT[j][piv_local[k]]  (within-panel factor rows of the chosen pivots)
__global__ void gather_lpp_kernel(const double* __restrict__ LppT, const int* __restrict__ piv_local,
                                  int nb, double* __restrict__ LppP)
{
    const int k = blockIdx.x, j = threadIdx.x;
    if (k < nb && j < nb) LppP[k * NB + j] = (j < k) ? LppT[(long long)j * MCAP + piv_local[k]] : 0.0;
}

// Candidate block row (doubles): [0] global grid index, [1] residual d, [2],[3] aux,
// [4 .. 4+nao) phi[:, g], [4+naop .. 4+naop+k0) Lt[:k0, g].  Row `nrows` (trailer):
// [0] = [1] = -1, [2] = number of real candidates, [3] = largest residual outside the set.
constexpr int META = 4;

__global__ void pack_candidates_kernel(const double* __restrict__ phi, long long ldphi, int nao, int naop,
                                       const double* __restrict__ Lt, long long ldl, int k0,
                                       const int* __restrict__ cand, const double* __restrict__ dC,
                                       const SelState* __restrict__ st, long long g_offset,
                                       int nrows, int rl, double* __restrict__ block)
{
    const int i = blockIdx.y;                       // row
    const int count = min(st->m, nrows);
    double* row = block + (long long)i * rl;
    const int ncol = rl - META;
    if (i >= count) {
        for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < rl; c += gridDim.x * blockDim.x) {
            double v = 0.0;
            if (c < 2) v = -1.0;
            if (i == nrows && c == 2) v = (double)count;
            if (i == nrows && c == 3) v = st->theta_key ? __longlong_as_double((long long)st->theta_key) : 0.0;
            row[c] = v;
        }
        return;
    }
    const long long g = cand[i];
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ncol; c += gridDim.x * blockDim.x) {
        double v = 0.0;
        if (c < nao) v = phi[(long long)c * ldphi + g];
        else if (c >= naop && c - naop < k0) v = Lt[(long long)(c - naop) * ldl + g];
        row[META + c] = v;
    }
    if (blockIdx.x == 0 && threadIdx.x < META) {
        const int c = threadIdx.x;
        row[c] = c == 0 ? (double)(g + g_offset) : (c == 1 ? dC[i] : 0.0);
    }
}

// gathered blocks of all ranks -> candidate index / residual arrays, theta = max over ranks
__global__ void panel_prepare_kernel(const double* __restrict__ blocks, int nranks, int nrows, int rl,
                                     int* __restrict__ cand, double* __restrict__ dC, SelState* st)
{
    const int m = nranks * (nrows + 1);
    for (int c = threadIdx.x; c < m; c += blockDim.x) {
        const double* row = blocks + (long long)c * rl;
        cand[c] = (int)row[0];
        dC[c] = row[1];
    }
    if (threadIdx.x == 0) {
        double theta = 0.0;
        for (int r = 0; r < nranks; ++r)
            theta = fmax(theta, blocks[((long long)r * (nrows + 1) + nrows) * rl + 3]);
        st->theta_key = dkey(theta);
        st->m = m;
        st->barrier = 0;
    }
}

// PivRows[k][:] = blocks[piv_local[k]][:]; pivots' global indices recorded
__global__ void gather_rows_kernel(const double* __restrict__ blocks, int rl, const int* __restrict__ piv_local,
                                   int nb, int k0, double* __restrict__ PivRows, long long* __restrict__ ip_dev)
{
    const int k = blockIdx.y;
    if (k >= nb) return;
    const double* src = blocks + (long long)piv_local[k] * rl;
    double* dst = PivRows + (long long)k * rl;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < rl; c += gridDim.x * blockDim.x) dst[c] = src[c];
    if (blockIdx.x == 0 && threadIdx.x == 0) ip_dev[k0 + k] = (long long)src[0];
}

// U[j, k0+k] = L[ip_(k0+k), j]: the pivot rows of the factor, transposed (upper triangular)
__global__ void assemble_U_kernel(const double* __restrict__ PivRows, int rl, int off_l,
                                  const double* __restrict__ LppT, const int* __restrict__ piv_local,
                                  int nb, int k0, int npts, double* __restrict__ U, long long ldu)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;      // pivot within the panel (fast index)
    if (k >= nb) return;
    for (int j = blockIdx.y; j < npts; j += gridDim.y) {
        double v = 0.0;
        if (j < k0) v = PivRows[(long long)k * rl + off_l + j];
        else if (j - k0 <= k) v = LppT[(long long)(j - k0) * MCAP + piv_local[k]];
        U[(long long)j * ldu + k0 + k] = v;
    }
}

__global__ void mark_selected_kernel(const long long* __restrict__ ip_dev, int k0, int nb,
                                     long long g_offset, long long ngl, double* __restrict__ d)
{
    const int k = threadIdx.x;
    if (k < nb) {
        const long long g = ip_dev[k0 + k] - g_offset;
        if (g >= 0 && g < ngl) d[g] = -1.0;
    }
}

// Per-grid-point triangular sweep over the panel columns + residual update.
// rows: Lt_panel[k * ldl + g], k < nb.  Divisor of column k is sqrt(resid[k]).
// A thread carries PT grid points (g, g + 128, ...): every triangle entry read from shared memory (a broadcast
// load, the kernel's bound: one per FMA with PT = 1) feeds PT independent chains.  The chain of each point is
// unchanged (ascending column index), so the bits are those of the one-point form.
constexpr int SWEEP_PT = 2;
__global__ void __launch_bounds__(128)
panel_sweep_kernel(double* __restrict__ Ltp, long long ldl, long long ng, int row0, int nb,
                   const double* __restrict__ LppP, const double* __restrict__ resid,
                   double* __restrict__ d)
{
    // Rows row0 .. row0 + nb of the panel: the terms of the earlier rows j < row0 have already been subtracted
    // (by the GEMM of isdf_select_flush, in ascending j), the chain of every entry continues here with
    // j = row0, row0 + 1, ...: same order, same bits as one sweep over the whole panel.
    constexpr int PT = SWEEP_PT;
    extern __shared__ double s_l[];            // packed lower triangle of the window, then nb pivots
    double* s_piv = s_l + NB * (NB + 1) / 2;
    Ltp += (long long)row0 * ldl;
    for (int i = threadIdx.x; i < nb * nb; i += blockDim.x) {
        const int k = i / nb, j = i - k * nb;
        if (j < k) s_l[k * (k + 1) / 2 + j] = LppP[(row0 + k) * NB + row0 + j];
    }
    for (int i = threadIdx.x; i < nb; i += blockDim.x) s_piv[i] = sqrt(resid[row0 + i]);
    __syncthreads();
    const long long g0 = (long long)blockIdx.x * (blockDim.x * PT) + threadIdx.x;
    if (g0 >= ng) return;
    // points beyond the grid are clamped to the thread's first point and never stored
    long long gp[PT]; bool ok[PT]; double dd[PT];
#pragma unroll
    for (int p = 0; p < PT; ++p) {
        const long long g = g0 + (long long)p * blockDim.x;
        ok[p] = g < ng; gp[p] = ok[p] ? g : g0; dd[p] = d[gp[p]];
    }
    constexpr int SB = 16;      // rows per register block: earlier rows are re-read nb/(2*SB) times
    for (int kb = 0; kb < nb; kb += SB) {
        double acc[PT][SB], res[PT][SB];
#pragma unroll
        for (int p = 0; p < PT; ++p)
#pragma unroll
            for (int i = 0; i < SB; ++i) acc[p][i] = (kb + i < nb) ? Ltp[(long long)(kb + i) * ldl + gp[p]] : 0.0;
        // kb is a multiple of SB: fetch 8 earlier rows at a time so the loads overlap
        for (int j0 = 0; j0 < kb; j0 += 8) {
            double x[PT][8];
#pragma unroll
            for (int p = 0; p < PT; ++p)
#pragma unroll
                for (int u = 0; u < 8; ++u) x[p][u] = -Ltp[(long long)(j0 + u) * ldl + gp[p]];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
#pragma unroll
                for (int i = 0; i < SB; ++i)
                    if (kb + i < nb) {
                        const double l = s_l[(kb + i) * (kb + i + 1) / 2 + j0 + u];
#pragma unroll
                        for (int p = 0; p < PT; ++p) acc[p][i] = fma(x[p][u], l, acc[p][i]);
                    }
            }
        }
#pragma unroll
        for (int i = 0; i < SB; ++i) {
            if (kb + i < nb) {
#pragma unroll
                for (int jj = 0; jj < i; ++jj) {
                    const double l = s_l[(kb + i) * (kb + i + 1) / 2 + kb + jj];
#pragma unroll
                    for (int p = 0; p < PT; ++p) acc[p][i] = fma(-res[p][jj], l, acc[p][i]);
                }
                const double pv = s_piv[kb + i];
#pragma unroll
                for (int p = 0; p < PT; ++p) {
                    res[p][i] = acc[p][i] / pv;
                    if (ok[p]) Ltp[(long long)(kb + i) * ldl + gp[p]] = res[p][i];
                    dd[p] = fma(-res[p][i], res[p][i], dd[p]);
                }
            }
        }
    }
#pragma unroll
    for (int p = 0; p < PT; ++p)
        if (ok[p]) d[gp[p]] = dd[p];
}

struct SelWs {
    double* d; unsigned int* hist; int* cand; double* dC; double* T;
    double* LppT; int* piv_local; double* PivRows; double* LppP; double* resid_dev;
    long long* ip_dev; Partial* partials; SelState* st; double* block;
    long long* chunk_cnt; long long* chunk_base; unsigned long long* chunk_theta;
};

inline int pad2(int n) { return (n + 1) & ~1; }
inline int row_len(int nao, int k0) { return META + pad2(nao) + pad2(k0); }

size_t carve(SelWs& w, void* ws, size_t bytes, int nao, long long ngl, int npts)
{
    WsAlloc a(ws, bytes);
    const size_t rl = row_len(nao, npts);
    w.d = a.get<double>(ngl);
    w.hist = a.get<unsigned int>(2048);
    w.cand = a.get<int>(MCAP);
    w.dC = a.get<double>(MCAP);
    w.T = a.get<double>((size_t)MCAP * MCAP);
    w.LppT = a.get<double>((size_t)NB * MCAP);
    w.piv_local = a.get<int>(NB);
    w.PivRows = a.get<double>((size_t)NB * rl);
    w.LppP = a.get<double>((size_t)NB * NB);
    w.resid_dev = a.get<double>(npts);
    w.ip_dev = a.get<long long>(npts);
    w.partials = a.get<Partial>(2 * PANEL_CTAS);
    w.st = a.get<SelState>(1);
    w.block = a.get<double>((size_t)MCAP * rl);       // local candidate block (quota+1 <= MCAP rows)
    const size_t nchunks = (size_t)((ngl + COMPACT_CHUNK - 1) / COMPACT_CHUNK);
    w.chunk_cnt = a.get<long long>(2 * nchunks);
    w.chunk_base = a.get<long long>(2 * nchunks);
    w.chunk_theta = a.get<unsigned long long>(nchunks);
    return a.off + 256;
}

}  // namespace

// Host-side plan of one selection run (dimensions + carved workspace pointers).
struct isdf_select_plan {
    isdf_handle* h;
    int nao, npts;
    long long ngl, g_offset, ng_total;
    SelWs w;
};

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

extern "C" int isdf_select_max_rows(void) { return MCAP; }
extern "C" int isdf_select_panel_width(void) { return NB; }
static thread_local int g_last_panel_mode = 0;   // per calling thread (one handle per thread)
extern "C" int isdf_select_last_panel_mode(void) { return g_last_panel_mode; }
extern "C" int isdf_select_row_len(int nao, int k0) { return row_len(nao, k0); }

// ---- stepwise interface (one slab of grid points per rank) -----------------------------------
extern "C" int isdf_select_create(isdf_handle* h, int nao, long long ngl, long long g_offset,
                                  long long ng_total, int npts, void* ws, size_t ws_bytes,
                                  isdf_select_plan** out)
{
    ARG_CHECK(h && ws && out && nao > 0 && ngl > 0 && npts > 0 && g_offset >= 0);
    ARG_CHECK(ng_total >= g_offset + ngl && ng_total < 0x7fffffffLL && npts <= ng_total);
    isdf_select_plan* pl = new isdf_select_plan();
    pl->h = h; pl->nao = nao; pl->npts = npts; pl->ngl = ngl; pl->g_offset = g_offset; pl->ng_total = ng_total;
    const size_t need = carve(pl->w, ws, ws_bytes, nao, ngl, npts);
    if (need > ws_bytes) {
        isdf_set_error("select_ip workspace too small: need %zu have %zu", need, ws_bytes);
        delete pl;
        return ISDF_ERR_WS;
    }
    *out = pl;
    return ISDF_OK;
}

extern "C" int isdf_select_destroy(isdf_select_plan* pl)
{
    delete pl;
    return ISDF_OK;
}

// d_g = (sum_mu phi^2)^2 on the local slab; *h_dmax = its maximum (host sync)
extern "C" int isdf_select_begin(isdf_select_plan* pl, const double* phi, long long ldphi, double* h_dmax)
{
    ARG_CHECK(pl && phi && h_dmax && ldphi >= pl->ngl);
    isdf_handle* h = pl->h;
    cudaStream_t s = h->stream;
    SelWs& w = pl->w;
    init_diag_kernel<<<(unsigned)((pl->ngl + 255) / 256), 256, 0, s>>>(phi, ldphi, pl->nao, pl->ngl, w.d);
    state_init_kernel<<<1, 1024, 0, s>>>(w.d, pl->ngl, 0.0, w.st);
    CUDA_TRY(cudaMemsetAsync(w.hist, 0, 2048 * sizeof(unsigned int), s));
    CUDA_TRY(cudaGetLastError());
    h->other_launches += 2;
    SelState* hst = (SelState*)h->h_pinned;
    CUDA_TRY(cudaMemcpyAsync(hst, w.st, sizeof(SelState), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    *h_dmax = hst->dmax0;
    return ISDF_OK;
}

// stop threshold on the residual diagonal (absolute); the default rule is
// tol = ng_total * eps * max_g d_g (pyscf/lib/scipy_helper.py:88-90)
extern "C" int isdf_select_set_tol(isdf_select_plan* pl, double tol)
{
    ARG_CHECK(pl != nullptr);
    CUDA_TRY(cudaMemcpyAsync(&pl->w.st->tol, &tol, sizeof(double), cudaMemcpyHostToDevice, pl->h->stream));
    CUDA_TRY(cudaStreamSynchronize(pl->h->stream));
    return ISDF_OK;
}

// The `quota` largest local residuals -> block[(quota+1)][row_len(nao,k0)] (see pack_candidates_kernel).
// block == nullptr packs into the plan's own buffer.
extern "C" int isdf_select_candidates(isdf_select_plan* pl, const double* phi, long long ldphi,
                                      const double* Lt, long long ldl, int k0, int quota, double* block)
{
    ARG_CHECK(pl && phi && Lt && k0 >= 0 && quota >= 1 && quota < MCAP);
    isdf_handle* h = pl->h;
    cudaStream_t s = h->stream;
    SelWs& w = pl->w;
    if (!block) block = w.block;
    const int shifts[6] = {52, 41, 30, 20, 10, 0};
    const int bitsv[6] = {11, 11, 11, 10, 10, 10};
    const int hist_blocks = (int)std::min<long long>((pl->ngl + 255) / 256, 2 * h->num_sms);
    radix_begin_kernel<<<1, 1, 0, s>>>(w.st, (long long)quota);
    for (int pass = 0; pass < 6; ++pass) {
        radix_hist_kernel<<<hist_blocks, 256, 0, s>>>(w.d, pl->ngl, w.st, shifts[pass], bitsv[pass], w.hist);
        radix_scan_kernel<<<1, 1024, 0, s>>>(w.st, shifts[pass], bitsv[pass], w.hist, pass == 5);
    }
    {
        const int nchunks = (int)((pl->ngl + COMPACT_CHUNK - 1) / COMPACT_CHUNK);
        compact_chunk_kernel<false><<<nchunks, 1024, 0, s>>>(w.d, pl->ngl, w.st, w.chunk_cnt, w.chunk_base,
                                                             w.chunk_theta, w.cand, w.dC);
        compact_scan_kernel<<<1, 32, 0, s>>>(w.st, nchunks, w.chunk_cnt, w.chunk_base, w.chunk_theta);
        compact_chunk_kernel<true><<<nchunks, 1024, 0, s>>>(w.d, pl->ngl, w.st, w.chunk_cnt, w.chunk_base,
                                                            w.chunk_theta, w.cand, w.dC);
    }
    const int rl = row_len(pl->nao, k0);
    dim3 grid((rl + 255) / 256, quota + 1);
    pack_candidates_kernel<<<grid, 256, 0, s>>>(phi, ldphi, pl->nao, pad2(pl->nao), Lt, ldl, k0, w.cand, w.dC,
                                               w.st, pl->g_offset, quota, rl, block);
    CUDA_TRY(cudaGetLastError());
    h->other_launches += 17;
    return ISDF_OK;
}

// rows per block of the panel sweep (ISDF_SWEEP_BLOCK: 16 .. 128, a multiple of 16; 128 = one scalar sweep)
static int sweep_block()
{
    static const int v = [] {
        const char* e = getenv("ISDF_SWEEP_BLOCK");
        int b = e ? atoi(e) : 32;
        if (b < 16 || b > NB || (b & 15)) b = 32;
        return b;
    }();
    return v;
}

static int gemm_kc(isdf_handle* h, int M, int N, int K, const double* A, long long lda,
                   const double* B, long long ldb, int b_kmajor, double* C, long long ldc,
                   const double* Cin, int neg, int epi, int upper_only = 0)
{
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.A = A; p.lda = lda; p.a_kmajor = 0;
    p.B = B; p.ldb = ldb; p.b_kmajor = b_kmajor;
    p.C = C; p.ldc = ldc; p.Cin = Cin; p.ldcin = ldc;
    p.E = nullptr; p.lde = 0; p.alpha = 1.0; p.neg = neg; p.epi = epi; p.upper_only = upper_only;
    return gemm_launch(h, p);
}

// T[i][j] = T[j][i] for i > j (the candidate Gram is symmetric: only its upper tiles are computed)
__global__ void mirror_upper_kernel(double* __restrict__ T, long long ld, int n)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i < n && j < i) T[(long long)i * ld + j] = T[(long long)j * ld + i];
}

// Panel factorisation on the gathered candidate blocks of all ranks
// (blocks: [nranks * (quota+1)][row_len]); every rank runs it redundantly on
// identical data and gets identical pivots.  U (nullable) receives the new
// columns of the transposed pivot factor.  Host sync: returns nb, stop.
extern "C" int isdf_select_panel(isdf_select_plan* pl, const double* blocks, int nranks, int quota,
                                 int k0, int nbmax, double* U, long long ldu, int* h_nb, int* h_stop)
{
    ARG_CHECK(pl && h_nb && h_stop && nranks >= 1 && quota >= 1 && k0 >= 0 && nbmax >= 1 && nbmax <= NB);
    isdf_handle* h = pl->h;
    cudaStream_t s = h->stream;
    SelWs& w = pl->w;
    if (!blocks) blocks = w.block;
    const int m = nranks * (quota + 1);
    ARG_CHECK(m <= MCAP);
    const int rl = row_len(pl->nao, k0), naop = pad2(pl->nao);
    int rc;
    panel_prepare_kernel<<<1, 1024, 0, s>>>(blocks, nranks, quota, rl, w.cand, w.dC, w.st);
    h->other_launches++;
    // T = G[C,C] - Lc Lc^T  (same FMA chains as the flush GEMMs)
    // symmetric: upper tiles only, then mirrored (x*y == y*x, so the mirrored entries are the bits
    // the full product would have produced)
    if ((rc = gemm_kc(h, m, m, pl->nao, blocks + META, rl, blocks + META, rl, 0, w.T, MCAP, nullptr, 0, EPI_SQUARE, 1))) return rc;
    if (k0 > 0)
        if ((rc = gemm_kc(h, m, m, k0, blocks + META + naop, rl, blocks + META + naop, rl, 0, w.T, MCAP, w.T, 1, EPI_STORE, 1))) return rc;
    {
        dim3 blk(32, 8), grd((m + 31) / 32, (m + 7) / 8);
        mirror_upper_kernel<<<grd, blk, 0, s>>>(w.T, MCAP, m);
        h->other_launches++;
    }
    {
        SelState* st = w.st; const double* T = w.T; long long ldT = MCAP; const int* cand = w.cand;
        double* dC = w.dC; double* LppT = w.LppT; int* pvl = w.piv_local; double* rs = w.resid_dev;
        Partial* pa = w.partials; int k0a = k0, nbm = nbmax;
        // the CTA's own factor columns live in shared memory when they fit (cpb <= 192 candidates per CTA)
        const int cpb = (((m + PANEL_CTAS - 1) / PANEL_CTAS) + 31) & ~31;
        int use_smem = cpb <= 192 ? 1 : 0;
        const size_t psmem = use_smem ? sizeof(double) * NB * cpb : 0;
        if (psmem > 0) {
            int rc_;
            if ((rc_ = isdf_func_max_smem((const void*)panel_kernel, sizeof(double) * NB * 192))) return rc_;
        }
        // opt-in: the same factorisation as one 16-CTA thread-block cluster (DSMEM exchange, hardware barrier)
        // (the variable is read at every panel so that one process can compare the forms)
        int cluster_ok = -1;                     // 0 unavailable on this device, 1 usable
        int cluster_mode = 0;                    // 1: DSMEM reads + cluster barrier, 2: st.async pushes + mbarrier
        {
            // default since round 2: form 2 (st.async pushes), 34 ms faster per C64 build on every rank; passed its
            // bit-exactness tests on three boxes and the full C64 oracle comparison.  0 = grid-barrier kernel.
            const char* e = getenv("ISDF_PANEL_CLUSTER");
            cluster_mode = e ? atoi(e) : 2;
            if (cluster_mode < 0 || cluster_mode > 2) cluster_mode = 2;
        }
        if (cluster_mode) {                      // attributes: once per (kernel, device), remembered with their outcome
            cudaError_t e1 = isdf_func_attr_once((const void*)panel_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaError_t e2 = isdf_func_attr_once((const void*)panel_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(sizeof(double) * NB * 192));
            cudaError_t e3 = isdf_func_attr_once((const void*)panel_push_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaError_t e4 = isdf_func_attr_once((const void*)panel_push_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(sizeof(double) * NB * PUSH_CPB_MAX));
            cluster_ok = (e1 == cudaSuccess && e2 == cudaSuccess && e3 == cudaSuccess && e4 == cudaSuccess) ? 1 : 0;
        }
        if (cluster_ok != 1) cluster_mode = 0;
        bool launched = false;
        if (cluster_mode >= 1 && use_smem && (cluster_mode == 1 || cpb <= PUSH_CPB_MAX)) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(PANEL_CTAS); cfg.blockDim = dim3(PANEL_THREADS);
            cfg.dynamicSmemBytes = psmem; cfg.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = PANEL_CTAS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int nclusters = 0;
            auto kern = cluster_mode == 2 ? panel_push_kernel : panel_cluster_kernel;
            if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) == cudaSuccess && nclusters >= 1) {
                CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, st, T, ldT, cand, dC, LppT, pvl, rs, k0a, nbm));
                launched = true;
                g_last_panel_mode = cluster_mode;
            } else {
                cudaGetLastError();
                cluster_ok = 0;                  // a 16-CTA cluster does not fit on this device: stay on the grid-barrier kernel
            }
        }
        if (!launched) {
            void* args[] = {&st, &T, &ldT, &cand, &dC, &LppT, &pvl, &rs, &pa, &k0a, &nbm, &use_smem};
            CUDA_TRY(cudaLaunchCooperativeKernel((void*)panel_kernel, dim3(PANEL_CTAS), dim3(PANEL_THREADS), args, psmem, s));
            g_last_panel_mode = 0;
        }
        h->other_launches++;
    }
    SelState* hst = (SelState*)h->h_pinned;
    CUDA_TRY(cudaMemcpyAsync(hst, w.st, sizeof(SelState), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    const int nb = hst->nb_eff;
    *h_nb = nb; *h_stop = hst->stop;
    if (nb > 0) {
        dim3 g1((rl + 255) / 256, nb);
        gather_rows_kernel<<<g1, 256, 0, s>>>(blocks, rl, w.piv_local, nb, k0, w.PivRows, w.ip_dev);
        gather_lpp_kernel<<<nb, NB, 0, s>>>(w.LppT, w.piv_local, nb, w.LppP);
        h->other_launches += 2;
        if (U) {
            dim3 g2((nb + 127) / 128, std::min(pl->npts, 2048));
            assemble_U_kernel<<<g2, 128, 0, s>>>(w.PivRows, rl, META + naop, w.LppT, w.piv_local, nb, k0,
                                                 pl->npts, U, ldu);
            h->other_launches++;
        }
        CUDA_TRY(cudaGetLastError());
    }
    return ISDF_OK;
}

// New factor rows k0 .. k0+nb for the local slab, residual update, pivots masked.
extern "C" int isdf_select_flush(isdf_select_plan* pl, const double* phi, long long ldphi,
                                 double* Lt, long long ldl, int k0, int nb)
{
    ARG_CHECK(pl && phi && Lt && k0 >= 0 && nb >= 0 && nb <= NB && k0 + nb <= pl->npts);
    if (nb == 0) return ISDF_OK;
    isdf_handle* h = pl->h;
    cudaStream_t s = h->stream;
    SelWs& w = pl->w;
    const int rl = row_len(pl->nao, k0), naop = pad2(pl->nao);
    const long long ngl = pl->ngl;
    const size_t sweep_smem = sizeof(double) * (NB * (NB + 1) / 2 + NB);
    int rc;
    if ((rc = isdf_func_max_smem((const void*)panel_sweep_kernel, sweep_smem))) return rc;
    double* Ltp = Lt + (long long)k0 * ldl;
    if ((rc = gemm_kc(h, nb, (int)ngl, pl->nao, w.PivRows + META, rl, phi, ldphi, 1, Ltp, ldl, nullptr, 0, EPI_SQUARE))) return rc;
    if (k0 > 0)
        if ((rc = gemm_kc(h, nb, (int)ngl, k0, w.PivRows + META + naop, rl, Lt, ldl, 1, Ltp, ldl, Ltp, 1, EPI_STORE))) return rc;
    // Triangular sweep over the panel's own columns, in row blocks: the part of a block's chains that runs over
    // earlier blocks is a product with the panel factor -- on the FP64 tensor pipe, in place, ascending column
    // index -- and only the block's own triangle goes through the scalar kernel (a quarter of the FMAs at 32
    // rows per block).  Every entry's chain keeps its order: flush columns, earlier blocks, own block.
    const int bs = sweep_block();
    for (int r0 = 0; r0 < nb; r0 += bs) {
        const int nr = nb - r0 < bs ? nb - r0 : bs;
        if (r0 > 0)
            if ((rc = gemm_kc(h, nr, (int)ngl, r0, w.LppP + (long long)r0 * NB, NB, Ltp, ldl, 1,
                              Ltp + (long long)r0 * ldl, ldl, Ltp + (long long)r0 * ldl, 1, EPI_STORE))) return rc;
        panel_sweep_kernel<<<(unsigned)((ngl + 128 * SWEEP_PT - 1) / (128 * SWEEP_PT)), 128, sweep_smem, s>>>(
            Ltp, ldl, ngl, r0, nr, w.LppP, w.resid_dev + k0, w.d);
        h->other_launches++;
    }
    mark_selected_kernel<<<1, NB, 0, s>>>(w.ip_dev, k0, nb, pl->g_offset, ngl, w.d);
    CUDA_TRY(cudaGetLastError());
    h->other_launches += 1;
    return ISDF_OK;
}

extern "C" int isdf_select_finish(isdf_select_plan* pl, int rank, long long* h_ip, double* h_resid)
{
    ARG_CHECK(pl && h_ip && rank >= 0 && rank <= pl->npts);
    cudaStream_t s = pl->h->stream;
    if (rank > 0) {
        CUDA_TRY(cudaMemcpyAsync(h_ip, pl->w.ip_dev, sizeof(long long) * rank, cudaMemcpyDeviceToHost, s));
        if (h_resid) CUDA_TRY(cudaMemcpyAsync(h_resid, pl->w.resid_dev, sizeof(double) * rank, cudaMemcpyDeviceToHost, s));
    }
    CUDA_TRY(cudaStreamSynchronize(s));
    return ISDF_OK;
}

// Quota policy shared by the single-GPU driver below and the multi-rank host
// driver (pyscf_isdf_b200/dist.py): grow the candidate set when the exactness
// certificate ends panels early.
extern "C" int isdf_select_first_quota(int nranks);
extern "C" int isdf_select_next_quota(int quota, int nb, int nbmax, int nranks)
{
    // quota + 1 rows per rank (candidates + trailer) stays a power of two: whole GEMM tiles.
    // The redundant panel work grows with (nranks * quota)^2, so the set is enlarged only
    // while the certificate keeps ending panels early (ties across the set boundary) and
    // falls back to the base size after every full panel.
    const int base = isdf_select_first_quota(nranks);
    const int cap = std::min(MCAP, 2048) / nranks - 1;
    if (nb >= nbmax) return base;
    if (nb < nbmax / 2 && 2 * quota + 1 <= cap) return 2 * quota + 1;
    return quota;
}

extern "C" int isdf_select_first_quota(int nranks)
{
    return std::max(NB, 1024 / nranks) - 1;
}

// ---- single-GPU driver ----------------------------------------------------------------
extern "C" int isdf_select_ip(isdf_handle* h, const double* phi, int nao, long long ng, long long ldphi,
                              int npts, double tol, double* Lt, long long ldl,
                              long long* h_ip, double* h_resid, int* h_rank,
                              void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi && Lt && h_ip && h_rank && ws);
    ARG_CHECK(nao > 0 && ng > 0 && ng < 0x7fffffffLL && npts > 0 && npts <= ng && ldphi >= ng && ldl >= ng);
    isdf_select_plan* pl = nullptr;
    int rc;
    if ((rc = isdf_select_create(h, nao, ng, 0, ng, npts, ws, ws_bytes, &pl))) return rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_SELECT], s));
    double dmax = 0.0;
    rc = isdf_select_begin(pl, phi, ldphi, &dmax);
    if (!rc) rc = isdf_select_set_tol(pl, tol < 0.0 ? (double)ng * 2.220446049250313e-16 * dmax : tol);
    int quota = std::min(isdf_select_first_quota(1), (int)std::min<long long>(ng, MCAP - 1));
    int k0 = 0;
    while (!rc && k0 < npts) {
        const int nbmax = std::min(NB, npts - k0);
        int nb = 0, stop = 0;
        if ((rc = isdf_select_candidates(pl, phi, ldphi, Lt, ldl, k0, quota, nullptr))) break;
        if ((rc = isdf_select_panel(pl, nullptr, 1, quota, k0, nbmax, nullptr, 0, &nb, &stop))) break;
        if ((rc = isdf_select_flush(pl, phi, ldphi, Lt, ldl, k0, nb))) break;
        k0 += nb;
        if (stop || nb == 0) break;
        quota = isdf_select_next_quota(quota, nb, nbmax, 1);
    }
    if (!rc) {
        CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_SELECT], s));
        h->ev_used[ISDF_STAGE_SELECT] = true;
        rc = isdf_select_finish(pl, k0, h_ip, h_resid);
        *h_rank = k0;
    }
    isdf_select_destroy(pl);
    return rc;
}
