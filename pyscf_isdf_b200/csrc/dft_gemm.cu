// Stage 4, forward transforms as dense DFTs on the FP64 tensor pipe (DMMA m8n8k4).
//
// Why: the Stockham / Rader line FFTs of fft_core.cuh move every point through shared memory once per
// radix stage with run-time index arithmetic -- ncu (profiles/r02_ncu_fft.md) shows ~450 thread
// instructions per complex point and pass, issue-bound at 20 % of HBM peak.  For the mesh sizes of this
// path (n <= 128) a DFT written as a small GEMM costs 2 n^2 flops per line once both symmetries of the
// cosine / sine matrices are used, runs on the DMMA pipe at one instruction per 512 flops, needs no
// twiddle bookkeeping, and works for ANY n (71 is prime: Rader needed two 70-point sub-FFTs).
//
//   x_j, j = 0..n-1;  e_j = x_j + x_{n-j},  o_j = x_j - x_{n-j}   (j = 1..hp = (n-1)/2;  e_0 = x_0, e_{n/2} = x_{n/2})
//   P_k = sum_{j in 0..n/2} cos(2 pi k j/n) e_j ,   Q_k = sum_{j in 1..hp} sin(2 pi k j/n) o_j ,   k = 0..n/2
//   X_k = P_k - i Q_k ,   X_{n-k} = P_k + i Q_k                    (forward sign, as pyscf/pbc/tools/pbc.py:149-170)
//
// P = C e and Q = S o are two real GEMMs [ne x ne] . [ne x columns] (ne = n/2 + 1) over complex data
// viewed as pairs of real columns; C and S sit in shared memory, the data tile too (e and o overwrite
// x in place), accumulators in registers, the combine and the fused sqrt(w coulG) weight in the epilogue.
// Three kernels: z pass (real lines -> half spectrum), column pass (y, x), Coulomb-weight table.
#include "common.cuh"
#include "fft_kernels.cuh"
#include <vector>
#include <cmath>
#include <cstdlib>
#include <algorithm>

namespace {

constexpr int DFT_THREADS = 256;
constexpr int MG = 5;                       // m-tiles (8 output frequencies each) accumulated at once

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c[0]), "+d"(c[1])
                 : "d"(a), "d"(b));
}

// Asynchronous global -> shared copies (LDGSTS): the whole tile of a CTA is in flight at once; a synchronous
// load loop kept only ~16 KB per SM in flight and ran the passes at 19 GB/s per SM (profiles/r02_ncu_fft.md).
__device__ __forceinline__ void cp_async8(void* dst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Table layout in memory (global and shared): tab[(i * KS + s) * 64 + 2 * lane + {0, 1}] = (C, S)[8 i + lane / 4][4 s + lane % 4]
// for m-tile i < MT and k-step s < KS = KP / 4: the DMMA A fragments of both matrices for one (tile, step) are one
// 16-byte load per lane, 512 contiguous bytes per warp.
__device__ __forceinline__ void load_tables(double* tab_s, const DftAxis& ax)
{
    const int tot2 = ax.MT * (ax.KP / 4) * 32;
    for (int i = threadIdx.x; i < tot2; i += DFT_THREADS) cp_async16(tab_s + 2 * i, ax.C + 2 * i);
}

// The two GEMMs of a pair of 8-column tiles: p += C e, q += S o for MG m-tiles starting at mg.
// pe0/pe1 (po0/po1) point at this lane's column of the folded tile: element j of e (o) is pe[offE(j)] (po[offO(j)]).
// COLS = true: rows of T are the transform index (column pass); false: columns are (z pass).
template <bool COLS, int NT>
__device__ __forceinline__ void dft_mma_group(double (&p)[MG][NT][2], double (&q)[MG][NT][2], const double* __restrict__ tab_s,
                                              const double* __restrict__ b0, const double* __restrict__ b1, int LDT,
                                              const DftAxis& ax, int mg, int lane)
{
    const int n = ax.n, ne = ax.ne, hp = ax.hp, KS = ax.KP / 4, t = lane & 3;
    const double2* tl = (const double2*)tab_s + lane;
#pragma unroll
    for (int i = 0; i < MG; ++i)
#pragma unroll
        for (int a = 0; a < NT; ++a) { p[i][a][0] = p[i][a][1] = 0.0; q[i][a][0] = q[i][a][1] = 0.0; }
    for (int s = 0; s < KS; ++s) {
        const int j = 4 * s + t;
        const int je = min(j, ne - 1);
        const int jo = (j >= 1 && j <= hp) ? n - j : 0;
        const int oe = COLS ? je * LDT : je, oo = COLS ? jo * LDT : jo;
        const double e0 = b0[oe], o0 = b0[oo];
        double e1 = 0.0, o1 = 0.0;
        if (NT == 2) { e1 = b1[oe]; o1 = b1[oo]; }
#pragma unroll
        for (int i = 0; i < MG; ++i)
            if (mg + i < ax.MT) {
                const double2 cs = tl[((mg + i) * KS + s) * 32];
                dmma(p[i][0], cs.x, e0); dmma(q[i][0], cs.y, o0);
                if (NT == 2) { dmma(p[i][NT - 1], cs.x, e1); dmma(q[i][NT - 1], cs.y, o1); }
            }
    }
}

// ---- z pass: nlines real lines of length n (contiguous) -> nlines x (n/2+1) complex -----------------
template <int NT>
__global__ void __launch_bounds__(DFT_THREADS, NT == 2 ? 2 : 3)
dft_z_r2c_kernel(const double* __restrict__ in, long long ldin, double2* __restrict__ out, long long ldout,
                 int nlines, int LW, int LDT, DftAxis ax)
{
    extern __shared__ __align__(16) double smd[];
    const int n = ax.n, ne = ax.ne, hp = ax.hp;
    double* tab_s = smd;
    double* T = tab_s + ax.MT * (ax.KP / 4) * 64;
    const int tid = threadIdx.x;
    const int l0 = blockIdx.x * LW;
    const int nl = min(LW, nlines - l0);
    const double* src = in + (long long)blockIdx.y * ldin + (long long)l0 * n;
    load_tables(tab_s, ax);
    {
        // contiguous chunk of nl * n doubles; (line, z) of element idx advanced incrementally
        const int dl = DFT_THREADS / n, dz = DFT_THREADS - dl * n;
        int line = tid / n, z = tid - line * n;
        for (int idx = tid; idx < LW * n; idx += DFT_THREADS) {
            if (line < nl) cp_async8(T + line * LDT + z, src + idx);
            else T[line * LDT + z] = 0.0;
            line += dl; z += dz;
            if (z >= n) { z -= n; ++line; }
        }
    }
    cp_async_wait_all();
    __syncthreads();
    if (hp > 0) {
        const int dl = DFT_THREADS / hp, dj = DFT_THREADS - dl * hp;
        int line = tid / hp, jj = tid - line * hp;
        for (int idx = tid; idx < LW * hp; idx += DFT_THREADS) {
            double* row = T + line * LDT;
            const double a = row[jj + 1], b = row[n - 1 - jj];
            row[jj + 1] = a + b; row[n - 1 - jj] = a - b;
            line += dl; jj += dj;
            if (jj >= hp) { jj -= hp; ++line; }
        }
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    double2* dst = out + (long long)blockIdx.y * ldout + (long long)l0 * ne;
    for (int lb = warp * 8 * NT; lb < LW; lb += (DFT_THREADS / 32) * 8 * NT) {
        if (lb >= nl) break;
        const double* r0 = T + (lb + g) * LDT;
        const double* r1 = T + (lb + 8 * (NT - 1) + g) * LDT;
        for (int mg = 0; mg < ax.MT; mg += MG) {
            double p[MG][NT][2], q[MG][NT][2];
            dft_mma_group<false, NT>(p, q, tab_s, r0, r1, LDT, ax, mg, lane);
#pragma unroll
            for (int i = 0; i < MG; ++i) {
                const int k = 8 * (mg + i) + g;
                if (mg + i < ax.MT && k < ne) {
#pragma unroll
                    for (int a = 0; a < NT; ++a)
#pragma unroll
                        for (int r = 0; r < 2; ++r) {
                            const int line = lb + 8 * a + 2 * t + r;
                            if (line < nl) dst[(long long)line * ne + k] = make_double2(p[i][a][r], -q[i][a][r]);
                        }
                }
            }
        }
    }
}

// ---- column pass: rows j = 0..n-1 at stride row_stride, NV "virtual" columns per vector -----------------
// virtual column v -> (outer o = v / ncols, c = v % ncols) at offset o * outer_stride + c (complex units).
// x pass: one outer block, ncols = row_stride = columns of the slab.  y pass: outer = x planes, ncols = nz/2+1.
// MODE 1 multiplies row k, column v by wtab[k * wld + v] (x pass of the W build: sqrt(w coulG)).
template <int MODE, int NT>
__global__ void __launch_bounds__(DFT_THREADS, NT == 2 ? 2 : 3)
dft_cols_kernel(double2* __restrict__ data, long long vec_stride, long long outer_stride, int row_stride,
                int ncols, int NV, int CW, DftAxis ax, const double* __restrict__ wtab, int wld)
{
    extern __shared__ __align__(16) double smd[];
    const int n = ax.n, ne = ax.ne, hp = ax.hp;
    const int LDT = 2 * CW + 4;
    double* tab_s = smd;
    double* T = tab_s + ax.MT * (ax.KP / 4) * 64;
    const int tid = threadIdx.x;
    const int v0 = blockIdx.x * CW;
    const int nv = min(CW, NV - v0);
    double2* base = data + (long long)blockIdx.y * vec_stride;
    load_tables(tab_s, ax);
    const float inv_nc = 1.0f / (float)ncols;
    // every thread owns one column c of the tile and the rows j0, j0 + rstep, ...  (CW divides DFT_THREADS)
    const int c = tid % CW, j0 = tid / CW, rstep = DFT_THREADS / CW;
    {
        double* dstp = T + j0 * LDT + 2 * c;
        if (c < nv) {
            const int v = v0 + c;
            const int o = fast_div(v, ncols, inv_nc);
            const double2* srcp = base + (long long)o * outer_stride + (v - o * ncols) + (long long)j0 * row_stride;
            const long long sstep = (long long)rstep * row_stride;
            for (int j = j0; j < n; j += rstep, dstp += rstep * LDT, srcp += sstep) cp_async16(dstp, srcp);
        } else {
            for (int j = j0; j < n; j += rstep, dstp += rstep * LDT) *(double2*)dstp = make_double2(0.0, 0.0);
        }
    }
    cp_async_wait_all();
    __syncthreads();
    // fold in place: row j <- x_j + x_{n-j}, row n-j <- x_j - x_{n-j}
    for (int j = 1 + j0; j <= hp; j += rstep) {
        double2* pa = (double2*)(T + j * LDT + 2 * c);
        double2* pb = (double2*)(T + (n - j) * LDT + 2 * c);
        const double2 a = *pa, b = *pb;
        *pa = make_double2(a.x + b.x, a.y + b.y);
        *pb = make_double2(a.x - b.x, a.y - b.y);
    }
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int ntiles = CW / 4;                                  // 4 complex = 8 real columns each
    for (int tb = warp * NT; tb < ntiles; tb += (DFT_THREADS / 32) * NT) {
        if (tb * 4 >= nv) break;
        const int cb0 = tb * 8, cb1 = min(tb + NT - 1, ntiles - 1) * 8;   // first real column of the tile(s)
        // this lane's output column(s): offsets once per tile group
        long long off[NT]; bool ok[NT]; int vv[NT];
#pragma unroll
        for (int a = 0; a < NT; ++a) {
            const int cc = (a == 0 ? cb0 : cb1) / 2 + t;
            ok[a] = cc < nv && (a == 0 || tb + 1 < ntiles);
            const int v = v0 + min(cc, nv - 1);
            const int o = fast_div(v, ncols, inv_nc);
            off[a] = (long long)o * outer_stride + (v - o * ncols);
            vv[a] = v;
        }
        for (int mg = 0; mg < ax.MT; mg += MG) {
            double p[MG][NT][2], q[MG][NT][2];
            dft_mma_group<true, NT>(p, q, tab_s, T + cb0 + g, T + cb1 + g, LDT, ax, mg, lane);
#pragma unroll
            for (int i = 0; i < MG; ++i) {
                const int k = 8 * (mg + i) + g;
                if (mg + i < ax.MT && k < ne) {
                    const bool mirror = (k >= 1 && k <= hp);
#pragma unroll
                    for (int a = 0; a < NT; ++a) {
                        if (ok[a]) {
                            // X_k = P_k - i Q_k,  X_{n-k} = P_k + i Q_k
                            double2 lo = make_double2(p[i][a][0] + q[i][a][1], p[i][a][1] - q[i][a][0]);
                            double2 hi = make_double2(p[i][a][0] - q[i][a][1], p[i][a][1] + q[i][a][0]);
                            if (MODE == 1) {
                                const double w = wtab[(long long)k * wld + vv[a]];
                                lo.x *= w; lo.y *= w;
                                if (mirror) { const double w2 = wtab[(long long)(n - k) * wld + vv[a]]; hi.x *= w2; hi.y *= w2; }
                            }
                            double2* col = base + off[a];
                            col[(long long)k * row_stride] = lo;
                            if (mirror) col[(long long)(n - k) * row_stride] = hi;
                        }
                    }
                }
            }
        }
    }
}

// sqrt(w coulG) of the W build for the column slab [col0, col0 + ncl): w[ix * ncl + c]
__global__ void coul_weight_kernel(CoulGeom geo, int nx, int ncl, double* __restrict__ w)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int ix = blockIdx.y;
    if (c >= ncl || ix >= nx) return;
    const int col = geo.col0 + c;
    const int iy = col / geo.nzh, iz = col - iy * geo.nzh;
    const double f = coulG_at(geo, ix, iy, iz);
    const bool self = (iz == 0) || (2 * iz == geo.mesh[2]);
    w[(long long)ix * ncl + c] = sqrt(self ? f : 2.0 * f);
}

size_t table_doubles(const DftAxis& ax) { return (size_t)ax.MT * (ax.KP / 4) * 64; }

}  // namespace

namespace isdf_fft {

// environment ISDF_FFT_DFT: 0 = never, 1 = every axis with n <= DFT_NMAX, unset = where it measured faster:
// n <= 112 (48^3 1.5x, 64^3 1.7x, 71^3 3.2x, 96^3 / 108^3 on par with the Stockham kernels), and up to DFT_NMAX
// when n has a prime factor above 7 (the Stockham path handles those by Rader or a direct O(r^2) stage);
// 128 = 2^7 stays on the radix-4/2 Stockham kernels (9.6 against 15.5 ms for 190 vectors of 128^3).
bool dft_enabled(int n)
{
    static const int mode = [] { const char* e = getenv("ISDF_FFT_DFT"); return e ? atoi(e) : -1; }();
    if (mode == 0 || n < 2 || n > DFT_NMAX) return false;
    if (mode > 0 || n <= 112) return true;
    int m = n;
    for (int f : {2, 3, 5, 7}) while (m % f == 0) m /= f;
    return m > 1;
}

void dft_axis_shape(int n, DftAxis& ax)
{
    ax.n = n; ax.ne = n / 2 + 1; ax.hp = (n - 1) / 2;
    ax.MT = (ax.ne + 7) / 8;
    ax.KP = (ax.ne + 3) / 4 * 4;
    ax.LDA = ax.KP;
    while (ax.LDA % 16 != 4 && ax.LDA % 16 != 12) ++ax.LDA;     // conflict-free DMMA fragment loads
    ax.C = ax.S = nullptr;
}

// host table in the fragment layout of load_tables (C and S interleaved, zero outside the ne x ne block);
// appended to `tab`; off_c is its offset, off_s is set to the same value (one table)
void dft_build_tables(int n, DftAxis& ax, std::vector<double>& tab, size_t& off_c, size_t& off_s)
{
    dft_axis_shape(n, ax);
    const size_t nd = table_doubles(ax);
    off_c = off_s = tab.size();
    tab.resize(tab.size() + nd, 0.0);
    const long double PI2 = 2.0L * 3.141592653589793238462643383279502884L;
    const int KS = ax.KP / 4;
    for (int i = 0; i < ax.MT; ++i)
        for (int s = 0; s < KS; ++s)
            for (int lane = 0; lane < 32; ++lane) {
                const int k = 8 * i + lane / 4, j = 4 * s + lane % 4;
                if (k >= ax.ne || j >= ax.ne) continue;
                const long double ang = PI2 * (long double)(((long long)k * j) % n) / (long double)n;
                const size_t at = off_c + ((size_t)(i * KS + s) * 32 + lane) * 2;
                tab[at] = (double)cosl(ang);
                if (j >= 1 && j <= ax.hp) tab[at + 1] = (double)sinl(ang);
            }
}

static size_t smem_z(const DftAxis& ax, int LW, int LDT) { return sizeof(double) * (table_doubles(ax) + (size_t)LW * LDT + 16); }
static size_t smem_cols(const DftAxis& ax, int CW) { return sizeof(double) * (table_doubles(ax) + (size_t)ax.n * (2 * CW + 4) + 16); }

constexpr size_t SMEM_TWO_CTAS = 113 * 1024, SMEM_THREE_CTAS = 75 * 1024, SMEM_ONE_CTA = 226 * 1024;

// ISDF_DFT_NT (tuning knob): 8-column tiles a warp accumulates at once.  2 = fewer table loads per DMMA, 128
// registers, 2 CTAs per SM; 1 = 80 registers and half-size tiles, 3 CTAs per SM (more overlap of the load /
// compute / store phases of different CTAs)
// Default: 1 while three CTAs of the smaller tile fit (n <= ~80: 4.5 against 5.1 ms for 1024 vectors of 71^3),
// else 2 (n = 96, 108: the tables alone are ~50 KB; 10.9 against 16.6 ms for 384 vectors of 108^3).
int dft_nt(const DftAxis& ax)
{
    static const int forced = [] { const char* e = getenv("ISDF_DFT_NT"); return e ? atoi(e) : 0; }();
    if (forced == 1 || forced == 2) return forced;
    return smem_cols(ax, 32) <= SMEM_THREE_CTAS ? 1 : 2;
}

int dft_z_r2c(cudaStream_t s, const double* in, long long ldin, double2* out, long long ldout, int nvec,
              int nlines, const DftAxis& ax)
{
    int LDT = ax.n;
    while (LDT % 16 != 4 && LDT % 16 != 12) ++LDT;
    const int NT = dft_nt(ax);
    const size_t want = NT == 2 ? SMEM_TWO_CTAS : SMEM_THREE_CTAS;
    int LW = NT == 2 ? 128 : 64;
    while (LW > 16 && smem_z(ax, LW, LDT) > want) LW /= 2;
    const size_t smem = smem_z(ax, LW, LDT);
    if (smem > SMEM_ONE_CTA) { isdf_set_error("dft_z_r2c: n=%d does not fit shared memory", ax.n); return ISDF_ERR_ARG; }
    int rc;
    const void* kern = NT == 2 ? (const void*)dft_z_r2c_kernel<2> : (const void*)dft_z_r2c_kernel<1>;
    if ((rc = isdf_func_max_smem(kern, smem))) return rc;
    for (int v0 = 0; v0 < nvec; v0 += 65535) {
        const int nv = std::min(65535, nvec - v0);
        dim3 grid((nlines + LW - 1) / LW, nv);
        if (NT == 2) dft_z_r2c_kernel<2><<<grid, DFT_THREADS, smem, s>>>(in + (long long)v0 * ldin, ldin, out + (long long)v0 * ldout, ldout, nlines, LW, LDT, ax);
        else dft_z_r2c_kernel<1><<<grid, DFT_THREADS, smem, s>>>(in + (long long)v0 * ldin, ldin, out + (long long)v0 * ldout, ldout, nlines, LW, LDT, ax);
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

int dft_cols(cudaStream_t s, double2* data, long long vec_stride, int nvec, int nouter, long long outer_stride,
             int row_stride, int ncols, const DftAxis& ax, const double* wtab, int wld)
{
    const long long NVll = (long long)nouter * ncols;
    ARG_CHECK(NVll < (1LL << 24));                      // fast_div range
    const int NV = (int)NVll;
    const int NT = dft_nt(ax);
    const size_t want = NT == 2 ? SMEM_TWO_CTAS : SMEM_THREE_CTAS;
    int CW = NT == 2 ? 64 : 32;
    while (CW > 8 && (smem_cols(ax, CW) > want || CW / 2 >= NV)) CW /= 2;
    const size_t smem = smem_cols(ax, CW);
    if (smem > SMEM_ONE_CTA) { isdf_set_error("dft_cols: n=%d does not fit shared memory", ax.n); return ISDF_ERR_ARG; }
    int rc;
    const void* kern = wtab ? (NT == 2 ? (const void*)dft_cols_kernel<1, 2> : (const void*)dft_cols_kernel<1, 1>)
                            : (NT == 2 ? (const void*)dft_cols_kernel<0, 2> : (const void*)dft_cols_kernel<0, 1>);
    if ((rc = isdf_func_max_smem(kern, smem))) return rc;
    for (int v0 = 0; v0 < nvec; v0 += 65535) {
        const int nv = std::min(65535, nvec - v0);
        dim3 grid((NV + CW - 1) / CW, nv);
        double2* d = data + (long long)v0 * vec_stride;
        if (wtab) {
            if (NT == 2) dft_cols_kernel<1, 2><<<grid, DFT_THREADS, smem, s>>>(d, vec_stride, outer_stride, row_stride, ncols, NV, CW, ax, wtab, wld);
            else dft_cols_kernel<1, 1><<<grid, DFT_THREADS, smem, s>>>(d, vec_stride, outer_stride, row_stride, ncols, NV, CW, ax, wtab, wld);
        } else {
            if (NT == 2) dft_cols_kernel<0, 2><<<grid, DFT_THREADS, smem, s>>>(d, vec_stride, outer_stride, row_stride, ncols, NV, CW, ax, nullptr, 0);
            else dft_cols_kernel<0, 1><<<grid, DFT_THREADS, smem, s>>>(d, vec_stride, outer_stride, row_stride, ncols, NV, CW, ax, nullptr, 0);
        }
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

int dft_coul_weights(cudaStream_t s, const CoulGeom& geo, int nx, int ncl, double* w)
{
    dim3 grid((ncl + 127) / 128, nx);
    coul_weight_kernel<<<grid, 128, 0, s>>>(geo, nx, ncl, w);
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

}  // namespace isdf_fft
