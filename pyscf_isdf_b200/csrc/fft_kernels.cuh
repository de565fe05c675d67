// The FFT kernels of stage 4 (z pass R2C / C2R, column passes with the fused Coulomb multiply) as
// templates on FUSED, included by fft.cu (FUSED = false: the kernels measured in round 1) and by
// fft_fused.cu (FUSED = true: Rader lines through rader_fused_smem).  Two translation units on purpose:
// with both sets of instantiations in one module nvcc stops inlining the line-FFT routines into the
// default kernels (the plan is then copied to local memory: 200-byte stack frames).
#pragma once
#include "fft_core.cuh"

namespace isdf_fft {

struct CoulGeom {
    int mesh[3];
    int nzh;
    double b[9];        // reciprocal vectors (rows)
    double scale;       // multiplies the kernel (1/ng for V)
    int col0;           // first (y,z)-column of this rank's slab in the x pass (0 when not sharded)
};

// launchers of the FUSED = true instantiations (defined in fft_fused.cu); they set the dynamic
// shared-memory attribute and launch on `s`; return 0 or ISDF_ERR_CUDA
int fused_z_r2c(dim3 grid, size_t smem, cudaStream_t s, const double* in, long long ldin, double2* out,
                long long ldout, FftPlan pl, int nlines, int LB);
int fused_z_c2r(dim3 grid, size_t smem, cudaStream_t s, const double2* in, long long ldin, double* out,
                long long ldout, FftPlan pl, int nlines, int LB);
int fused_cols(int mode, dim3 grid, size_t smem, cudaStream_t s, double2* data, long long vec_stride,
               long long outer_stride, FftPlan pl, int ncols, int CH, CoulGeom geo);

// ---- forward passes as dense DFTs on the FP64 tensor pipe (dft_gemm.cu) ----
constexpr int DFT_NMAX = 128;
struct DftAxis {
    int n, ne, hp;          // length, n/2 + 1 output / folded-input rows, (n-1)/2 mirrored pairs
    int MT, KP, LDA;        // 8-row tiles of the tables, inner dimension padded to 4, (unused) row stride
    const double* C;        // device table in DMMA fragment order: (cos, sin)(2 pi k j / n) interleaved, see dft_gemm.cu
    const double* S;        // same pointer (one table)
};
bool dft_enabled(int n);
void dft_build_tables(int n, DftAxis& ax, std::vector<double>& tab, size_t& off_c, size_t& off_s);
int dft_z_r2c(cudaStream_t s, const double* in, long long ldin, double2* out, long long ldout, int nvec,
              int nlines, const DftAxis& ax);
int dft_cols(cudaStream_t s, double2* data, long long vec_stride, int nvec, int nouter, long long outer_stride,
             int row_stride, int ncols, const DftAxis& ax, const double* wtab, int wld);
int dft_coul_weights(cudaStream_t s, const CoulGeom& geo, int nx, int ncl, double* w);

}  // namespace isdf_fft

namespace {


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
template <bool FUSED>
__global__ void __launch_bounds__(FFT_THREADS, FUSED ? 3 : 0)
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
    double2* r = fft_smem<-1, FUSED>(a, b, pl, np, LS, 1, false);
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
template <bool FUSED>
__global__ void __launch_bounds__(FFT_THREADS, FUSED ? 3 : 0)
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
    double2* r = fft_smem<+1, FUSED>(a, b, pl, nl, LS, 1, false);
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
template <int MODE, bool FUSED>
__global__ void __launch_bounds__(FFT_THREADS, FUSED ? 3 : 0)
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
    if (MODE == 3) r = fft_smem<+1, FUSED>(a, b, pl, nc, 1, CH, true);
    else r = fft_smem<-1, FUSED>(a, b, pl, nc, 1, CH, true);
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
            r = fft_smem<+1, FUSED>(r, other, pl, nc, 1, CH, true);
        }
    }
    for (int idx = threadIdx.x; idx < n * nc; idx += blockDim.x) {
        const int i = fast_div(idx, nc, inv_nc), c = idx - i * nc;
        base[(long long)i * ncols + c] = r[i * CH + c];
    }
}

}  // namespace
