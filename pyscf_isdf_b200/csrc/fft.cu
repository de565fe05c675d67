// Stage 4: batched 3-D real FFT of the auxiliary functions fused with the
// Coulomb kernel 4 pi/|G|^2, and the Coulomb metric W.
//
// Conventions restated from the reference: forward DFT unnormalised, inverse
// carries 1/N (pyscf/pbc/tools/pbc.py:149-211); G vectors in FFT index order
// (pyscf/pbc/gto/cell.py:523-596); coulG = 4 pi/|G|^2 with G=0 -> 0
// (pyscf/pbc/tools/pbc.py:304,352-356).
//
// Hand-written mixed-radix Stockham FFTs in shared memory (radix 4/2/3/5/7 and
// a direct DFT for any other prime factor, so odd and prime meshes such as
// 35^3 or 71^3 work).  A 3-D transform is three passes, each reading and
// writing coalesced row segments:
//   z: real lines -> half spectrum (nz/2+1)          [R2C]
//   y: columns of every (x) plane
//   x: columns of the whole [nx, ny*nzh] matrix, with the Coulomb-kernel
//      multiply fused (and, for V, the inverse x transform in the same kernel).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <vector>
#include <cstdlib>
#include <cmath>
#include <algorithm>
#include "fft_kernels.cuh"

namespace {

__global__ void symmetrize_upper_kernel(double* __restrict__ W, long long ld, int n)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i < n && j < n && j < i) W[(long long)i * ld + j] = W[(long long)j * ld + i];
}

// ---- host side -----------------------------------------------------------------
struct Plans {
    FftPlan ax[3];
    double2* d_tw;      // one allocation: twiddles, Rader kernels, permutations, DFT tables
    DftAxis dft[3];     // dense-DFT form of an axis (dft_gemm.cu); use_dft[d] says whether the forward pass takes it
    bool use_dft[3];
};

// ISDF_FFT_FUSED=1 (opt-in): Rader lines through rader_fused_smem (fft_core.cuh)
int fft_fused_requested()
{
    const char* e = getenv("ISDF_FFT_FUSED");
    return (e && atoi(e) > 0) ? 1 : 0;
}

// What a handle keeps between calls: the plans of the last mesh (device tables allocated once with cudaMalloc)
// and the last Coulomb-weight table.  A build calls the transforms dozens of times on one mesh; allocating,
// uploading and synchronising per call showed up as occasional 0.3 s stalls in the stage (round 1).
struct FftCache {
    int mesh[3] = {0, 0, 0};
    int fused = -1;
    Plans P;
    void* d_tab = nullptr; size_t tab_cap = 0;
    double* d_w = nullptr; size_t w_cap = 0;          // sqrt(w coulG) table of the x pass
    CoulGeom w_geo; int w_nx = 0, w_ncl = 0; bool w_valid = false;
};

void fft_cache_free(void* p)
{
    FftCache* c = (FftCache*)p;
    if (c->d_tab) cudaFree(c->d_tab);
    if (c->d_w) cudaFree(c->d_w);
    delete c;
}

FftCache* cache_of(isdf_handle* h)
{
    if (!h->fft_cache) { h->fft_cache = new FftCache(); h->fft_cache_free = fft_cache_free; }
    return (FftCache*)h->fft_cache;
}

int make_plans(isdf_handle* h, const int* mesh, Plans& P)
{
    FftCache* fc = cache_of(h);
    const int fused = fft_fused_requested();
    if (fc->d_tab && fc->fused == fused && fc->mesh[0] == mesh[0] && fc->mesh[1] == mesh[1] && fc->mesh[2] == mesh[2]) {
        P = fc->P;
        return ISDF_OK;
    }
    std::vector<double2> tw;          // complex tables, offsets in double2 units
    std::vector<int> iv;              // integer tables
    AxisOffsets off[3];
    for (int d = 0; d < 3; ++d) {
        const int n = mesh[d];
        ARG_CHECK(n >= 1 && n <= 1024);
        if (build_axis_tables(n, fused, P.ax[d], tw, iv, off[d])) {
            isdf_set_error("mesh size %d has an unsupported prime factor", n);
            return ISDF_ERR_ARG;
        }
    }
    std::vector<double> dtab;         // cosine / sine tables of the dense-DFT passes
    size_t off_c[3] = {0, 0, 0}, off_s[3] = {0, 0, 0};
    for (int d = 0; d < 3; ++d) {
        P.use_dft[d] = dft_enabled(mesh[d]);
        if (P.use_dft[d]) dft_build_tables(mesh[d], P.dft[d], dtab, off_c[d], off_s[d]);
    }
    const size_t bytes_tw = align_up(sizeof(double2) * tw.size(), 256);
    const size_t bytes_iv = align_up(sizeof(int) * std::max<size_t>(iv.size(), 1), 256);
    const size_t bytes_dft = sizeof(double) * std::max<size_t>(dtab.size(), 1);
    const size_t need = bytes_tw + bytes_iv + bytes_dft;
    // the previous tables may still be in use by kernels in flight on this handle's stream
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    fc->fused = -1;
    if (need > fc->tab_cap) {
        if (fc->d_tab) CUDA_TRY(cudaFree(fc->d_tab));
        fc->d_tab = nullptr; fc->tab_cap = 0;
        CUDA_TRY(cudaMalloc(&fc->d_tab, need));
        fc->tab_cap = need;
    }
    P.d_tw = (double2*)fc->d_tab;
    int* d_iv = (int*)((char*)P.d_tw + bytes_tw);
    double* d_dft = (double*)((char*)P.d_tw + bytes_tw + bytes_iv);
    CUDA_TRY(cudaMemcpyAsync(P.d_tw, tw.data(), sizeof(double2) * tw.size(), cudaMemcpyHostToDevice, h->stream));
    if (!iv.empty())
        CUDA_TRY(cudaMemcpyAsync(d_iv, iv.data(), sizeof(int) * iv.size(), cudaMemcpyHostToDevice, h->stream));
    if (!dtab.empty())
        CUDA_TRY(cudaMemcpyAsync(d_dft, dtab.data(), sizeof(double) * dtab.size(), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));          // the host vectors are locals
    for (int d = 0; d < 3; ++d) {
        bind_axis_tables(P.ax[d], P.d_tw, d_iv, off[d]);
        if (P.use_dft[d]) { P.dft[d].C = d_dft + off_c[d]; P.dft[d].S = d_dft + off_s[d]; }
    }
    fc->P = P; fc->fused = fused;
    fc->mesh[0] = mesh[0]; fc->mesh[1] = mesh[1]; fc->mesh[2] = mesh[2];
    fc->w_valid = false;
    return ISDF_OK;
}

// sqrt(w coulG) for the x pass over the column slab geo.col0 .. +ncl: the handle's cached table when the
// geometry is the one it was filled for, else refilled (stream-ordered; the old content has no reader left
// once the fill kernel is enqueued behind them on the same stream)
int coul_weight_table(isdf_handle* h, const CoulGeom& geo, int nx, int ncl, double** out)
{
    FftCache* fc = cache_of(h);
    const size_t need = sizeof(double) * (size_t)nx * ncl;
    if (fc->w_valid && fc->w_nx == nx && fc->w_ncl == ncl && memcmp(&fc->w_geo, &geo, sizeof(geo)) == 0) {
        *out = fc->d_w;
        return ISDF_OK;
    }
    fc->w_valid = false;
    if (need > fc->w_cap) {
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        if (fc->d_w) CUDA_TRY(cudaFree(fc->d_w));
        fc->d_w = nullptr; fc->w_cap = 0;
        CUDA_TRY(cudaMalloc((void**)&fc->d_w, need));
        fc->w_cap = need;
    }
    int rc;
    if ((rc = dft_coul_weights(h->stream, geo, nx, ncl, fc->d_w))) return rc;
    h->other_launches++;
    memcpy(&fc->w_geo, &geo, sizeof(geo));
    fc->w_nx = nx; fc->w_ncl = ncl; fc->w_valid = true;
    *out = fc->d_w;
    return ISDF_OK;
}

template <typename K>
int set_smem(K kern, size_t smem)
{
    // per (kernel, device), raised only: several handles may launch different widths at once
    return isdf_func_max_smem((const void*)kern, smem);
}

// launch the FUSED = false instantiation, or hand a fused plan (FftPlan::rader >= 2) to fft_fused.cu
#define FFT_RUN_Z(fused_, K_, FK_, grid_, smem_, stream_, ...)                                               \
    do {                                                                                                     \
        int rc_;                                                                                             \
        if (fused_) { if ((rc_ = FK_(grid_, smem_, stream_, __VA_ARGS__))) return rc_; }                     \
        else { if ((rc_ = set_smem(K_<false>, smem_))) return rc_;                                           \
               K_<false><<<grid_, FFT_THREADS, smem_, stream_>>>(__VA_ARGS__); }                             \
    } while (0)
#define FFT_RUN_COLS(fused_, MODE_, grid_, smem_, stream_, ...)                                              \
    do {                                                                                                     \
        int rc_;                                                                                             \
        if (fused_) { if ((rc_ = fused_cols(MODE_, grid_, smem_, stream_, __VA_ARGS__))) return rc_; }       \
        else { if ((rc_ = set_smem(fft_cols_kernel<MODE_, false>, smem_))) return rc_;                       \
               fft_cols_kernel<MODE_, false><<<grid_, FFT_THREADS, smem_, stream_>>>(__VA_ARGS__); }         \
    } while (0)

void fill_geom(CoulGeom& g, const int* mesh, const double* a, double scale)
{
    memset(&g, 0, sizeof(g));             // padding bytes too: the weight-table cache compares the struct bytewise
    // reciprocal vectors b = 2 pi inv(a^T): rows b_i with a_i . b_j = 2 pi delta_ij
    const double a00 = a[0], a01 = a[1], a02 = a[2], a10 = a[3], a11 = a[4], a12 = a[5], a20 = a[6], a21 = a[7], a22 = a[8];
    const double det = a00 * (a11 * a22 - a12 * a21) - a01 * (a10 * a22 - a12 * a20) + a02 * (a10 * a21 - a11 * a20);
    const double f = 2.0 * 3.14159265358979323846 / det;
    // b_0 = 2pi (a1 x a2)/det, etc.
    g.b[0] = f * (a11 * a22 - a12 * a21); g.b[1] = f * (a12 * a20 - a10 * a22); g.b[2] = f * (a10 * a21 - a11 * a20);
    g.b[3] = f * (a21 * a02 - a22 * a01); g.b[4] = f * (a22 * a00 - a20 * a02); g.b[5] = f * (a20 * a01 - a21 * a00);
    g.b[6] = f * (a01 * a12 - a02 * a11); g.b[7] = f * (a02 * a10 - a00 * a12); g.b[8] = f * (a00 * a11 - a01 * a10);
    for (int d = 0; d < 3; ++d) g.mesh[d] = mesh[d];
    g.nzh = mesh[2] / 2 + 1;
    g.scale = scale;
    g.col0 = 0;
}

// shared memory per CTA of the FFT kernels (two ping-pong line buffers); smaller = more CTAs per SM.
// Tuning knob: environment ISDF_FFT_SMEM_KB.
// shared memory per CTA: column passes (y, x) want long contiguous row segments (wide CTAs), the z
// pass reads whole lines and prefers more resident CTAs (tuning knobs ISDF_FFT_SMEM_KB / _Z_KB)
size_t fft_smem_budget(bool zpass = false)
{
    struct Budget {
        size_t v[2];
        Budget() {
            const char* e = getenv("ISDF_FFT_SMEM_KB");
            const char* ez = getenv("ISDF_FFT_SMEM_Z_KB");
            const long kb = e ? atol(e) : 64, kbz = ez ? atol(ez) : 64;
            v[0] = (size_t)std::min(200L, std::max(8L, kb)) * 1024;
            v[1] = (size_t)std::min(200L, std::max(8L, kbz)) * 1024;
        }
    };
    static const Budget b;                  // initialised once, thread-safe (C++11 function-local static)
    return b.v[zpass ? 1 : 0];
}

// lines (or columns) per CTA such that two ping-pong buffers of `nalloc` slots per line fit the budget
int pick_lb(int nalloc, size_t budget) { return (int)std::max<size_t>(1, std::min<size_t>(64, budget / (2 * sizeof(double2) * nalloc))); }

// forward 3-D transform of nvec real vectors into half spectra (spec[vec*lds + ...]);
// xmode selects what the x pass fuses (0 plain, 1 sqrt(w coulG), 2 coulG*scale + inverse x)
constexpr int MAX_GRID_VEC = 65535;       // the vector index is a grid y / z dimension

int forward_passes_batch(isdf_handle* h, const Plans& P, const int* mesh, const double* in, long long ldin,
                         int nvec, double2* spec, long long lds, int xmode, const CoulGeom& geo);

int forward_passes(isdf_handle* h, const Plans& P, const int* mesh, const double* in, long long ldin,
                   int nvec, double2* spec, long long lds, int xmode, const CoulGeom& geo)
{
    for (int v0 = 0; v0 < nvec; v0 += MAX_GRID_VEC) {
        const int rc = forward_passes_batch(h, P, mesh, in + (long long)v0 * ldin, ldin, std::min(MAX_GRID_VEC, nvec - v0),
                                            spec + (long long)v0 * lds, lds, xmode, geo);
        if (rc) return rc;
    }
    return ISDF_OK;
}

int forward_passes_batch(isdf_handle* h, const Plans& P, const int* mesh, const double* in, long long ldin,
                         int nvec, double2* spec, long long lds, int xmode, const CoulGeom& geo)
{
    const int nx = mesh[0], ny = mesh[1], nz = mesh[2], nzh = nz / 2 + 1;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    int rc;
    if (P.use_dft[2]) {
        if ((rc = dft_z_r2c(s, in, ldin, spec, lds, nvec, nx * ny, P.dft[2]))) return rc;
    } else {
        const int LB = pick_lb(P.ax[2].lalloc(), fft_smem_budget(true));
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        dim3 grid((nx * ny + 2 * LB - 1) / (2 * LB), nvec);
        FFT_RUN_Z((P.ax[2].rader >= 2), fft_z_r2c_kernel, fused_z_r2c, grid, smem, s, in, ldin, spec, lds, P.ax[2], nx * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    if (P.use_dft[1]) {
        if ((rc = dft_cols(s, spec, lds, nvec, nx, (long long)ny * nzh, nzh, nzh, P.dft[1], nullptr, 0))) return rc;
    } else {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        dim3 grid((nzh + CH - 1) / CH, nx, nvec);
        FFT_RUN_COLS((P.ax[1].rader >= 2), 0, grid, smem, s, spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    if (P.use_dft[0] && xmode != 2) {
        const int ncols = ny * nzh;
        double* wtab = nullptr;
        if (xmode == 1 && (rc = coul_weight_table(h, geo, nx, ncols, &wtab))) return rc;
        if ((rc = dft_cols(s, spec, lds, nvec, 1, 0, ncols, ncols, P.dft[0], wtab, ncols))) return rc;
    } else {
        const int ncols = ny * nzh;
        const int CH = std::min(pick_lb(P.ax[0].lalloc(), budget), ncols);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[0].lalloc();
        dim3 grid((ncols + CH - 1) / CH, 1, nvec);
        if (xmode == 1) FFT_RUN_COLS((P.ax[0].rader >= 2), 1, grid, smem, s, spec, lds, 0, P.ax[0], ncols, CH, geo);
        else if (xmode == 2) FFT_RUN_COLS((P.ax[0].rader >= 2), 2, grid, smem, s, spec, lds, 0, P.ax[0], ncols, CH, geo);
        else FFT_RUN_COLS((P.ax[0].rader >= 2), 0, grid, smem, s, spec, lds, 0, P.ax[0], ncols, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 3;
    return ISDF_OK;
}

// inverse y and z passes (x already inverted): spectra -> real vectors
int inverse_yz_batch(isdf_handle* h, const Plans& P, const int* mesh, double2* spec, long long lds,
                     int nvec, double* out, long long ldout, const CoulGeom& geo);

int inverse_yz(isdf_handle* h, const Plans& P, const int* mesh, double2* spec, long long lds,
               int nvec, double* out, long long ldout, const CoulGeom& geo)
{
    for (int v0 = 0; v0 < nvec; v0 += MAX_GRID_VEC) {
        const int rc = inverse_yz_batch(h, P, mesh, spec + (long long)v0 * lds, lds, std::min(MAX_GRID_VEC, nvec - v0),
                                        out + (long long)v0 * ldout, ldout, geo);
        if (rc) return rc;
    }
    return ISDF_OK;
}

int inverse_yz_batch(isdf_handle* h, const Plans& P, const int* mesh, double2* spec, long long lds,
                     int nvec, double* out, long long ldout, const CoulGeom& geo)
{
    const int nx = mesh[0], ny = mesh[1], nz = mesh[2], nzh = nz / 2 + 1;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        dim3 grid((nzh + CH - 1) / CH, nx, nvec);
        FFT_RUN_COLS((P.ax[1].rader >= 2), 3, grid, smem, s, spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    {
        const int LB = pick_lb(P.ax[2].lalloc(), budget);
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        dim3 grid((nx * ny + LB - 1) / LB, nvec);
        FFT_RUN_Z((P.ax[2].rader >= 2), fft_z_c2r_kernel, fused_z_c2r, grid, smem, s, spec, lds, out, ldout, P.ax[2], nx * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 2;
    return ISDF_OK;
}

}  // namespace

extern "C" size_t isdf_build_W_ws_bytes(int np, const int* h_mesh)
{
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    return sizeof(double2) * ngh * (size_t)np + 4096;
}

extern "C" int isdf_coulomb_apply(isdf_handle* h, const double* in, long long ldin, int nvec,
                                  const int* h_mesh, const double* h_a,
                                  double* out, long long ldout, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && in && out && h_mesh && h_a && ws && nvec > 0);
    const long long ng = (long long)h_mesh[0] * h_mesh[1] * h_mesh[2];
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    ARG_CHECK(ldin >= ng && ldout >= ng);
    const size_t per_vec = sizeof(double2) * ngh;
    const int batch = (int)std::min<size_t>(nvec, ws_bytes / per_vec);
    if (batch < 1) { isdf_set_error("coulomb_apply workspace too small: need >= %zu", per_vec); return ISDF_ERR_WS; }
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0 / (double)ng);
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FFT], h->stream));
    for (int v0 = 0; v0 < nvec; v0 += batch) {
        const int nv = std::min(batch, nvec - v0);
        if ((rc = forward_passes(h, P, h_mesh, in + (long long)v0 * ldin, ldin, nv, (double2*)ws, (long long)ngh, 2, geo))) return rc;
        if ((rc = inverse_yz(h, P, h_mesh, (double2*)ws, (long long)ngh, nv, out + (long long)v0 * ldout, ldout, geo))) return rc;
    }
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FFT], h->stream));
    h->ev_used[ISDF_STAGE_FFT] = true;
    return ISDF_OK;
}

extern "C" int isdf_build_W(isdf_handle* h, const double* zeta, int np, long long ldz,
                            const int* h_mesh, const double* h_a,
                            double* W, double* V, long long ldv,
                            void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && zeta && h_mesh && h_a && W && ws && np > 0);
    const long long ng = (long long)h_mesh[0] * h_mesh[1] * h_mesh[2];
    const size_t ngh = (size_t)h_mesh[0] * h_mesh[1] * (h_mesh[2] / 2 + 1);
    ARG_CHECK(ldz >= ng);
    const size_t need = isdf_build_W_ws_bytes(np, h_mesh);
    if (need > ws_bytes) { isdf_set_error("build_W workspace too small: need %zu have %zu", need, ws_bytes); return ISDF_ERR_WS; }
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_COULOMB], s));
    if (V != nullptr) {
        ARG_CHECK(ldv >= ng);
        if ((rc = isdf_coulomb_apply(h, zeta, ldz, np, h_mesh, h_a, V, ldv, ws, ws_bytes))) return rc;
    }
    Plans P;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    const double det = h_a[0] * (h_a[4] * h_a[8] - h_a[5] * h_a[7]) - h_a[1] * (h_a[3] * h_a[8] - h_a[5] * h_a[6]) +
                       h_a[2] * (h_a[3] * h_a[7] - h_a[4] * h_a[6]);
    const double vol = fabs(det);
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0);
    double2* Y = (double2*)ws;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FFT], s));
    if ((rc = forward_passes(h, P, h_mesh, zeta, ldz, np, Y, (long long)ngh, 1, geo))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FFT], s));
    h->ev_used[ISDF_STAGE_FFT] = true;
    // W = (vol/ng^2) Y Y^T over the weighted half spectrum (upper triangle, then mirrored)
    GemmParams p;
    p.M = np; p.N = np; p.K = (int)(2 * ngh);
    p.A = (const double*)Y; p.lda = 2 * (long long)ngh; p.a_kmajor = 0;
    p.B = (const double*)Y; p.ldb = 2 * (long long)ngh; p.b_kmajor = 0;
    p.C = W; p.ldc = np; p.Cin = nullptr; p.ldcin = 0; p.E = nullptr; p.lde = 0;
    p.alpha = vol / ((double)ng * (double)ng); p.neg = 0; p.epi = EPI_STORE; p.upper_only = 1;
    if ((rc = gemm_launch(h, p))) return rc;
    dim3 blk(32, 8), grd((np + 31) / 32, (np + 7) / 8);
    symmetrize_upper_kernel<<<grd, blk, 0, s>>>(W, np, np);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_COULOMB], s));
    h->ev_used[ISDF_STAGE_COULOMB] = true;
    return ISDF_OK;
}

// ---- slab-decomposed 3-D transform (grid sharded over ranks by x-planes) --------------------
// Each rank holds x-planes [x_off[r], x_off[r+1]) of every vector.  z and y passes are local;
// one all-to-all (done by the caller, NCCL) re-shards the half spectra from x-planes to
// (y,z)-column slabs; the x pass with the fused Coulomb weight then runs on whole x lines.
namespace {

__global__ void fft_pack_kernel(const double2* __restrict__ spec, long long lds, int nxl, int ncols,
                                int c0, int ncl, double2* __restrict__ send)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncl) return;
    const int x = blockIdx.y, v = blockIdx.z;
    send[((long long)v * nxl + x) * ncl + c] = spec[(long long)v * lds + (long long)x * ncols + c0 + c];
}

__global__ void fft_unpack_kernel(const double2* __restrict__ recv, int nxl, int x0, int ncl,
                                  double2* __restrict__ Y, long long ldy)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncl) return;
    const int x = blockIdx.y, v = blockIdx.z;
    Y[(long long)v * ldy + (long long)(x0 + x) * ncl + c] = recv[((long long)v * nxl + x) * ncl + c];
}

}  // namespace

// local z (R2C) and y passes: in[nvec, ldin] real, x-planes of this rank -> spec[nvec][nxl][ny*nzh] complex
extern "C" int isdf_fft_zy(isdf_handle* h, const double* in, long long ldin, int nvec, int nxl,
                           const int* h_mesh, void* spec, long long lds)
{
    ARG_CHECK(h && in && h_mesh && spec && nvec > 0 && nxl > 0 && nxl <= h_mesh[0]);
    const int ny = h_mesh[1], nz = h_mesh[2], nzh = nz / 2 + 1;
    ARG_CHECK(ldin >= (long long)nxl * ny * nz && lds >= (long long)nxl * ny * nzh);
    if (nvec > MAX_GRID_VEC) {               // the vector index is a grid dimension: batches of <= 65535
        for (int v0 = 0; v0 < nvec; v0 += MAX_GRID_VEC) {
            const int rcb = isdf_fft_zy(h, in + (long long)v0 * ldin, ldin, std::min(MAX_GRID_VEC, nvec - v0), nxl, h_mesh,
                                        (double2*)spec + (long long)v0 * lds, lds);
            if (rcb) return rcb;
        }
        return ISDF_OK;
    }
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    cudaStream_t s = h->stream;
    const size_t budget = fft_smem_budget();
    CoulGeom geo; memset(&geo, 0, sizeof(geo));
    if (P.use_dft[2]) {
        if ((rc = dft_z_r2c(s, in, ldin, (double2*)spec, lds, nvec, nxl * ny, P.dft[2]))) return rc;
    } else {
        const int LB = pick_lb(P.ax[2].lalloc(), fft_smem_budget(true));
        const size_t smem = 2 * sizeof(double2) * LB * P.ax[2].lalloc();
        dim3 grid((nxl * ny + 2 * LB - 1) / (2 * LB), nvec);
        FFT_RUN_Z((P.ax[2].rader >= 2), fft_z_r2c_kernel, fused_z_r2c, grid, smem, s, in, ldin, (double2*)spec, lds, P.ax[2], nxl * ny, LB);
        CUDA_TRY(cudaGetLastError());
    }
    if (P.use_dft[1]) {
        if ((rc = dft_cols(s, (double2*)spec, lds, nvec, nxl, (long long)ny * nzh, nzh, nzh, P.dft[1], nullptr, 0))) return rc;
    } else {
        const int CH = std::min(pick_lb(P.ax[1].lalloc(), budget), nzh);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[1].lalloc();
        dim3 grid((nzh + CH - 1) / CH, nxl, nvec);
        FFT_RUN_COLS((P.ax[1].rader >= 2), 0, grid, smem, s, (double2*)spec, lds, (long long)ny * nzh, P.ax[1], nzh, CH, geo);
        CUDA_TRY(cudaGetLastError());
    }
    h->other_launches += 2;
    return ISDF_OK;
}

// send buffer for the all-to-all: block of rank r = [nvec][nxl][h_col_off[r+1]-h_col_off[r]], blocks in rank order
extern "C" int isdf_fft_pack(isdf_handle* h, const void* spec, long long lds, int nvec, int nxl, int ncols,
                             const int* h_col_off, int nranks, void* send)
{
    ARG_CHECK(h && spec && h_col_off && send && nranks >= 1 && h_col_off[0] == 0 && h_col_off[nranks] == ncols);
    for (int r = 0; r < nranks; ++r) {
        const int c0 = h_col_off[r], ncl = h_col_off[r + 1] - c0;
        if (ncl <= 0) continue;
        double2* dst = (double2*)send + (long long)nvec * nxl * c0;
        dim3 grid((ncl + 127) / 128, nxl, nvec);
        fft_pack_kernel<<<grid, 128, 0, h->stream>>>((const double2*)spec, lds, nxl, ncols, c0, ncl, dst);
        h->other_launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

// receive buffer (block of rank r = [nvec][nx_r][ncl]) -> Y[nvec][nx][ncl] (row stride ldy, complex units)
extern "C" int isdf_fft_unpack(isdf_handle* h, const void* recv, int nvec, const int* h_x_off, int nranks,
                               int ncl, void* Y, long long ldy)
{
    ARG_CHECK(h && recv && h_x_off && Y && nranks >= 1 && ncl > 0 && h_x_off[0] == 0);
    ARG_CHECK(ldy >= (long long)h_x_off[nranks] * ncl);
    for (int r = 0; r < nranks; ++r) {
        const int x0 = h_x_off[r], nxl = h_x_off[r + 1] - x0;
        if (nxl <= 0) continue;
        const double2* src = (const double2*)recv + (long long)nvec * ncl * x0;
        dim3 grid((ncl + 127) / 128, nxl, nvec);
        fft_unpack_kernel<<<grid, 128, 0, h->stream>>>(src, nxl, x0, ncl, (double2*)Y, ldy);
        h->other_launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

// x pass over the column slab [col0, col0+ncl) of the (y,z) plane, fused with the
// sqrt(weight * coulG) factor of the W build: Y[nvec][nx][ncl] in place.
extern "C" int isdf_fft_x_coulomb(isdf_handle* h, void* Y, long long ldy, int nvec,
                                  const int* h_mesh, const double* h_a, int col0, int ncl)
{
    ARG_CHECK(h && Y && h_mesh && h_a && nvec > 0 && ncl > 0 && col0 >= 0);
    const int nx = h_mesh[0];
    ARG_CHECK(col0 + ncl <= h_mesh[1] * (h_mesh[2] / 2 + 1) && ldy >= (long long)nx * ncl);
    if (nvec > MAX_GRID_VEC) {
        for (int v0 = 0; v0 < nvec; v0 += MAX_GRID_VEC) {
            const int rcb = isdf_fft_x_coulomb(h, (double2*)Y + (long long)v0 * ldy, ldy, std::min(MAX_GRID_VEC, nvec - v0),
                                               h_mesh, h_a, col0, ncl);
            if (rcb) return rcb;
        }
        return ISDF_OK;
    }
    Plans P; int rc;
    if ((rc = make_plans(h, h_mesh, P))) return rc;
    CoulGeom geo; fill_geom(geo, h_mesh, h_a, 1.0);
    geo.col0 = col0;
    if (P.use_dft[0]) {
        double* wtab = nullptr;
        if ((rc = coul_weight_table(h, geo, nx, ncl, &wtab))) return rc;
        if ((rc = dft_cols(h->stream, (double2*)Y, ldy, nvec, 1, 0, ncl, ncl, P.dft[0], wtab, ncl))) return rc;
    } else {
        const size_t budget = fft_smem_budget();
        const int CH = std::min(pick_lb(P.ax[0].lalloc(), budget), ncl);
        const size_t smem = 2 * sizeof(double2) * CH * P.ax[0].lalloc();
        dim3 grid((ncl + CH - 1) / CH, 1, nvec);
        FFT_RUN_COLS((P.ax[0].rader >= 2), 1, grid, smem, h->stream, (double2*)Y, ldy, 0, P.ax[0], ncl, CH, geo);
    }
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}

extern "C" int isdf_symmetrize_upper(isdf_handle* h, double* W, long long ld, int n)
{
    ARG_CHECK(h && W && n > 0 && ld >= n);
    dim3 blk(32, 8), grd((n + 31) / 32, (n + 7) / 8);
    symmetrize_upper_kernel<<<grd, blk, 0, h->stream>>>(W, ld, n);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}
