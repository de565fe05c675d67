// Stage 5: ISDF Coulomb / exchange contraction, once per SCF iteration.
//   T = D X                      [nao, np]      X = Phi[:, ip]
//   J = X diag(W rho) X^T        rho_P = sum_mu X[mu,P] T[mu,P]
//   K = X ((X^T T) o W) X^T      (Hadamard with W fused into the GEMM epilogue)
// Exact-grid limit: pyscf/pbc/df/fft_jk.py:33-109 (J), :177-302 (K).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <algorithm>

namespace {

// rho[P] = sum_mu X[mu,P] * T[mu,P]
__global__ void coldot_kernel(const double* __restrict__ X, const double* __restrict__ T,
                              int nao, int np, long long ld, double* __restrict__ rho)
{
    const int P = blockIdx.x * blockDim.x + threadIdx.x;
    if (P >= np) return;
    double s = 0.0;
    for (int mu = 0; mu < nao; ++mu) s = fma(X[(long long)mu * ld + P], T[(long long)mu * ld + P], s);
    rho[P] = s;
}

// v = W rho for a block of rows, one warp per row
__global__ void gemv_rows_kernel(const double* __restrict__ W, long long ld, int nrows, int n,
                                 const double* __restrict__ x, double* __restrict__ y)
{
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= nrows) return;
    const double* w = W + (long long)row * ld;
    double s = 0.0;
    for (int j = lane; j < n; j += 32) s = fma(w[j], x[j], s);
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[row] = s;
}

// Y[mu,P] = X[mu,P] * v[P]
__global__ void scale_cols_kernel(const double* __restrict__ X, const double* __restrict__ v,
                                  int nao, int np, long long ldx, long long ldy, double* __restrict__ Y)
{
    const int P = blockIdx.x * blockDim.x + threadIdx.x;
    if (P >= np) return;
    const double f = v[P];
    for (int mu = blockIdx.y; mu < nao; mu += gridDim.y) Y[(long long)mu * ldy + P] = X[(long long)mu * ldx + P] * f;
}

int gemm6(isdf_handle* h, int M, int N, int K, const double* A, long long lda, int ak,
          const double* B, long long ldb, int bk, double* C, long long ldc,
          const double* E, long long lde, int epi)
{
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.A = A; p.lda = lda; p.a_kmajor = ak;
    p.B = B; p.ldb = ldb; p.b_kmajor = bk; p.C = C; p.ldc = ldc; p.Cin = nullptr; p.ldcin = 0;
    p.E = E; p.lde = lde; p.alpha = 1.0; p.neg = 0; p.epi = epi; p.upper_only = 0;
    return gemm_launch(h, p);
}

}  // namespace

extern "C" size_t isdf_get_jk_block_ws_bytes(int nao, int np, int npl)
{
    return sizeof(double) * ((size_t)nao * np + 2 * (size_t)nao * npl + (size_t)npl * np + (size_t)np + npl) + 8192;
}

extern "C" size_t isdf_get_jk_ws_bytes(int nao, int np) { return isdf_get_jk_block_ws_bytes(nao, np, np); }

namespace {

// J and K contributions of the row block once the two [r, np] factors of M = L^T R are known
// (M = X^T D X: L = X, R = D X, r = nao;  occupied orbitals: L = R = T' = (C sqrt(n))^T X, r = nocc):
//   rho_P = sum_i L[i,P] R[i,P];  J_part = X_b diag(W_b rho) X_b^T;  K_part = X_b ((L_b^T R) o W_b) X^T
int jk_from_factors(isdf_handle* h, const double* X, const double* Wrows, long long ldw, int p0, int npl,
                    int nao, int np, const double* L, const double* R, int r, int with_j, int with_k,
                    double* vj, double* vk, double* Ub, double* Vb, double* Nm, double* rho, double* v)
{
    cudaStream_t s = h->stream;
    const double* Xb = X + p0;                         // columns p0.. of X (row stride np)
    int rc;
    if (with_j) {
        coldot_kernel<<<(np + 255) / 256, 256, 0, s>>>(L, R, r, np, np, rho);
        gemv_rows_kernel<<<(npl + 7) / 8, 256, 0, s>>>(Wrows, ldw, npl, np, rho, v);
        dim3 g((npl + 255) / 256, std::min(nao, 1024));
        scale_cols_kernel<<<g, 256, 0, s>>>(Xb, v, nao, npl, np, npl, Ub);
        CUDA_TRY(cudaGetLastError());
        h->other_launches += 3;
        // J_part = Ub X_b^T
        if ((rc = gemm6(h, nao, nao, npl, Ub, npl, 0, Xb, np, 0, vj, nao, nullptr, 0, EPI_STORE))) return rc;
    }
    if (with_k) {
        // N_b = (L_b^T R) o W_b ; V_b = N_b X^T ; K_part = X_b V_b
        if ((rc = gemm6(h, npl, np, r, L + p0, np, 1, R, np, 1, Nm, np, Wrows, ldw, EPI_MUL))) return rc;
        if ((rc = gemm6(h, npl, nao, np, Nm, np, 0, X, np, 0, Vb, nao, nullptr, 0, EPI_STORE))) return rc;
        if ((rc = gemm6(h, nao, nao, npl, Xb, np, 0, Vb, nao, 1, vk, nao, nullptr, 0, EPI_STORE))) return rc;
    }
    return ISDF_OK;
}

}  // namespace

// Contribution of the interpolation points [p0, p0+npl) (a row block of W) to J and K:
//   J_part = X_b diag(W_b rho) X_b^T ,  K_part = X_b ((X_b^T D X) o W_b) X^T
// Summed over disjoint blocks this is the full J/K (the caller all-reduces).
extern "C" int isdf_get_jk_block(isdf_handle* h, const double* X, const double* Wrows, long long ldw,
                                 int p0, int npl, const double* dm, int nset, int nao, int np,
                                 int with_j, int with_k, double* vj, double* vk,
                                 void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && X && Wrows && dm && ws && nset > 0 && nao > 0 && np > 0);
    ARG_CHECK(p0 >= 0 && npl > 0 && p0 + npl <= np && ldw >= np);
    ARG_CHECK((!with_j || vj) && (!with_k || vk));
    if (isdf_get_jk_block_ws_bytes(nao, np, npl) > ws_bytes) {
        isdf_set_error("get_jk workspace too small: need %zu have %zu", isdf_get_jk_block_ws_bytes(nao, np, npl), ws_bytes);
        return ISDF_ERR_WS;
    }
    WsAlloc a(ws, ws_bytes);
    double* T = a.get<double>((size_t)nao * np);       // D X
    double* Ub = a.get<double>((size_t)nao * npl);     // X_b diag(v)
    double* Vb = a.get<double>((size_t)npl * nao);     // N_b X^T
    double* Nm = a.get<double>((size_t)npl * np);      // (X_b^T T) o W_b
    double* rho = a.get<double>(np);
    double* v = a.get<double>(npl);
    cudaStream_t s = h->stream;
    int rc;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_JK], s));
    for (int i = 0; i < nset; ++i) {
        const double* D = dm + (size_t)i * nao * nao;
        // T = D X
        if ((rc = gemm6(h, nao, np, nao, D, nao, 0, X, np, 1, T, np, nullptr, 0, EPI_STORE))) return rc;
        if ((rc = jk_from_factors(h, X, Wrows, ldw, p0, npl, nao, np, X, T, nao, with_j, with_k,
                                  with_j ? vj + (size_t)i * nao * nao : nullptr, with_k ? vk + (size_t)i * nao * nao : nullptr,
                                  Ub, Vb, Nm, rho, v))) return rc;
    }
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_JK], s));
    h->ev_used[ISDF_STAGE_JK] = true;
    return ISDF_OK;
}

// The same contribution for a density matrix given by its occupied orbitals, D = C diag(n) C^T (the
// mo_coeff / mo_occ tags the reference's K build reads, pyscf/pbc/df/fft_jk.py:206-210, :235-238):
// Cs[nocc, nao] = sqrt(n_i) C[:, i]^T.  M = X^T D X = T'^T T' with T' = Cs X, so the first product shrinks
// from nao x nao x np to nocc x nao x np and the Hadamard product from K = nao to K = nocc.
extern "C" int isdf_get_jk_block_occ(isdf_handle* h, const double* X, const double* Wrows, long long ldw,
                                     int p0, int npl, const double* Cs, int nocc, int nao, int np,
                                     int with_j, int with_k, double* vj, double* vk,
                                     void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && X && Wrows && Cs && ws && nocc > 0 && nocc <= nao && nao > 0 && np > 0);
    ARG_CHECK(p0 >= 0 && npl > 0 && p0 + npl <= np && ldw >= np);
    ARG_CHECK((!with_j || vj) && (!with_k || vk));
    if (isdf_get_jk_block_ws_bytes(nao, np, npl) > ws_bytes) {
        isdf_set_error("get_jk workspace too small: need %zu have %zu", isdf_get_jk_block_ws_bytes(nao, np, npl), ws_bytes);
        return ISDF_ERR_WS;
    }
    WsAlloc a(ws, ws_bytes);
    double* T = a.get<double>((size_t)nao * np);       // T' = Cs X uses the first nocc rows
    double* Ub = a.get<double>((size_t)nao * npl);
    double* Vb = a.get<double>((size_t)npl * nao);
    double* Nm = a.get<double>((size_t)npl * np);
    double* rho = a.get<double>(np);
    double* v = a.get<double>(npl);
    cudaStream_t s = h->stream;
    int rc;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_JK], s));
    if ((rc = gemm6(h, nocc, np, nao, Cs, nao, 0, X, np, 1, T, np, nullptr, 0, EPI_STORE))) return rc;
    if ((rc = jk_from_factors(h, X, Wrows, ldw, p0, npl, nao, np, T, T, nocc, with_j, with_k, vj, vk, Ub, Vb, Nm, rho, v))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_JK], s));
    h->ev_used[ISDF_STAGE_JK] = true;
    return ISDF_OK;
}

extern "C" int isdf_get_jk(isdf_handle* h, const double* X, const double* W,
                           const double* dm, int nset, int nao, int np,
                           int with_j, int with_k, double* vj, double* vk,
                           void* ws, size_t ws_bytes)
{
    return isdf_get_jk_block(h, X, W, np, 0, np, dm, nset, nao, np, with_j, with_k, vj, vk, ws, ws_bytes);
}

// ---- J/K for callers that must not allocate ------------------------------------------------------
// Same contraction as isdf_get_jk, with its scratch in a fixed file-scope device array instead of a
// caller workspace: no cudaMalloc / cudaFree / allocating library routine anywhere on this path (for
// code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The array
// exists once per GPU that loads the library and is held until the process exits.  Every call on a
// device shares it: calls that overlap on the GPU (two streams, two handles of one device) race.
// It keeps the previous call's contents; nothing in it has to start at zero, because every buffer
// carved from it (T, Ub, Vb, Nm, rho, v) is written in full by the kernels of a call before it is read.
#define ISDF_JK_STATIC_SCRATCH_BYTES (32u << 20)
static __device__ __align__(256) unsigned char g_jk_scratch[ISDF_JK_STATIC_SCRATCH_BYTES];

namespace {

constexpr int JK_MAX_DEVICES = 64;
void* g_jk_scratch_ptr[JK_MAX_DEVICES];      // device address of g_jk_scratch per device, set under the set-up lock

// In host code `g_jk_scratch` is only a stand-in for the symbol: the device address comes from the runtime,
// once per device, with that device current.
int resolve_jk_scratch(void* arg)
{
    const int dev = *(const int*)arg;
    void* ptr = nullptr;
    CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_jk_scratch));
    if (ptr == nullptr) { isdf_set_error("cudaGetSymbolAddress returned a null scratch address on device %d", dev); return ISDF_ERR_CUDA; }
    g_jk_scratch_ptr[dev] = ptr;
    return ISDF_OK;
}

}  // namespace

extern "C" size_t isdf_get_jk_static_max_bytes(void) { return (size_t)ISDF_JK_STATIC_SCRATCH_BYTES; }

extern "C" int isdf_get_jk_static(isdf_handle* h, const double* X, const double* W,
                                  const double* dm, int nset, int nao, int np,
                                  int with_j, int with_k, double* vj, double* vk)
{
    ARG_CHECK(h && nao > 0 && np > 0);
    const size_t need = isdf_get_jk_ws_bytes(nao, np);
    if (need > (size_t)ISDF_JK_STATIC_SCRATCH_BYTES) {
        isdf_set_error("isdf_get_jk_static: nao=%d np=%d needs %zu bytes of scratch, the built-in array holds %zu; "
                       "use isdf_get_jk with a caller workspace", nao, np, need, (size_t)ISDF_JK_STATIC_SCRATCH_BYTES);
        return ISDF_ERR_WS;
    }
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev != h->device || dev < 0 || dev >= JK_MAX_DEVICES) {
        isdf_set_error("isdf_get_jk_static: current device %d is not the handle's device %d", dev, h->device);
        return ISDF_ERR_ARG;
    }
    static const char key = 0;
    int rc;
    if ((rc = isdf_once_per_device(&key, resolve_jk_scratch, &dev))) return rc;
    return isdf_get_jk_block(h, X, W, np, 0, np, dm, nset, nao, np, with_j, with_k, vj, vk,
                             g_jk_scratch_ptr[dev], (size_t)ISDF_JK_STATIC_SCRATCH_BYTES);
}

// ---- exact-density J on the FFT grid (the FFTDF route, pyscf/pbc/df/fft_jk.py:33-109) ----------
// rho_g = sum_{mu nu} phi[mu,g] D[mu,nu] phi[nu,g] on a slab of grid points;
// the caller solves vR = ifft(coulG * fft(rho)).real (isdf_coulomb_apply on the whole grid) and
// J = weight * phi diag(vR) phi^T comes from isdf_j_from_potential.  Workspace: nao * ldphi doubles.
extern "C" size_t isdf_grid_j_ws_bytes(int nao, long long ldphi)
{
    return (size_t)nao * (size_t)ldphi * sizeof(double) + 256;
}

extern "C" int isdf_density_on_grid(isdf_handle* h, const double* phi, long long ldphi, int nao, long long ngl,
                                    const double* dm, int nset, double* rho, long long ldr,
                                    void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi && dm && rho && ws && nao > 0 && ngl > 0 && nset > 0 && ldphi >= ngl && ldr >= ngl);
    ARG_CHECK(ngl < 0x7fffffffLL);
    if (ws_bytes < isdf_grid_j_ws_bytes(nao, ldphi)) { isdf_set_error("isdf_density_on_grid: workspace too small"); return ISDF_ERR_WS; }
    double* T = (double*)align_up((size_t)ws, 256);
    cudaStream_t s = h->stream;
    for (int i = 0; i < nset; ++i) {
        int rc;      // T[mu,g] = sum_nu D[mu,nu] phi[nu,g]
        if ((rc = gemm6(h, nao, (int)ngl, nao, dm + (size_t)i * nao * nao, nao, 0, phi, ldphi, 1, T, ldphi, nullptr, 0, EPI_STORE))) return rc;
        coldot_kernel<<<(unsigned)((ngl + 255) / 256), 256, 0, s>>>(phi, T, nao, (int)ngl, ldphi, rho + (size_t)i * ldr);
        h->other_launches++;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

extern "C" int isdf_j_from_potential(isdf_handle* h, const double* phi, long long ldphi, int nao, long long ngl,
                                     const double* vR, long long ldv, int nset, double weight,
                                     double* vj, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi && vR && vj && ws && nao > 0 && ngl > 0 && nset > 0 && ldphi >= ngl && ldv >= ngl);
    ARG_CHECK(ngl < 0x7fffffffLL);
    if (ws_bytes < isdf_grid_j_ws_bytes(nao, ldphi)) { isdf_set_error("isdf_j_from_potential: workspace too small"); return ISDF_ERR_WS; }
    double* Y = (double*)align_up((size_t)ws, 256);
    cudaStream_t s = h->stream;
    for (int i = 0; i < nset; ++i) {
        dim3 g((unsigned)((ngl + 255) / 256), (unsigned)std::min(nao, 64));
        scale_cols_kernel<<<g, 256, 0, s>>>(phi, vR + (size_t)i * ldv, nao, (int)ngl, ldphi, ldphi, Y);
        h->other_launches++;
        GemmParams p;       // J = weight * Y phi^T
        p.M = nao; p.N = nao; p.K = (int)ngl; p.A = Y; p.lda = ldphi; p.a_kmajor = 0;
        p.B = phi; p.ldb = ldphi; p.b_kmajor = 0; p.C = vj + (size_t)i * nao * nao; p.ldc = nao;
        p.Cin = nullptr; p.ldcin = 0; p.E = nullptr; p.lde = 0; p.alpha = weight; p.neg = 0; p.epi = EPI_STORE; p.upper_only = 0;
        int rc;
        if ((rc = gemm_launch(h, p))) return rc;
    }
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}
