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

// v = W rho, one warp per row
__global__ void gemv_rows_kernel(const double* __restrict__ W, long long ld, int n,
                                 const double* __restrict__ x, double* __restrict__ y)
{
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const double* w = W + (long long)row * ld;
    double s = 0.0;
    for (int j = lane; j < n; j += 32) s = fma(w[j], x[j], s);
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[row] = s;
}

// Y[mu,P] = X[mu,P] * v[P]
__global__ void scale_cols_kernel(const double* __restrict__ X, const double* __restrict__ v,
                                  int nao, int np, long long ld, double* __restrict__ Y)
{
    const int P = blockIdx.x * blockDim.x + threadIdx.x;
    if (P >= np) return;
    const double f = v[P];
    for (int mu = blockIdx.y; mu < nao; mu += gridDim.y) Y[(long long)mu * ld + P] = X[(long long)mu * ld + P] * f;
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

extern "C" size_t isdf_get_jk_ws_bytes(int nao, int np)
{
    return sizeof(double) * ((size_t)2 * nao * np + (size_t)np * np + 2 * (size_t)np) + 8192;
}

extern "C" int isdf_get_jk(isdf_handle* h, const double* X, const double* W,
                           const double* dm, int nset, int nao, int np,
                           int with_j, int with_k, double* vj, double* vk,
                           void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && X && W && dm && ws && nset > 0 && nao > 0 && np > 0);
    ARG_CHECK((!with_j || vj) && (!with_k || vk));
    if (isdf_get_jk_ws_bytes(nao, np) > ws_bytes) {
        isdf_set_error("get_jk workspace too small: need %zu have %zu", isdf_get_jk_ws_bytes(nao, np), ws_bytes);
        return ISDF_ERR_WS;
    }
    WsAlloc a(ws, ws_bytes);
    double* T = a.get<double>((size_t)nao * np);
    double* U = a.get<double>((size_t)nao * np);
    double* Nm = a.get<double>((size_t)np * np);
    double* rho = a.get<double>(np);
    double* v = a.get<double>(np);
    cudaStream_t s = h->stream;
    int rc;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_JK], s));
    for (int i = 0; i < nset; ++i) {
        const double* D = dm + (size_t)i * nao * nao;
        // T = D X
        if ((rc = gemm6(h, nao, np, nao, D, nao, 0, X, np, 1, T, np, nullptr, 0, EPI_STORE))) return rc;
        if (with_j) {
            coldot_kernel<<<(np + 255) / 256, 256, 0, s>>>(X, T, nao, np, np, rho);
            gemv_rows_kernel<<<(np + 7) / 8, 256, 0, s>>>(W, np, np, rho, v);
            dim3 g((np + 255) / 256, std::min(nao, 1024));
            scale_cols_kernel<<<g, 256, 0, s>>>(X, v, nao, np, np, U);
            CUDA_TRY(cudaGetLastError());
            h->other_launches += 3;
            if ((rc = gemm6(h, nao, nao, np, U, np, 0, X, np, 0, vj + (size_t)i * nao * nao, nao, nullptr, 0, EPI_STORE))) return rc;
        }
        if (with_k) {
            // N = (X^T T) o W ; U = X N ; K = U X^T
            if ((rc = gemm6(h, np, np, nao, X, np, 1, T, np, 1, Nm, np, W, np, EPI_MUL))) return rc;
            if ((rc = gemm6(h, nao, np, np, X, np, 0, Nm, np, 1, U, np, nullptr, 0, EPI_STORE))) return rc;
            if ((rc = gemm6(h, nao, nao, np, U, np, 0, X, np, 0, vk + (size_t)i * nao * nao, nao, nullptr, 0, EPI_STORE))) return rc;
        }
    }
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_JK], s));
    h->ev_used[ISDF_STAGE_JK] = true;
    return ISDF_OK;
}
