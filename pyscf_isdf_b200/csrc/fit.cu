// Stage 3: auxiliary-basis fitting  zeta = A^-1 B,
//   A = (X^T X)^2   [np, np]   (Hadamard square of the AO Gram at the IPs)
//   B = (X^T Phi)^2 [np, ng]   (pair densities IP x grid), X = Phi[:, ip].
// A = U^T U by a blocked right-looking Cholesky whose panel/trailing updates
// are the library's FP64 GEMM; the two triangular solves on the [np, ng]
// right-hand side are blocked the same way (diagonal blocks are inverted once,
// everything else is GEMM).  Same Cholesky-then-TRSM pattern the reference
// uses for its Gaussian DF fit (pyscf/pbc/df/rsdf_builder.py:226, :827).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <vector>
#include <algorithm>

namespace {

constexpr int CB = 128;   // diagonal block
static inline long long pad8(long long n) { return (n + 7) & ~7LL; }

// Factor one diagonal block (upper, A = U^T U) in shared memory and invert it.
// A: [cb, cb] block inside a row-major matrix (ld); Vinv: [CB, CB] block, ld CB.
__global__ void __launch_bounds__(256)
chol_diag_kernel(double* __restrict__ A, long long ld, int cb, double* __restrict__ Vinv, int* __restrict__ info,
                 int block_id, int do_factor, int batched_np)
{
    extern __shared__ double S[];      // [cb][CB+1]
    constexpr int LD = CB + 1;
    const int tid = threadIdx.x;
    if (batched_np > 0) {              // one CTA per diagonal block of a [batched_np]^2 matrix starting at A
        const int r0 = blockIdx.x * CB;
        A += (long long)r0 * ld + r0;
        Vinv += (size_t)blockIdx.x * CB * CB;
        cb = min(CB, batched_np - r0);
        block_id = blockIdx.x;
    }
    for (int i = tid; i < cb * cb; i += blockDim.x) {
        const int r = i / cb, c = i - r * cb;
        S[r * LD + c] = (c >= r) ? A[(long long)r * ld + c] : 0.0;
    }
    __syncthreads();
    if (do_factor) {
        for (int j = 0; j < cb; ++j) {
            if (tid == 0) {
                const double piv = S[j * LD + j];
                if (!(piv > 0.0)) { if (*info == 0) *info = block_id * CB + j + 1; }
                S[j * LD + j] = sqrt(piv > 0.0 ? piv : 1.0);
            }
            __syncthreads();
            const double inv = 1.0 / S[j * LD + j];
            for (int c = j + 1 + tid; c < cb; c += blockDim.x) S[j * LD + c] *= inv;
            __syncthreads();
            const int nrem = cb - j - 1;
            for (int i = tid; i < nrem * nrem; i += blockDim.x) {
                const int r = j + 1 + i / nrem, c = j + 1 + i % nrem;
                if (c >= r) S[r * LD + c] -= S[j * LD + r] * S[j * LD + c];
            }
            __syncthreads();
        }
        for (int i = tid; i < cb * cb; i += blockDim.x) {
            const int r = i / cb, c = i - r * cb;
            A[(long long)r * ld + c] = S[r * LD + c];      // upper factor, zeros below
        }
    }
    // inverse of the upper-triangular block, one column per thread
    for (int i = tid; i < CB * CB; i += blockDim.x) Vinv[i] = 0.0;
    __syncthreads();
    if (tid < cb) {
        const int c = tid;
        const double dcc = S[c * LD + c];
        if (!(fabs(dcc) > 0.0) || !isfinite(dcc)) atomicCAS(info, 0, block_id * CB + c + 1);   // zero / NaN pivot: reported by the caller
        Vinv[c * CB + c] = 1.0 / dcc;
        for (int i = c - 1; i >= 0; --i) {
            double s = 0.0;
            for (int j = i + 1; j <= c; ++j) s += S[i * LD + j] * Vinv[j * CB + c];
            Vinv[i * CB + c] = -s / S[i * LD + i];
        }
    }
}

int gemm_call(isdf_handle* h, int M, int N, int K,
              const double* A, long long lda, int ak, const double* B, long long ldb, int bk,
              double* C, long long ldc, const double* Cin, int neg, int epi, int upper)
{
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.A = A; p.lda = lda; p.a_kmajor = ak;
    p.B = B; p.ldb = ldb; p.b_kmajor = bk; p.C = C; p.ldc = ldc; p.Cin = Cin; p.ldcin = ldc;
    p.E = nullptr; p.lde = 0; p.alpha = 1.0; p.neg = neg; p.epi = epi; p.upper_only = upper;
    return gemm_launch(h, p);
}

int diag_kernel_cfg()
{
    const size_t smem = sizeof(double) * CB * (CB + 1);
    return isdf_func_max_smem((const void*)chol_diag_kernel, smem);
}

// In-place blocked Cholesky of the upper triangle of A[np, np]; Vinv gets the
// inverted diagonal blocks (block i at Vinv + i*CB*CB); P is a [CB, np] scratch panel.
int cholesky_upper(isdf_handle* h, double* A, long long lda_, int np, double* Vinv, double* P, int* d_info)
{
    const long long ldp = pad8(np);
    int rc;
    if ((rc = diag_kernel_cfg())) return rc;
    const size_t smem = sizeof(double) * CB * (CB + 1);
    cudaStream_t s = h->stream;
    for (int r0 = 0, ib = 0; r0 < np; r0 += CB, ++ib) {
        const int cb = std::min(CB, np - r0);
        double* Aii = A + (long long)r0 * lda_ + r0;
        chol_diag_kernel<<<1, 256, smem, s>>>(Aii, lda_, cb, Vinv + (size_t)ib * CB * CB, d_info, ib, 1, 0);
        CUDA_TRY(cudaGetLastError());
        h->other_launches++;
        const int r1 = r0 + cb, nrem = np - r1;
        if (nrem <= 0) break;
        // row panel: U[r0:r1, r1:] = U_ii^-T A[r0:r1, r1:] = Vinv^T * A_panel  (into P, then back)
        if ((rc = gemm_call(h, cb, nrem, cb, Vinv + (size_t)ib * CB * CB, CB, 1,
                            A + (long long)r0 * lda_ + r1, lda_, 1, P, ldp, nullptr, 0, EPI_STORE, 0))) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(A + (long long)r0 * lda_ + r1, sizeof(double) * lda_, P, sizeof(double) * ldp,
                                   sizeof(double) * nrem, cb, cudaMemcpyDeviceToDevice, s));
        // trailing update (upper triangle only): A[r1:, r1:] -= P^T P
        double* A22 = A + (long long)r1 * lda_ + r1;
        if ((rc = gemm_call(h, nrem, nrem, cb, P, ldp, 1, P, ldp, 1, A22, lda_, A22, 1, EPI_STORE, 1))) return rc;
    }
    return ISDF_OK;
}

// Diagonal-block inverses of an already-triangular U
int invert_diag_blocks(isdf_handle* h, double* U, long long ldu, int np, double* Vinv, int* d_info)
{
    int rc;
    if ((rc = diag_kernel_cfg())) return rc;
    const size_t smem = sizeof(double) * CB * (CB + 1);
    // the blocks are independent: one launch, one CTA per block
    chol_diag_kernel<<<(np + CB - 1) / CB, 256, smem, h->stream>>>(U, ldu, 0, Vinv, d_info, 0, 0, np);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}

// Z <- U^-T Z  (forward substitution, blocked left-looking); tmp: [CB, ng]
int solve_upper_trans(isdf_handle* h, const double* U, long long ldu, int np, const double* Vinv,
                      double* Z, long long ldz, long long ng, double* tmp)
{
    int rc;
    cudaStream_t s = h->stream;
    for (int r0 = 0, ib = 0; r0 < np; r0 += CB, ++ib) {
        const int cb = std::min(CB, np - r0);
        double* Zi = Z + (long long)r0 * ldz;
        if (r0 > 0)   // Z_i -= U[0:r0, r0:r0+cb]^T Z[0:r0]
            if ((rc = gemm_call(h, cb, (int)ng, r0, U + r0, ldu, 1, Z, ldz, 1, Zi, ldz, Zi, 1, EPI_STORE, 0))) return rc;
        // Z_i = Vinv^T Z_i
        if ((rc = gemm_call(h, cb, (int)ng, cb, Vinv + (size_t)ib * CB * CB, CB, 1, Zi, ldz, 1,
                            tmp, pad8(ng), nullptr, 0, EPI_STORE, 0))) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(Zi, sizeof(double) * ldz, tmp, sizeof(double) * pad8(ng),
                                   sizeof(double) * ng, cb, cudaMemcpyDeviceToDevice, s));
    }
    return ISDF_OK;
}

// Z <- U^-1 Z  (backward substitution, blocked)
int solve_upper(isdf_handle* h, const double* U, long long ldu, int np, const double* Vinv,
                double* Z, long long ldz, long long ng, double* tmp)
{
    int rc;
    cudaStream_t s = h->stream;
    const int nblk = (np + CB - 1) / CB;
    for (int ib = nblk - 1; ib >= 0; --ib) {
        const int r0 = ib * CB;
        const int cb = std::min(CB, np - r0);
        const int r1 = r0 + cb;
        double* Zi = Z + (long long)r0 * ldz;
        if (r1 < np)  // Z_i -= U[r0:r1, r1:] Z[r1:]
            if ((rc = gemm_call(h, cb, (int)ng, np - r1, U + (long long)r0 * ldu + r1, ldu, 0,
                                Z + (long long)r1 * ldz, ldz, 1, Zi, ldz, Zi, 1, EPI_STORE, 0))) return rc;
        // Z_i = Vinv Z_i
        if ((rc = gemm_call(h, cb, (int)ng, cb, Vinv + (size_t)ib * CB * CB, CB, 0, Zi, ldz, 1,
                            tmp, pad8(ng), nullptr, 0, EPI_STORE, 0))) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(Zi, sizeof(double) * ldz, tmp, sizeof(double) * pad8(ng),
                                   sizeof(double) * ng, cb, cudaMemcpyDeviceToDevice, s));
    }
    return ISDF_OK;
}

// Z <- U^-1 Z, right-looking: after a block row is solved, ALL rows above are updated with one
// GEMM (M = r0, K = 128).  Same flops as solve_upper, but the parallelism is r0 x ng instead of
// 128 x ng -- the better shape when Z is narrow (few right-hand sides: the column-sharded
// congruence W = U^-1 M U^-T), at the price of re-streaming Z once per block row.
int solve_upper_right(isdf_handle* h, const double* U, long long ldu, int np, const double* Vinv,
                      double* Z, long long ldz, long long ng, double* tmp)
{
    int rc;
    cudaStream_t s = h->stream;
    const int nblk = (np + CB - 1) / CB;
    for (int ib = nblk - 1; ib >= 0; --ib) {
        const int r0 = ib * CB;
        const int cb = std::min(CB, np - r0);
        double* Zi = Z + (long long)r0 * ldz;
        // Z_i = Vinv Z_i
        if ((rc = gemm_call(h, cb, (int)ng, cb, Vinv + (size_t)ib * CB * CB, CB, 0, Zi, ldz, 1,
                            tmp, pad8(ng), nullptr, 0, EPI_STORE, 0))) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(Zi, sizeof(double) * ldz, tmp, sizeof(double) * pad8(ng),
                                   sizeof(double) * ng, cb, cudaMemcpyDeviceToDevice, s));
        // Z[0:r0] -= U[0:r0, r0:r0+cb] Z_i
        if (r0 > 0)
            if ((rc = gemm_call(h, r0, (int)ng, cb, U + r0, ldu, 0, Zi, ldz, 1, Z, ldz, Z, 1, EPI_STORE, 0))) return rc;
    }
    return ISDF_OK;
}

struct FitWs { double* X; double* U; double* Vinv; double* tmp; int* info; };

size_t carve_fit(FitWs& w, void* ws, size_t bytes, int nao, long long ng, int np)
{
    WsAlloc a(ws, bytes);
    const size_t nblk = (np + CB - 1) / CB;
    w.X = a.get<double>((size_t)nao * pad8(np));
    w.U = a.get<double>((size_t)np * pad8(np));
    w.Vinv = a.get<double>(nblk * CB * CB);
    w.tmp = a.get<double>((size_t)CB * pad8(std::max<long long>(ng, np)));
    w.info = a.get<int>(4);
    return a.off + 256;
}

int check_info(isdf_handle* h, int* d_info, const char* what)
{
    int* hi = (int*)h->h_pinned;
    CUDA_TRY(cudaMemcpyAsync(hi, d_info, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (*hi != 0) {
        isdf_set_error("%s: matrix not positive definite at pivot %d", what, *hi);
        return ISDF_ERR_NUMERIC;
    }
    return ISDF_OK;
}

}  // namespace

// W = U^-1 M U^-T in place on the symmetric M[np, np] (row stride ldm), U upper triangular: the congruence of
// fit='implicit' on one GPU.  First Z = U^-1 M (blocked back substitution over all np columns, np^3 flops); then
// W = Z U^-T column block by column block from the right, LEFT-looking and for the rows at or above the block only
// (W is symmetric, the upper block-trapezoid determines it): per block one GEMM with M = rows above, N = 128,
// K = columns to the right -- np^3 / 3 flops instead of the np^3 of a second full solve, large-M shapes, and no
// transposed copies.  The lower triangle is left to isdf_symmetrize_upper (called here).
// Workspace: isdf_fit_from_U_ws_bytes(np, np).
extern "C" int isdf_congruence(isdf_handle* h, double* U, long long ldu, int np, double* M, long long ldm,
                               void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && U && M && ws && np > 0 && ldu >= np && ldm >= np);
    if (isdf_fit_from_U_ws_bytes(np, np) > ws_bytes) {
        isdf_set_error("isdf_congruence workspace too small: need %zu have %zu", isdf_fit_from_U_ws_bytes(np, np), ws_bytes);
        return ISDF_ERR_WS;
    }
    WsAlloc a(ws, ws_bytes);
    const int nblk = (np + CB - 1) / CB;
    double* Vinv = a.get<double>((size_t)nblk * CB * CB);
    double* tmp = a.get<double>((size_t)CB * pad8(np));
    int* info = a.get<int>(4);
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FIT], s));
    CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int) * 4, s));
    if ((rc = invert_diag_blocks(h, U, ldu, np, Vinv, info))) return rc;
    if ((rc = check_info(h, info, "isdf_congruence: zero or non-finite pivot in the selection factor"))) return rc;
    if ((rc = solve_upper_right(h, U, ldu, np, Vinv, M, ldm, np, tmp))) return rc;        // M <- Z = U^-1 M
    for (int jb = nblk - 1; jb >= 0; --jb) {
        const int r0 = jb * CB, cb = std::min(CB, np - r0), r1 = r0 + cb;
        double* Zj = M + r0;                                  // rows 0..r1 of block column jb
        if (r1 < np)    // Z[0:r1, jb] -= W[0:r1, r1:] U[jb, r1:]^T
            if ((rc = gemm_call(h, r1, cb, np - r1, M + r1, ldm, 0, U + (long long)r0 * ldu + r1, ldu, 0,
                                Zj, ldm, Zj, 1, EPI_STORE, 0))) return rc;
        // W[0:r1, jb] = Z[0:r1, jb] (U_jj^-1)^T   (tmp is [r1, cb])
        if ((rc = gemm_call(h, r1, cb, cb, Zj, ldm, 0, Vinv + (size_t)jb * CB * CB, CB, 0,
                            tmp, cb, nullptr, 0, EPI_STORE, 0))) return rc;
        CUDA_TRY(cudaMemcpy2DAsync(Zj, sizeof(double) * ldm, tmp, sizeof(double) * cb,
                                   sizeof(double) * cb, r1, cudaMemcpyDeviceToDevice, s));
    }
    if ((rc = isdf_symmetrize_upper(h, M, ldm, np))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FIT], s));
    h->ev_used[ISDF_STAGE_FIT] = true;
    return ISDF_OK;
}

namespace {

// out[c, r] = in[r, c] through a 32 x 33 shared-memory tile (both sides coalesced)
__global__ void transpose_kernel(const double* __restrict__ in, long long ldin, int rows, int cols,
                                 double* __restrict__ out, long long ldout)
{
    __shared__ double tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[i][threadIdx.x] = in[(long long)r * ldin + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (r < rows && c < cols) out[(long long)c * ldout + r] = tile[threadIdx.x][i];
    }
}

}  // namespace

// out[cols, rows] (row stride ldout) = in[rows, cols]^T (row stride ldin); used between the two triangular
// solves of the congruence W = U^-1 M U^-T and for the final W[cols, :] block (pyscf_isdf_b200/dist.py)
extern "C" int isdf_transpose(isdf_handle* h, const double* in, long long ldin, int rows, int cols,
                              double* out, long long ldout)
{
    ARG_CHECK(h && in && out && rows >= 0 && cols >= 0 && ldin >= cols && ldout >= rows && in != out);
    if (rows == 0 || cols == 0) return ISDF_OK;
    dim3 blk(32, 8);
    for (int r0 = 0; r0 < rows; r0 += 32 * 65535) {            // grid.y limit
        const int nr = std::min(rows - r0, 32 * 65535);
        dim3 grd((cols + 31) / 32, (nr + 31) / 32);
        transpose_kernel<<<grd, blk, 0, h->stream>>>(in + (long long)r0 * ldin, ldin, nr, cols, out + r0, ldout);
    }
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}

extern "C" size_t isdf_fit_aux_ws_bytes(int nao, long long ng, int np)
{
    FitWs w;
    return carve_fit(w, nullptr, 0, nao, ng, np);
}

extern "C" int isdf_fit_aux(isdf_handle* h, const double* phi, int nao, long long ng, long long ldphi,
                            const long long* h_ip, int np, double* zeta, long long ldz,
                            void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi && h_ip && zeta && ws && nao > 0 && ng > 0 && np > 0 && ldphi >= ng && ldz >= ng);
    FitWs w;
    const size_t need = carve_fit(w, ws, ws_bytes, nao, ng, np);
    if (need > ws_bytes) { isdf_set_error("fit_aux workspace too small: need %zu have %zu", need, ws_bytes); return ISDF_ERR_WS; }
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FIT], s));
    CUDA_TRY(cudaMemsetAsync(w.info, 0, sizeof(int) * 4, s));
    const long long ldu = pad8(np);
    if ((rc = isdf_gather_cols(h, phi, ldphi, nao, h_ip, np, w.X, ldu))) return rc;
    // A = (X^T X)^2 (upper triangle), B = (X^T Phi)^2
    if ((rc = gemm_call(h, np, np, nao, w.X, ldu, 1, w.X, ldu, 1, w.U, ldu, nullptr, 0, EPI_SQUARE, 1))) return rc;
    if ((rc = gemm_call(h, np, (int)ng, nao, w.X, ldu, 1, phi, ldphi, 1, zeta, ldz, nullptr, 0, EPI_SQUARE, 0))) return rc;
    if ((rc = cholesky_upper(h, w.U, ldu, np, w.Vinv, w.tmp, w.info))) return rc;
    if ((rc = check_info(h, w.info, "isdf_fit_aux Cholesky"))) return rc;
    if ((rc = solve_upper_trans(h, w.U, ldu, np, w.Vinv, zeta, ldz, ng, w.tmp))) return rc;
    if ((rc = solve_upper(h, w.U, ldu, np, w.Vinv, zeta, ldz, ng, w.tmp))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FIT], s));
    h->ev_used[ISDF_STAGE_FIT] = true;
    return ISDF_OK;
}

extern "C" int isdf_fit_aux_from_factor(isdf_handle* h, double* Lt, long long ldl, long long ng,
                                        const long long* h_ip, int np, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && Lt && h_ip && ws && ng > 0 && np > 0 && ldl >= ng);
    FitWs w;
    const size_t need = carve_fit(w, ws, ws_bytes, 1, ng, np);
    if (need > ws_bytes) { isdf_set_error("fit_aux workspace too small: need %zu have %zu", need, ws_bytes); return ISDF_ERR_WS; }
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FIT], s));
    CUDA_TRY(cudaMemsetAsync(w.info, 0, sizeof(int) * 4, s));
    // U[k, i] = Lt[k, ip_i]: transpose of the (lower-triangular) pivot rows of L
    const long long ldu = pad8(np);
    if ((rc = isdf_gather_cols(h, Lt, ldl, np, h_ip, np, w.U, ldu))) return rc;
    if ((rc = invert_diag_blocks(h, w.U, ldu, np, w.Vinv, w.info))) return rc;
    if ((rc = check_info(h, w.info, "isdf_fit_aux_from_factor: zero or non-finite pivot in the selection factor"))) return rc;
    if ((rc = solve_upper(h, w.U, ldu, np, w.Vinv, Lt, ldl, ng, w.tmp))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FIT], s));
    h->ev_used[ISDF_STAGE_FIT] = true;
    return ISDF_OK;
}

// zeta = U^-1 Z for the local slab Z[np, ngl] (in place), U[j,i] = L[ip_i, j] the transposed
// pivot factor assembled by isdf_select_panel.  Columns (grid points) are independent, so a
// grid-sharded run needs no communication here.
extern "C" size_t isdf_fit_from_U_ws_bytes(long long ngl, int np)
{
    FitWs w;
    WsAlloc a(nullptr, 0);
    const size_t nblk = (np + CB - 1) / CB;
    w.Vinv = a.get<double>(nblk * CB * CB);
    w.tmp = a.get<double>((size_t)CB * pad8(ngl));
    w.info = a.get<int>(4);
    return a.off + 256;
}

extern "C" int isdf_fit_aux_from_U(isdf_handle* h, double* U, long long ldu, int np,
                                   double* Z, long long ldz, long long ngl, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && U && Z && ws && np > 0 && ngl > 0 && ldu >= np && ldz >= ngl);
    if (isdf_fit_from_U_ws_bytes(ngl, np) > ws_bytes) {
        isdf_set_error("fit_aux_from_U workspace too small: need %zu have %zu", isdf_fit_from_U_ws_bytes(ngl, np), ws_bytes);
        return ISDF_ERR_WS;
    }
    WsAlloc a(ws, ws_bytes);
    const size_t nblk = (np + CB - 1) / CB;
    double* Vinv = a.get<double>(nblk * CB * CB);
    double* tmp = a.get<double>((size_t)CB * pad8(ngl));
    int* info = a.get<int>(4);
    int rc;
    cudaStream_t s = h->stream;
    CUDA_TRY(cudaEventRecord(h->ev0[ISDF_STAGE_FIT], s));
    CUDA_TRY(cudaMemsetAsync(info, 0, sizeof(int) * 4, s));
    if ((rc = invert_diag_blocks(h, U, ldu, np, Vinv, info))) return rc;
    if ((rc = check_info(h, info, "isdf_fit_aux_from_U: zero or non-finite pivot in the selection factor"))) return rc;
    if (ngl <= 4LL * np) {
        if ((rc = solve_upper_right(h, U, ldu, np, Vinv, Z, ldz, ngl, tmp))) return rc;
    } else {
        if ((rc = solve_upper(h, U, ldu, np, Vinv, Z, ldz, ngl, tmp))) return rc;
    }
    CUDA_TRY(cudaEventRecord(h->ev1[ISDF_STAGE_FIT], s));
    h->ev_used[ISDF_STAGE_FIT] = true;
    return ISDF_OK;
}
