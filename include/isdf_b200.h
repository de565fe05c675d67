/* C-ABI of libisdf_b200.so -- the drop-in boundary of the ISDF hot path.
 *
 * The reference (stock PySCF 2.5.0 as mounted) reaches its native code for
 * this path through ctypes with host-owned buffers and raw pointers
 * (pyscf/pbc/gto/eval_gto.py:139-151 -> libpbc.PBCGTOval_sph_deriv0,
 * pyscf/lib/pbc/grid_ao.c:524-534).  This header keeps that contract:
 * plain pointers and sizes, no framework types, caller owns every buffer.
 *
 * Conventions
 *  - pointers prefixed h_ are HOST memory, all others are DEVICE memory of the
 *    GPU the handle was created on; every matrix is row-major FP64.
 *  - every function returns 0 on success, <0 on error; isdf_last_error()
 *    then holds a message (the reference's C entry points are void and trust
 *    their inputs; the Python wrapper turns a non-zero code into RuntimeError,
 *    the convention of pyscf/pbc/df/fft.py:216-220).
 *  - calls are asynchronous on the handle's stream unless they return host
 *    scalars (isdf_select_ip, isdf_stage_time_ms).
 *  - workspaces are caller-provided (LAPACK style): call the *_ws_bytes
 *    function first.
 */
#ifndef ISDF_B200_H
#define ISDF_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct isdf_handle isdf_handle;

enum isdf_stage {
    ISDF_STAGE_AO = 0,       /* stage 1 */
    ISDF_STAGE_SELECT = 1,   /* stage 2 */
    ISDF_STAGE_FIT = 2,      /* stage 3 */
    ISDF_STAGE_COULOMB = 3,  /* stage 4 (FFT + W) */
    ISDF_STAGE_JK = 4,       /* stage 5 */
    ISDF_STAGE_FFT = 5,      /* FFT part of stage 4 alone */
    ISDF_STAGE_GEMM = 6      /* last isdf_dgemm_timed call */
};

int isdf_create(int device, isdf_handle** out);
int isdf_destroy(isdf_handle* h);
int isdf_set_stream(isdf_handle* h, void* cuda_stream);
const char* isdf_last_error(void);
int isdf_num_sms(isdf_handle* h);
/* kernels launched so far: which=0 GEMM kernels, which=1 all other kernels */
long long isdf_launch_count(isdf_handle* h, int which);
/* CUDA-event time of the last run of a stage (ms), -1 if never run */
int isdf_stage_time_ms(isdf_handle* h, int stage, float* h_out);

/* Stage 1.  Replaces numint.eval_ao -> eval_gto -> PBCGTOval_sph_deriv0
 * (pyscf/pbc/dft/numint.py:70-93, pyscf/pbc/gto/eval_gto.py:31-165,
 * pyscf/lib/pbc/grid_ao.c:524-534) at the Gamma point.
 * atm/bas/env: libcint tables (pyscf/gto/mole.py:59-89); a: lattice rows (Bohr);
 * Ls: images sorted by |L|; rcut: per-shell radius.  Writes
 * phi[mu, g-g0] = sum_L chi_mu(r_g - R - L) for grid points g0 <= g < g1. */
int isdf_eval_ao(isdf_handle* h,
                 const int* h_atm, int natm, const int* h_bas, int nbas,
                 const double* h_env, int nenv,
                 const double* h_a, const int* h_mesh,
                 const double* h_Ls, int nimg, const double* h_rcut,
                 long long g0, long long g1,
                 double* phi, long long ldphi);

/* Stage 2 [spec; nearest reference code pyscf/lib/scipy_helper.py:71-110].
 * Greedy pivoted Cholesky of G[g,g'] = (sum_mu phi[mu,g] phi[mu,g'])^2.
 * Lt[npts, ldl] receives the factor rows (Lt[k, g] = L[g, k]); h_ip / h_resid
 * get the pivots and the residual diagonal at selection time; *h_rank <= npts
 * (stops early when the largest residual <= tol; tol < 0 -> ng*eps*max d). */
size_t isdf_select_ip_ws_bytes(int nao, long long ng, int npts);
int isdf_select_ip(isdf_handle* h, const double* phi, int nao, long long ng, long long ldphi,
                   int npts, double tol, double* Lt, long long ldl,
                   long long* h_ip, double* h_resid, int* h_rank,
                   void* ws, size_t ws_bytes);

/* Stage 3 [spec].  A = (X^T X)^2, B = (X^T phi)^2 with X = phi[:, ip];
 * zeta[np, ldz] = A^-1 B by Cholesky (A = U^T U) and two triangular solves. */
size_t isdf_fit_aux_ws_bytes(int nao, long long ng, int np);
int isdf_fit_aux(isdf_handle* h, const double* phi, int nao, long long ng, long long ldphi,
                 const long long* h_ip, int np, double* zeta, long long ldz,
                 void* ws, size_t ws_bytes);
/* Same result from the selection factor: zeta = Lp^-T Lt with Lp = Lt[:, ip]
 * (in place on Lt); mathematically identical to isdf_fit_aux. */
int isdf_fit_aux_from_factor(isdf_handle* h, double* Lt, long long ldl, long long ng,
                             const long long* h_ip, int np, void* ws, size_t ws_bytes);

/* Stage 4.  Conventions of pyscf/pbc/tools/pbc.py:149-211 (fft/ifft) and
 * :269-356 (coulG = 4 pi/|G|^2, G=0 -> 0), normalisation of
 * pyscf/pbc/df/fft_jk.py:72-74,93-94:
 *   V_P = ifft(coulG * fft(zeta_P)).real ;  W = (vol/ng) zeta V^T .
 * W is formed in reciprocal space, W = (vol/ng^2) Re[zhat^H diag(coulG) zhat]
 * (Parseval; no inverse FFT needed).  V (nullable) is produced on request. */
size_t isdf_build_W_ws_bytes(int np, const int* h_mesh);
int isdf_build_W(isdf_handle* h, const double* zeta, int np, long long ldz,
                 const int* h_mesh, const double* h_a,
                 double* W, double* V, long long ldv,
                 void* ws, size_t ws_bytes);
/* batched 3-D FFT Coulomb solve alone: out_P = ifft(coulG*fft(in_P)).real */
int isdf_coulomb_apply(isdf_handle* h, const double* in, long long ldin, int nvec,
                       const int* h_mesh, const double* h_a,
                       double* out, long long ldout, void* ws, size_t ws_bytes);

/* Stage 5 [spec; exact limit pyscf/pbc/df/fft_jk.py:33-109,177-302].
 * X[nao, np] = phi[:, ip];  J = X diag(W rho) X^T, rho_P = (X^T D X)_PP;
 * K = X ((X^T D X) o W) X^T.  dm/vj/vk: [nset, nao, nao]. */
size_t isdf_get_jk_ws_bytes(int nao, int np);
int isdf_get_jk(isdf_handle* h, const double* X, const double* W,
                const double* dm, int nset, int nao, int np,
                int with_j, int with_k, double* vj, double* vk,
                void* ws, size_t ws_bytes);

/* Generic FP64 GEMM used by every stage (exposed for tests / bench):
 * C[m,n] = epi( Cin[m,n] -/+ sum_k opA(k,m) opB(k,n) ), one FMA chain per
 * element in ascending k.  a_kmajor: A[k*lda+m] else A[m*lda+k]; same for B.
 * epi: 0 C=alpha*acc, 1 C=acc^2, 2 C=acc*E[m,n]. */
int isdf_dgemm(isdf_handle* h, int M, int N, int K,
               const double* A, long long lda, int a_kmajor,
               const double* B, long long ldb, int b_kmajor,
               double* C, long long ldc,
               const double* Cin, long long ldcin,
               const double* E, long long lde,
               double alpha, int neg, int epi, int upper_only);

/* Tuning knob: CTA tile of the GEMM kernel, 0 = automatic (default),
 * 2 = 64x128, 3 = 64x64. */
int isdf_set_gemm_tile(isdf_handle* h, int which);

/* Per-launch timing of the GEMM kernel (bench.py roofline): enable=1 resets and
 * starts bracketing every GEMM launch with CUDA events on the handle's stream,
 * enable=0 stops; read returns algorithmic flops (2*K per computed element),
 * summed kernel milliseconds and the launch count since the last enable. */
int isdf_gemm_profile(isdf_handle* h, int enable);
int isdf_gemm_profile_read(isdf_handle* h, double* h_flops, double* h_ms, long long* h_count);

/* gather columns: out[r, j] = in[r, idx[j]] (idx on host) */
int isdf_gather_cols(isdf_handle* h, const double* in, long long ldin, int nrows,
                     const long long* h_idx, int ncols, double* out, long long ldout);

#ifdef __cplusplus
}
#endif
#endif
