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

/* Stage 2, stepwise form used when the grid is sharded over several GPUs (one
 * process per GPU, each owning the slab [g_offset, g_offset+ngl) of grid points,
 * its columns of phi and of the factor Lt).  Per panel of <= isdf_select_panel_width()
 * pivots:  candidates (local: the `quota` largest residuals, packed with their phi
 * and factor columns into rows of isdf_select_row_len(nao,k0) doubles, quota+1 rows)
 * -> [all-gather of the blocks, done by the caller] -> panel (every rank factors the
 * gathered candidate set redundantly; pivots are exact while the largest candidate
 * residual exceeds the largest residual outside the set) -> flush (local: new factor
 * rows for the slab).  isdf_select_ip is this loop with one rank.  U (nullable,
 * [npts, ldu]) receives U[j,i] = L[ip_i, j], the transposed pivot factor that
 * isdf_fit_aux_from_U solves with. */
typedef struct isdf_select_plan isdf_select_plan;
int isdf_select_max_rows(void);
int isdf_select_panel_width(void);
/* Which form of the pivot kernel the last isdf_select_panel call launched: 0 = cooperative grid
 * (global-memory exchange, atomic barrier; the default), 1 = one 16-CTA cluster with DSMEM reads and
 * the cluster barrier, 2 = one 16-CTA cluster with st.async pushes and mbarriers.  1 and 2 are
 * requested with the environment variable ISDF_PANEL_CLUSTER and fall back to 0 when the candidate
 * set is too large for their shared-memory layout or the device cannot host the cluster.  All three
 * produce bit-identical pivots. */
int isdf_select_last_panel_mode(void);
int isdf_select_row_len(int nao, int k0);
int isdf_select_first_quota(int nranks);
int isdf_select_next_quota(int quota, int nb, int nbmax, int nranks);
int isdf_select_create(isdf_handle* h, int nao, long long ngl, long long g_offset,
                       long long ng_total, int npts, void* ws, size_t ws_bytes,
                       isdf_select_plan** out);
int isdf_select_destroy(isdf_select_plan* pl);
int isdf_select_begin(isdf_select_plan* pl, const double* phi, long long ldphi, double* h_dmax);
int isdf_select_set_tol(isdf_select_plan* pl, double tol);
int isdf_select_candidates(isdf_select_plan* pl, const double* phi, long long ldphi,
                           const double* Lt, long long ldl, int k0, int quota, double* block);
int isdf_select_panel(isdf_select_plan* pl, const double* blocks, int nranks, int quota,
                      int k0, int nbmax, double* U, long long ldu, int* h_nb, int* h_stop);
int isdf_select_flush(isdf_select_plan* pl, const double* phi, long long ldphi,
                      double* Lt, long long ldl, int k0, int nb);
int isdf_select_finish(isdf_select_plan* pl, int rank, long long* h_ip, double* h_resid);

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

/* Same solve on one slab of grid points, with the transposed pivot factor U
 * from isdf_select_panel: Z[np, ngl] <- U^-1 Z.  No communication between slabs. */
size_t isdf_fit_from_U_ws_bytes(long long ngl, int np);
int isdf_fit_aux_from_U(isdf_handle* h, double* U, long long ldu, int np,
                        double* Z, long long ldz, long long ngl, void* ws, size_t ws_bytes);

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

/* Stage 4 with the grid sharded by x-planes over ranks: local z/y passes, pack for the
 * all-to-all that re-shards the half spectra to (y,z)-column slabs (the exchange itself
 * is the caller's NCCL all_to_all), unpack, x pass fused with sqrt(weight*coulG).
 * The rank's partial W is then isdf_dgemm over its spectral slab (K = 2*nx*ncl,
 * alpha = vol/ng^2, upper_only) + isdf_symmetrize_upper, summed over ranks by the caller.
 * spec/send/recv/Y are complex128 buffers; lds/ldy count complex elements. */
int isdf_fft_zy(isdf_handle* h, const double* in, long long ldin, int nvec, int nxl,
                const int* h_mesh, void* spec, long long lds);
int isdf_fft_pack(isdf_handle* h, const void* spec, long long lds, int nvec, int nxl, int ncols,
                  const int* h_col_off, int nranks, void* send);
int isdf_fft_unpack(isdf_handle* h, const void* recv, int nvec, const int* h_x_off, int nranks,
                    int ncl, void* Y, long long ldy);
int isdf_fft_x_coulomb(isdf_handle* h, void* Y, long long ldy, int nvec,
                       const int* h_mesh, const double* h_a, int col0, int ncl);
int isdf_symmetrize_upper(isdf_handle* h, double* W, long long ld, int n);

/* Upper triangle of C[n,n] = alpha * A A^T (A [n,K] row-major FP64)
 * computed on the int8 tensor cores (tcgen05 kind::i8, exact int32 accumulation) from
 * at most `nslices` signed 7-bit digit planes per row-scaled entry (error-free splitting; 8
 * planes reproduce FP64-GEMM accuracy; K chunks that are small against the whole rows take
 * fewer planes for the same normwise accuracy 2^-(7 nslices - 6), see csrc/syrk_i8.cu).  Replaces the FP64 isdf_dgemm call of
 * the W build [spec; the exact-FFTDF analogue is the einsum of
 * pyscf/pbc/df/fft_jk.py:93-94]; follow with isdf_symmetrize_upper. */
size_t isdf_syrk_i8_ws_bytes(int n, long long K, int nslices);
/* int8 operations (2 per multiply-add, all tiles and digit products) issued by the last
 * isdf_syrk_i8 call of this process: the numerator of its roofline figure in bench.py. */
double isdf_syrk_i8_last_ops(void);
int isdf_syrk_i8(isdf_handle* h, int n, long long K, const double* A, long long lda,
                 double* C, long long ldc, double alpha, int nslices, void* ws, size_t ws_bytes);

/* Exact-density J on the FFT grid, the reference's FFTDF route
 * [pyscf/pbc/df/fft_jk.py:33-109: rhoR = einsum(ao, dm, ao); vR = ifft(coulG fft(rhoR)).real;
 * vj = weight * ao^T diag(vR) ao, weight = vol/ngrids].  Slab form for the grid-sharded driver:
 * rho on ngl local grid points, then isdf_coulomb_apply on the gathered density, then the J
 * contribution of the slab.  Workspace of both calls: isdf_grid_j_ws_bytes. */
size_t isdf_grid_j_ws_bytes(int nao, long long ldphi);
int isdf_density_on_grid(isdf_handle* h, const double* phi, long long ldphi, int nao, long long ngl,
                         const double* dm, int nset, double* rho, long long ldr,
                         void* ws, size_t ws_bytes);
int isdf_j_from_potential(isdf_handle* h, const double* phi, long long ldphi, int nao, long long ngl,
                          const double* vR, long long ldv, int nset, double weight,
                          double* vj, void* ws, size_t ws_bytes);

/* GGA exchange-correlation on the device-resident AO values ("next" row 1 of SURVEY 8f): the pieces
 * of NumInt.nr_rks [pyscf/pbc/dft/numint.py:283-381] for one slab of ngl grid points at the Gamma
 * point, closed shell.
 *   isdf_eval_ao_deriv1  AO values and Cartesian gradients [PBCGTOval_sph_deriv1,
 *                        pyscf/lib/pbc/grid_ao.c:548-558]: four planes (value, d/dx, d/dy, d/dz), each
 *                        [nao, ldphi], comp_stride doubles apart; other arguments as isdf_eval_ao.
 *   isdf_xc_density      rho4[4, ldr] = (rho, grad rho) for a symmetric dm[nao, nao]
 *                        [eval_rho, numint.py:138-145, hermi = 1].
 *   isdf_xc_potential    functional at every grid point [libxc through pyscf/dft/libxc.py:eval_xc in the
 *                        reference; published closed forms here, csrc/xc_core.cuh]; writes
 *                        wv4[4, ldr] = weight (vrho/2, 2 vsigma grad rho) [numint.py:366, :987-994] and
 *                        d_sums[2] (device) = (sum rho weight, sum rho weight exc) [numint.py:361-365].
 *                        code from isdf_xc_code("lda," | "lda,vwn" | "blyp" | "pbe" | "pbe0" | "b3lyp"), -1 if unknown;
 *                        the two hybrids give their semilocal part, the caller adds -fraction/2 K [pbc/dft/rks.py:103-118].
 *   isdf_xc_matrix       vmat[nao, nao] = M + M^T, M = Phi (sum_c wv_c Phi_c)^T [_vxc_mat, numint.py:812-815, :374].
 * A sharded grid sums vmat and d_sums over ranks.  Workspace of the three xc calls: isdf_xc_ws_bytes. */
int isdf_eval_ao_deriv1(isdf_handle* h,
                        const int* h_atm, int natm, const int* h_bas, int nbas,
                        const double* h_env, int nenv,
                        const double* h_a, const int* h_mesh,
                        const double* h_Ls, int nimg, const double* h_rcut,
                        long long g0, long long g1,
                        double* phi4, long long ldphi, long long comp_stride);
int isdf_xc_code(const char* name);
double isdf_xc_hybrid_fraction(int code);   /* exact-exchange fraction of "pbe0" (0.25) / "b3lyp" (0.20), else 0 */
size_t isdf_xc_ws_bytes(int nao, long long ldphi);
int isdf_xc_density(isdf_handle* h, const double* phi4, long long ldphi, long long comp_stride,
                    int nao, long long ngl, const double* dm, double* rho4, long long ldr,
                    void* ws, size_t ws_bytes);
int isdf_xc_potential(isdf_handle* h, int code, const double* rho4, long long ldr, long long ngl,
                      double weight, double* wv4, double* d_sums, void* ws, size_t ws_bytes);
int isdf_xc_matrix(isdf_handle* h, const double* phi4, long long ldphi, long long comp_stride,
                   int nao, long long ngl, const double* wv4, long long ldr, double* vmat,
                   void* ws, size_t ws_bytes);

/* Stage 5 [spec; exact limit pyscf/pbc/df/fft_jk.py:33-109,177-302].
 * X[nao, np] = phi[:, ip];  J = X diag(W rho) X^T, rho_P = (X^T D X)_PP;
 * K = X ((X^T D X) o W) X^T.  dm/vj/vk: [nset, nao, nao]. */
size_t isdf_get_jk_ws_bytes(int nao, int np);
int isdf_get_jk(isdf_handle* h, const double* X, const double* W,
                const double* dm, int nset, int nao, int np,
                int with_j, int with_k, double* vj, double* vk,
                void* ws, size_t ws_bytes);

/* Contribution of the interpolation points [p0, p0+npl) to J and K, given the row
 * block Wrows[npl, ldw] = W[p0:p0+npl, :]; the full J/K is the sum over disjoint
 * blocks (one block per rank, all-reduced by the caller).  p0 must be even. */
size_t isdf_get_jk_block_ws_bytes(int nao, int np, int npl);
int isdf_get_jk_block(isdf_handle* h, const double* X, const double* Wrows, long long ldw,
                      int p0, int npl, const double* dm, int nset, int nao, int np,
                      int with_j, int with_k, double* vj, double* vk,
                      void* ws, size_t ws_bytes);

/* isdf_get_jk_block for ONE density matrix given by its occupied orbitals, D = C diag(n) C^T: the mo_coeff /
 * mo_occ tags the reference's exchange build uses to shrink its work from nao to nocc
 * (pyscf/pbc/df/fft_jk.py:206-210, :235-238).  Cs[nocc, nao] (row-major) = sqrt(n_i) C[:, i]^T, nocc <= nao.
 * Same workspace size and outputs ([nao, nao]) as isdf_get_jk_block with nset = 1. */
int isdf_get_jk_block_occ(isdf_handle* h, const double* X, const double* Wrows, long long ldw,
                          int p0, int npl, const double* Cs, int nocc, int nao, int np,
                          int with_j, int with_k, double* vj, double* vk,
                          void* ws, size_t ws_bytes);

/* isdf_get_jk for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and result as isdf_get_jk without the workspace: the
 * scratch is a fixed device array inside the library (isdf_get_jk_static_max_bytes() = 32 MiB per
 * GPU that loads the library, held until exit), and the call makes no cudaMalloc / cudaFree /
 * allocating library call.  Sizes with isdf_get_jk_ws_bytes(nao, np) above that maximum are
 * rejected with ISDF_ERR_WS (-3) and a message naming both numbers.  The handle's device must be
 * current.  All calls on one device share the array: calls that overlap on the GPU race, so
 * serialise them (one stream, or events between streams).  It keeps the last call's contents;
 * nothing in it needs clearing, every part is rewritten before it is read. */
size_t isdf_get_jk_static_max_bytes(void);
int isdf_get_jk_static(isdf_handle* h, const double* X, const double* W,
                       const double* dm, int nset, int nao, int np,
                       int with_j, int with_k, double* vj, double* vk);

/* Non-local GTH projectors of FFTDF.get_pp (pyscf/pbc/df/fft.py:86-141), the part that scales with the number of atoms:
 * out[r, g] = tab[r, g] * exp(sign * i G_g . R) for one atom at R (host, 3 doubles) from its species' table
 * tab[nrow, ldt] (device; real plane and optional imaginary plane) over ng reciprocal vectors Gv[ng, 3] (device).
 * sign = -1: AO transforms times SI (fft.py:100, ft_ao phase); sign = +1: projectors times conj(SI) (fft.py:129).
 * The contraction over G that follows is isdf_dgemm on the re / im planes. */
int isdf_phase_expand(isdf_handle* h, const double* tab_re, const double* tab_im, long long ldt, int nrow,
                      const double* Gv, long long ng, const double* h_R, double sign,
                      double* out_re, double* out_im, long long ldo);

/* W = U^-1 M U^-T in place on the symmetric M[np, np]: both triangular solves of fit='implicit' on one GPU, the second
 * one restricted to the upper block-trapezoid (4/3 np^3 flops in all), result symmetrised.  Pattern of
 * pyscf/pbc/df/rsdf_builder.py:226,827 (Cholesky factor applied from both sides).  Workspace:
 * isdf_fit_from_U_ws_bytes(np, np). */
int isdf_congruence(isdf_handle* h, double* U, long long ldu, int np, double* M, long long ldm,
                    void* ws, size_t ws_bytes);

/* out[cols, rows] = in[rows, cols]^T (row strides ldout, ldin; no overlap): the re-orientation between the two
 * triangular solves of the congruence W = U^-1 M U^-T (no reference counterpart: NumPy's `.T`). */
int isdf_transpose(isdf_handle* h, const double* in, long long ldin, int rows, int cols,
                   double* out, long long ldout);

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
 * 2 = 64x128 (four 64x32 warps), 3 = 64x64 (two 64x32 warps), 4 = 64x64 (four 32x32 warps),
 * 5 = 64x128 / 6 = 128x128 (eight / sixteen 32x32 warps); 7..10 = the warp-specialised form fed by 2-D
 * tensor-map TMA (csrc/gemm_tma.cu: 64x128, 128x64, 128x128, 64x64 CTA tiles), bit-identical results, used
 * for operands with 16-byte aligned bases and row strides, the default kernel otherwise. */
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
