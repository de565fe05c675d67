/* CPU oracle in C -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 *
 * Bit-exact statement of the interpolation-point selection: the unblocked
 * greedy pivoted Cholesky of pyscf/lib/scipy_helper.py:84-105 applied to the
 * implicit pair-density Gram matrix G[g,g'] = (sum_mu phi[mu,g] phi[mu,g'])^2,
 * with every sum written as ONE fused-multiply-add chain in ascending index:
 *
 *   s_g  = fma-chain_mu phi[mu,g]^2 from 0 ;             d_g = s_g * s_g
 *   step k: p = first argmax d ;  stop if d_p <= tol ;   piv = sqrt(d_p)
 *           t   = fma-chain_mu phi[mu,g] phi[mu,p] from 0 ; acc = t * t
 *           acc = fma(-Lt[j,g], Lt[j,p], acc)  for j = 0..k-1
 *           Lt[k,g] = acc / piv ;  d_g = fma(-Lt[k,g], Lt[k,g], d_g) ; d_p = -1
 *
 * The CUDA path (pyscf_isdf_b200/csrc/select.cu) runs the same chains in a
 * blocked order, so pivots AND factor must agree bit for bit.
 * PARITY UNPINNED: the mounted reference has no ISDF code to check this against.
 *
 * Build: gcc -O2 -mfma -ffp-contract=off -shared -fPIC (see Makefile).
 */
#include <math.h>
#include <stdlib.h>
#include <float.h>

long oracle_select_ip(const double *phi, int nao, long ng, long ldphi, int npts,
                      double tol, double *Lt, long ldl, long *ip, double *resid)
{
    double *d = (double *)malloc(sizeof(double) * ng);
    double dmax = 0.0;
    long g;
#pragma omp parallel for
    for (g = 0; g < ng; g++) {
        double s = 0.0;
        for (int mu = 0; mu < nao; mu++) {
            double v = phi[(long)mu * ldphi + g];
            s = fma(v, v, s);
        }
        d[g] = s * s;
    }
    for (g = 0; g < ng; g++) if (d[g] > dmax) dmax = d[g];
    if (tol < 0) tol = (double)ng * DBL_EPSILON * dmax;
    long rank = 0;
    for (int k = 0; k < npts; k++) {
        long p = 0;
        double best = d[0];
        for (g = 1; g < ng; g++) if (d[g] > best) { best = d[g]; p = g; }   /* first maximum */
        if (!(best > tol)) break;
        double piv = sqrt(best);
        double *row = Lt + (long)k * ldl;
#pragma omp parallel for
        for (g = 0; g < ng; g++) {
            double t = 0.0;
            for (int mu = 0; mu < nao; mu++)
                t = fma(phi[(long)mu * ldphi + g], phi[(long)mu * ldphi + p], t);
            double acc = t * t;
            for (int j = 0; j < k; j++)
                acc = fma(-Lt[(long)j * ldl + g], Lt[(long)j * ldl + p], acc);
            double L = acc / piv;
            row[g] = L;
            d[g] = fma(-L, L, d[g]);
        }
        d[p] = -1.0;
        ip[k] = p;
        resid[k] = best;
        rank++;
    }
    free(d);
    return rank;
}

/* C[m,n] = epi( Cin[m,n] -/+ chain_k A(k,m) B(k,n) ): the GEMM contract of
 * pyscf_isdf_b200/csrc/gemm.cu, for bit-exact kernel tests.
 * a_kmajor: A[k*lda+m] else A[m*lda+k]; epi 0: alpha*acc, 1: acc^2, 2: acc*E */
void oracle_gemm_chain(int M, int N, int K, const double *A, long lda, int a_kmajor,
                       const double *B, long ldb, int b_kmajor, double *C, long ldc,
                       const double *Cin, long ldcin, const double *E, long lde,
                       double alpha, int neg, int epi)
{
    long m;
#pragma omp parallel for
    for (m = 0; m < M; m++) {
        for (long n = 0; n < N; n++) {
            double acc = Cin ? Cin[m * ldcin + n] : 0.0;
            for (long k = 0; k < K; k++) {
                double a = a_kmajor ? A[k * lda + m] : A[m * lda + k];
                double b = b_kmajor ? B[k * ldb + n] : B[n * ldb + k];
                acc = fma(neg ? -a : a, b, acc);
            }
            if (epi == 1) acc = acc * acc;
            else if (epi == 2) acc = acc * E[m * lde + n];
            else acc = alpha * acc;
            C[m * ldc + n] = acc;
        }
    }
}
