/* CPU oracle in C -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 *
 * The interpolation-point selection of isdf_oracle.c (greedy pivoted Cholesky of
 * pyscf/lib/scipy_helper.py:84-105 on the implicit pair-density Gram matrix, every sum ONE
 * fused-multiply-add chain in ascending index, ties -> lowest grid index) evaluated in panels,
 * threaded with OpenMP, so that the BASELINE sizes (C8: 0.1 TFLOP, C64: 63 TFLOP) finish in
 * seconds / minutes instead of hours.  Same arithmetic, different schedule:
 *
 *   - every factor entry Lt[k,g] is still  ( (chain_mu phi[mu,g] phi[mu,p_k])^2
 *     - chain_{j<k ascending} Lt[j,g] Lt[j,p_k] ) / sqrt(d_{p_k}),   d_g = fma(-Lt[k,g], Lt[k,g], d_g);
 *   - a panel of <= PANEL pivots is chosen on a candidate set C (the MCAND largest residuals):
 *     residuals only decrease, so while max_C d > theta = max_{not C} d the candidate argmax IS the
 *     global argmax.  When that certificate fails the panel ends; a panel whose first pivot cannot
 *     be certified takes it from a scan of all points.  The candidates' factor columns are built
 *     pivot by pivot, the other points' entries of the panel afterwards, streaming Lt[:k0] once.
 *
 * It must agree BIT FOR BIT with oracle_select_ip (tests/test_oracle_blocked_cpu.py) -- that test
 * is what ties this schedule to the textbook loop.  PARITY UNPINNED against the reference: the
 * mounted tree has no ISDF code.
 *
 * Build: gcc -O3 -mavx2 -mfma -ffp-contract=off -fopenmp (Makefile); only explicit fma() calls fuse.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define PANEL 128
#define MCAND 2048
#define GB 96 /* grid points per flush block */
#define JB 192 /* finished columns per L2 block of the flush */

typedef struct { double d; long g; } cand_t;

static int cand_cmp(const void *a, const void *b)
{
    const cand_t *x = (const cand_t *)a, *y = (const cand_t *)b;
    if (x->d > y->d) return -1;
    if (x->d < y->d) return 1;
    return (x->g > y->g) - (x->g < y->g);
}

/* acc[kk][g] for a block of grid points: pair density squared, then the chain over the finished columns
 * j < k0, then the in-panel part; writes Lt[k0+kk, g] and updates d[g].
 * PivPhi[mu*nb + kk] = phi[mu, p_kk];  PivL[j*nb + kk] = Lt[j, p_kk] (j < k0 + nb, in-panel rows included) */
__attribute__((target_clones("avx512f", "default")))
static void flush_block(const double *phi, long ldphi, int nao, double *Lt, long ldl, long g0, int gn,
                        int k0, int nb, const double *PivPhi, const double *PivL, const double *piv,
                        const long *pg, double *d)
{
    double acc[PANEL][GB];
    for (int kk = 0; kk < nb; kk++)
        for (int g = 0; g < gn; g++) acc[kk][g] = 0.0;
    for (int mu = 0; mu < nao; mu++) {
        const double *ph = phi + (long)mu * ldphi + g0;
        const double *pp = PivPhi + (long)mu * nb;
        for (int kk = 0; kk < nb; kk++) {
            const double b = pp[kk];
#pragma omp simd
            for (int g = 0; g < gn; g++) acc[kk][g] = fma(ph[g], b, acc[kk][g]);
        }
    }
    for (int kk = 0; kk < nb; kk++)
#pragma omp simd
        for (int g = 0; g < gn; g++) acc[kk][g] = acc[kk][g] * acc[kk][g];
    if (gn == GB && (nb % 4) == 0) {
        /* register tiles of 4 pivots x 16 points, j in blocks that stay in L2; each accumulator still sees
         * its products in ascending j */
        for (int jb = 0; jb < k0; jb += JB) {
            const int je = (jb + JB < k0) ? jb + JB : k0;
            for (int k4 = 0; k4 < nb; k4 += 4)
                for (int g16 = 0; g16 < GB; g16 += 16) {
                    double a0[16], a1[16], a2[16], a3[16];
#pragma omp simd
                    for (int v = 0; v < 16; v++) { a0[v] = acc[k4][g16 + v]; a1[v] = acc[k4 + 1][g16 + v];
                                                   a2[v] = acc[k4 + 2][g16 + v]; a3[v] = acc[k4 + 3][g16 + v]; }
                    for (int j = jb; j < je; j++) {
                        const double *lj = Lt + (long)j * ldl + g0 + g16;
                        const double *pl = PivL + (long)j * nb + k4;
                        const double b0 = -pl[0], b1 = -pl[1], b2 = -pl[2], b3 = -pl[3];
#pragma omp simd
                        for (int v = 0; v < 16; v++) {
                            const double x = lj[v];
                            a0[v] = fma(x, b0, a0[v]); a1[v] = fma(x, b1, a1[v]);
                            a2[v] = fma(x, b2, a2[v]); a3[v] = fma(x, b3, a3[v]);
                        }
                    }
#pragma omp simd
                    for (int v = 0; v < 16; v++) { acc[k4][g16 + v] = a0[v]; acc[k4 + 1][g16 + v] = a1[v];
                                                   acc[k4 + 2][g16 + v] = a2[v]; acc[k4 + 3][g16 + v] = a3[v]; }
                }
        }
    } else {
        for (int j = 0; j < k0; j++) {
            const double *lj = Lt + (long)j * ldl + g0;
            const double *pl = PivL + (long)j * nb;
            for (int kk = 0; kk < nb; kk++) {
                const double b = -pl[kk];
#pragma omp simd
                for (int g = 0; g < gn; g++) acc[kk][g] = fma(lj[g], b, acc[kk][g]);
            }
        }
    }
    for (int kk = 0; kk < nb; kk++) {
        /* columns k0 .. k0+kk-1 of this panel, already final in acc[0..kk-1] (as L values) */
        for (int j = 0; j < kk; j++) {
            const double b = -PivL[(long)(k0 + j) * nb + kk];
#pragma omp simd
            for (int g = 0; g < gn; g++) acc[kk][g] = fma(acc[j][g], b, acc[kk][g]);
        }
        const double pv = piv[kk];
        double *row = Lt + (long)(k0 + kk) * ldl + g0;
#pragma omp simd
        for (int g = 0; g < gn; g++) {
            const double L = acc[kk][g] / pv;
            acc[kk][g] = L;
            row[g] = L;
            d[g0 + g] = fma(-L, L, d[g0 + g]);
        }
        if (pg[kk] >= g0 && pg[kk] < g0 + gn) d[pg[kk]] = -1.0;      /* as the loop does right after its step */
    }
}

long oracle_select_ip_blocked(const double *phi, int nao, long ng, long ldphi, int npts,
                              double tol, double *Lt, long ldl, long *ip, double *resid)
{
    double *d = (double *)malloc(sizeof(double) * ng);
    cand_t *all = (cand_t *)malloc(sizeof(cand_t) * ng);
    const int mmax = ng < MCAND ? (int)ng : MCAND;
    long *cg = (long *)malloc(sizeof(long) * mmax);
    double *dC = (double *)malloc(sizeof(double) * mmax);
    double *phiC = (double *)malloc(sizeof(double) * (size_t)nao * mmax);
    double *Lc = (double *)malloc(sizeof(double) * (size_t)(npts + 1) * mmax);     /* Lc[j*m + c] = Lt[j, cg[c]] */
    double *PivPhi = (double *)malloc(sizeof(double) * (size_t)nao * PANEL);
    double *PivL = (double *)malloc(sizeof(double) * (size_t)(npts + PANEL) * PANEL);
    double piv[PANEL];
    long pg[PANEL];
    long g;
    double dmax = 0.0;
#pragma omp parallel for
    for (g = 0; g < ng; g++) {
        double s = 0.0;
        for (int mu = 0; mu < nao; mu++) {
            const double v = phi[(long)mu * ldphi + g];
            s = fma(v, v, s);
        }
        d[g] = s * s;
    }
    for (g = 0; g < ng; g++) if (d[g] > dmax) dmax = d[g];
    if (tol < 0) tol = (double)ng * DBL_EPSILON * dmax;
    int k0 = 0, stop = 0;
    while (k0 < npts && !stop) {
        /* candidates: the m largest residuals (ties by grid index), theta = largest residual outside */
        for (g = 0; g < ng; g++) { all[g].d = d[g]; all[g].g = g; }
        qsort(all, (size_t)ng, sizeof(cand_t), cand_cmp);
        const int m = mmax;
        double theta = (ng > m) ? all[m].d : -HUGE_VAL;
        for (int c = 0; c < m; c++) { cg[c] = all[c].g; dC[c] = all[c].d; }
#pragma omp parallel for
        for (int c = 0; c < m; c++) {
            for (int mu = 0; mu < nao; mu++) phiC[(long)mu * m + c] = phi[(long)mu * ldphi + cg[c]];
            for (int j = 0; j < k0; j++) Lc[(long)j * m + c] = Lt[(long)j * ldl + cg[c]];
        }
        const int nbmax = (npts - k0 < PANEL) ? npts - k0 : PANEL;
        int nb = 0;
        while (nb < nbmax) {
            /* first maximum over the candidates: largest d, lowest grid index among equals */
            int cb = 0;
            for (int c = 1; c < m; c++)
                if (dC[c] > dC[cb] || (dC[c] == dC[cb] && cg[c] < cg[cb])) cb = c;
            double best = dC[cb];
            if (!(best > theta)) {
                if (nb > 0) break;                       /* certificate failed: flush what we have */
                /* cannot certify the first pivot of this panel (ties with theta): the definition itself */
                long p = 0; double bb = d[0];
                for (g = 1; g < ng; g++) if (d[g] > bb) { bb = d[g]; p = g; }
                int found = -1;
                for (int c = 0; c < m; c++) if (cg[c] == p) { found = c; break; }
                if (found < 0) {                         /* bring p into the candidate set in place of the smallest */
                    int cs = 0;
                    for (int c = 1; c < m; c++) if (dC[c] < dC[cs]) cs = c;
                    if (dC[cs] > theta) theta = dC[cs];  /* the point we drop is outside the set from now on */
                    cg[cs] = p; dC[cs] = d[p];
                    for (int mu = 0; mu < nao; mu++) phiC[(long)mu * m + cs] = phi[(long)mu * ldphi + p];
                    for (int j = 0; j < k0; j++) Lc[(long)j * m + cs] = Lt[(long)j * ldl + p];
                    found = cs;
                }
                cb = found; best = bb;
            }
            if (!(best > tol)) { stop = 1; break; }
            const double pv = sqrt(best);
            const int k = k0 + nb;
            double *newrow = Lc + (long)k * m;
            /* candidates in chunks of 64, the candidate index innermost (contiguous, SIMD): every candidate's
             * chain still runs over mu, then over j, in ascending order */
#pragma omp parallel for schedule(static)
            for (int c0 = 0; c0 < m; c0 += 64) {
                const int cn = (m - c0 < 64) ? m - c0 : 64;
                double acc[64];
                for (int c = 0; c < cn; c++) acc[c] = 0.0;
                for (int mu = 0; mu < nao; mu++) {
                    const double b = phiC[(long)mu * m + cb];
                    const double *pc = phiC + (long)mu * m + c0;
#pragma omp simd
                    for (int c = 0; c < cn; c++) acc[c] = fma(pc[c], b, acc[c]);
                }
#pragma omp simd
                for (int c = 0; c < cn; c++) acc[c] = acc[c] * acc[c];
                for (int j = 0; j < k; j++) {
                    const double b = Lc[(long)j * m + cb];
                    const double *lc = Lc + (long)j * m + c0;
#pragma omp simd
                    for (int c = 0; c < cn; c++) acc[c] = fma(-lc[c], b, acc[c]);
                }
                for (int c = 0; c < cn; c++) {
                    const double L = acc[c] / pv;
                    newrow[c0 + c] = L;
                    dC[c0 + c] = fma(-L, L, dC[c0 + c]);
                }
            }
            dC[cb] = -1.0;
            piv[nb] = pv; pg[nb] = cg[cb];
            ip[k] = cg[cb]; resid[k] = best;
            for (int mu = 0; mu < nao; mu++) PivPhi[(long)mu * nbmax + nb] = phiC[(long)mu * m + cb];
            for (int j = 0; j < k; j++) PivL[(long)j * nbmax + nb] = Lc[(long)j * m + cb];
            nb++;
        }
        if (nb == 0) break;
        /* flush: the panel's factor rows for every grid point (candidates included: same chains, same bits) */
        if (nb != nbmax) {                                /* repack to stride nb */
            for (int mu = 0; mu < nao; mu++) for (int kk = 0; kk < nb; kk++) PivPhi[(long)mu * nb + kk] = PivPhi[(long)mu * nbmax + kk];
            for (int j = 0; j < k0 + nb; j++) for (int kk = 0; kk < nb; kk++) PivL[(long)j * nb + kk] = PivL[(long)j * nbmax + kk];
        }
        long nblk = (ng + GB - 1) / GB, b;
#pragma omp parallel for schedule(dynamic, 4)
        for (b = 0; b < nblk; b++) {
            const long g0 = b * GB;
            const int gn = (int)((ng - g0 < GB) ? ng - g0 : GB);
            flush_block(phi, ldphi, nao, Lt, ldl, g0, gn, k0, nb, PivPhi, PivL, piv, pg, d);
        }
        k0 += nb;
    }
    free(d); free(all); free(cg); free(dC); free(phiC); free(Lc); free(PivPhi); free(PivL);
    return k0;
}
