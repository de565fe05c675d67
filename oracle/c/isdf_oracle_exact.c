/* CPU oracle in C -- TEST INFRASTRUCTURE ONLY.  PROTOTYPE of the arithmetic proposed in DESIGN.md
 * section 8 item 3 (not used by the product yet): the interpolation-point selection with every long
 * dot product specified as an EXACT fixed-point sum instead of one FMA chain, so that it can be
 * evaluated on integer tensor cores in any order (tiling, split-K, rank count) with identical bits.
 *
 *   AO columns:  X_phi[mu,g] = llrint(phi[mu,g] / sphi_g * 2^53),  sphi_g = 2^e, max_mu |phi[mu,g]| = m 2^e, m in [.5,1)
 *   pair value:  t(g,p) = R( sum_mu X_phi[mu,g] X_phi[mu,p] ) * (sphi_g sphi_p 2^-106) ;  G(g,p) = t^2
 *   diagonal:    d_g = G(g,g) ;  factor scale s_g = 2^e with t(g,g) = m 2^e   (|L[g,j]| <= sqrt(d_g) < s_g)
 *   factor rows: X[j,g] = llrint(L[j,g] / s_g * 2^53) kept next to the FP64 values
 *   column k, pivot p, block b = k / 128 (columns are grouped in fixed blocks of 128 by index):
 *       acc = G(g,p) - R( sum_{j < 128 b} X[j,g] X[j,p] ) * (s_g s_p 2^-106)     [long part: exact integers;
 *             chunks of 32768 columns are subtracted one after the other]
 *       acc = fma(-L[j,g], L[j,p], acc)  for j = 128 b .. k-1                    [current block: FMA chain]
 *       L[k,g] = acc / sqrt(d_p) ;  d_g = fma(-L[k,g], L[k,g], d_g) ;  d_p = -1
 *   R(S): the integer S (|S| < 2^124) as FP64 through its centred mixed-radix digits over the 16 moduli of
 *         pyscf_isdf_b200/csrc/crt_core.cuh and the Horner sum s = fma(s, p_k, digit_k), k = 15 .. 0 -- the digits
 *         of an integer are unique, so any implementation that forms them gets the same bits.
 * PARITY UNPINNED (the mounted reference has no ISDF code).
 */
#include <math.h>
#include <stdlib.h>
#include <float.h>
#include <string.h>

typedef __int128 i128;

static const int P[16] = {256, 255, 253, 251, 247, 241, 239, 233, 229, 227, 223, 217, 211, 199, 197, 193};
static int Mtab[16][16], ctab[16], tab_ready = 0;

static int inv_mod(int x, int m)
{
    x %= m;
    for (int y = 1; y < m; y++) if ((x * y) % m == 1) return y;
    return 0;
}

static void make_tables(void)
{
    for (int k = 0; k < 16; k++) {
        int prod = 1 % P[k];
        for (int j = 0; j < 16; j++) {
            Mtab[k][j] = j < k ? prod : 0;
            if (j < k) prod = prod * (P[j] % P[k]) % P[k];
        }
        ctab[k] = k == 0 ? 1 : inv_mod(prod, P[k]);
    }
    tab_ready = 1;
}

/* R(S) */
double oracle_recon_i128(i128 S)
{
    if (!tab_ready) make_tables();
    int d[16];
    for (int k = 0; k < 16; k++) {
        const int p = P[k];
        int r = (int)(S % p);
        int acc = 0;
        for (int j = 0; j < k; j++) acc += d[j] * Mtab[k][j];
        int t = ((r - acc) % p) * ctab[k] % p;
        if (t > (p - 1) / 2) t -= p;
        else if (t < -(p / 2)) t += p;
        d[k] = t;
    }
    double s = 0.0;
    for (int k = 15; k >= 0; k--) s = fma(s, (double)P[k], (double)d[k]);
    return s;
}

/* test hook: R of hi * 2^64 + lo */
double oracle_recon_parts(long long hi, unsigned long long lo)
{
    return oracle_recon_i128(((i128)hi << 64) + (i128)lo);
}

static double pow2_scale(double m)          /* 2^e with m = f 2^e, f in [.5,1); 1 for m = 0 */
{
    int e = 0;
    if (!(m > 0.0)) return 1.0;
    frexp(m, &e);
    return ldexp(1.0, e);
}

#define BLK 128
#define CHUNK 32768

long oracle_select_ip_exact(const double *phi, int nao, long ng, long ldphi, int npts,
                            double tol, double *Lt, long ldl, long *ip, double *resid)
{
    if (!tab_ready) make_tables();
    const double two53 = 9007199254740992.0, twom106 = ldexp(1.0, -106);
    long long *Xphi = (long long *)malloc(sizeof(long long) * (size_t)nao * ng);
    long long *X = (long long *)malloc(sizeof(long long) * (size_t)npts * ng);
    double *sphi = (double *)malloc(sizeof(double) * ng), *sg = (double *)malloc(sizeof(double) * ng);
    double *d = (double *)malloc(sizeof(double) * ng);
    long g;
#pragma omp parallel for
    for (g = 0; g < ng; g++) {
        double m = 0.0;
        for (int mu = 0; mu < nao; mu++) m = fmax(m, fabs(phi[(long)mu * ldphi + g]));
        sphi[g] = pow2_scale(m);
        i128 S = 0;
        for (int mu = 0; mu < nao; mu++) {
            const long long x = llrint(phi[(long)mu * ldphi + g] / sphi[g] * two53);
            Xphi[(size_t)mu * ng + g] = x;
            S += (i128)x * x;
        }
        const double t = oracle_recon_i128(S) * (sphi[g] * sphi[g] * twom106);
        d[g] = t * t;
        sg[g] = pow2_scale(t);
    }
    double dmax = 0.0;
    for (g = 0; g < ng; g++) if (d[g] > dmax) dmax = d[g];
    if (tol < 0) tol = (double)ng * DBL_EPSILON * dmax;
    long rank = 0;
    for (int k = 0; k < npts; k++) {
        long p = 0;
        double best = d[0];
        for (g = 1; g < ng; g++) if (d[g] > best) { best = d[g]; p = g; }   /* first maximum */
        if (!(best > tol)) break;
        const double piv = sqrt(best);
        const int k0 = (k / BLK) * BLK;                     /* columns before the current block */
        double *row = Lt + (long)k * ldl;
        long long *xrow = X + (size_t)k * ng;
#pragma omp parallel for
        for (g = 0; g < ng; g++) {
            i128 S = 0;
            for (int mu = 0; mu < nao; mu++) S += (i128)Xphi[(size_t)mu * ng + g] * Xphi[(size_t)mu * ng + p];
            const double t = oracle_recon_i128(S) * (sphi[g] * sphi[p] * twom106);
            double acc = t * t;
            for (int c0 = 0; c0 < k0; c0 += CHUNK) {
                const int c1 = c0 + CHUNK < k0 ? c0 + CHUNK : k0;
                i128 T = 0;
                for (int j = c0; j < c1; j++) T += (i128)X[(size_t)j * ng + g] * X[(size_t)j * ng + p];
                acc = acc - oracle_recon_i128(T) * (sg[g] * sg[p] * twom106);
            }
            for (int j = k0; j < k; j++) acc = fma(-Lt[(long)j * ldl + g], Lt[(long)j * ldl + p], acc);
            const double L = acc / piv;
            row[g] = L;
            xrow[g] = llrint(L / sg[g] * two53);
            d[g] = fma(-L, L, d[g]);
        }
        d[p] = -1.0;
        ip[k] = p;
        resid[k] = best;
        rank++;
    }
    free(Xphi); free(X); free(sphi); free(sg); free(d);
    return rank;
}
