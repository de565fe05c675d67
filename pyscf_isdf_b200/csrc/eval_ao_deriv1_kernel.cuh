// Kernel of isdf_eval_ao_deriv1 (see eval_ao_deriv1.cu for what it replaces) together with the host
// routine that turns the libcint-style atm/bas/env arrays into its tables.  Kept in a header so that
// tests/emul/ao_kernel_emul.cpp can compile the very same source with g++ and run it block by block on
// the CPU (64 threads, barriers and warp exchanges emulated) against the oracle.
#pragma once
#include "ao_core.cuh"
#include <vector>
#include <algorithm>

namespace isdf_ao_k {

constexpr int AO_BLK = 56;
constexpr int AO_THREADS = 64;
constexpr int LIST_CAP = 2048;
constexpr int MAX_NCTR = 4;
constexpr int MAX_SHELL_AO = 28;   // nctr * (2l+1)

struct DevShell { int l, nprim, nctr, ao_off, exp_off, coef_off; double rcut; };
struct DevAtom { double x, y, z, rcmax; int sh0, sh1; };

struct AoParams {
    int mesh[3];
    double a[9];
    long long g0, g1, ng;
    int natm, nimg, ny;
    const DevAtom* atoms;
    const DevShell* shells;
    const double* env;
    const double* Ls;
    double* phi;
    long long ldphi, cstride;
};

__device__ __forceinline__ double fftfreq_frac(int i, int n)
{
    return (double)(i < (n + 1) / 2 ? i : i - n) / (double)n;
}

__global__ void __launch_bounds__(AO_THREADS) eval_ao_deriv1_kernel(AoParams p)
{
    __shared__ double s_pt[3][AO_THREADS];
    __shared__ int s_img[LIST_CAP];
    __shared__ double s_mind[LIST_CAP];
    __shared__ int s_cnt;
    __shared__ int s_wcnt[2];
    __shared__ double s_red[4][2];

    const int tid = threadIdx.x;
    const long long blk = p.g0 / AO_BLK + blockIdx.x;
    const long long g = blk * AO_BLK + tid;
    const bool valid = (tid < AO_BLK) && (g < p.ng);

    double px = 0, py = 0, pz = 0;
    if (valid) {
        const int nyz = p.mesh[1] * p.mesh[2];
        const int ix = (int)(g / nyz);
        const int rem = (int)(g - (long long)ix * nyz);
        const int iy = rem / p.mesh[2];
        const int iz = rem - iy * p.mesh[2];
        const double fx = fftfreq_frac(ix, p.mesh[0]);
        const double fy = fftfreq_frac(iy, p.mesh[1]);
        const double fz = fftfreq_frac(iz, p.mesh[2]);
        px = fx * p.a[0] + fy * p.a[3] + fz * p.a[6];
        py = fx * p.a[1] + fy * p.a[4] + fz * p.a[7];
        pz = fx * p.a[2] + fy * p.a[5] + fz * p.a[8];
    }
    s_pt[0][tid] = px; s_pt[1][tid] = py; s_pt[2][tid] = pz;
    const int npt = (int)min((long long)AO_BLK, p.ng - blk * AO_BLK);

    // bounding sphere of the block: centre = bbox midpoint
    double lo[3], hi[3];
    {
        double v[3] = {px, py, pz};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            double mn = valid ? v[d] : 1e300, mx = valid ? v[d] : -1e300;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if ((tid & 31) == 0) { s_red[d][tid >> 5] = mn; }
            __syncthreads();
            lo[d] = fmin(s_red[d][0], s_red[d][1]);
            __syncthreads();
            if ((tid & 31) == 0) { s_red[d][tid >> 5] = mx; }
            __syncthreads();
            hi[d] = fmax(s_red[d][0], s_red[d][1]);
            __syncthreads();
        }
    }
    const double cx = 0.5 * (lo[0] + hi[0]), cy = 0.5 * (lo[1] + hi[1]), cz = 0.5 * (lo[2] + hi[2]);
    const double rad = 0.5 * sqrt((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1]) +
                                  (hi[2] - lo[2]) * (hi[2] - lo[2])) * (1.0 + 1e-12) + 1e-12;

    const int ia0 = (int)((long long)p.natm * blockIdx.y / p.ny);
    const int ia1 = (int)((long long)p.natm * (blockIdx.y + 1) / p.ny);

    for (int ia = ia0; ia < ia1; ++ia) {
        const DevAtom at = p.atoms[ia];
        if (tid == 0) s_cnt = 0;
        __syncthreads();
        // ---- screening: one image per thread, exact block-minimum distance ----
        for (int m0 = 0; m0 < p.nimg; m0 += AO_THREADS) {
            const int m = m0 + tid;
            bool keep = false;
            double mind = 0.0;
            if (m < p.nimg) {
                const double rx = at.x + p.Ls[3 * m], ry = at.y + p.Ls[3 * m + 1], rz = at.z + p.Ls[3 * m + 2];
                const double dc = sqrt((cx - rx) * (cx - rx) + (cy - ry) * (cy - ry) + (cz - rz) * (cz - rz));
                if (dc - rad < at.rcmax) {
                    double d2min = 1e300;
                    for (int i = 0; i < npt; ++i) {
                        const double dx = s_pt[0][i] - rx, dy = s_pt[1][i] - ry, dz = s_pt[2][i] - rz;
                        d2min = fmin(d2min, dx * dx + dy * dy + dz * dz);
                    }
                    mind = sqrt(d2min);
                    keep = mind < at.rcmax;
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, keep);
            if ((tid & 31) == 0) s_wcnt[tid >> 5] = __popc(bal);
            __syncthreads();
            const int base = s_cnt + ((tid >> 5) ? s_wcnt[0] : 0);
            if (keep) {
                const int pos = base + __popc(bal & ((1u << (tid & 31)) - 1u));
                if (pos < LIST_CAP) { s_img[pos] = m; s_mind[pos] = mind; }
            }
            __syncthreads();
            if (tid == 0) s_cnt = min(LIST_CAP, s_cnt + s_wcnt[0] + s_wcnt[1]);
            __syncthreads();
        }
        const int cnt = s_cnt;
        if (cnt == 0) {
            // shells of this atom are identically zero on this block
            if (valid && g >= p.g0 && g < p.g1) {
                for (int is = at.sh0; is < at.sh1; ++is) {
                    const DevShell sh = p.shells[is];
                    const int dim = sh.nctr * (2 * sh.l + 1);
                    for (int c = 0; c < dim; ++c)
                        for (int k = 0; k < 4; ++k) p.phi[k * p.cstride + (long long)(sh.ao_off + c) * p.ldphi + (g - p.g0)] = 0.0;
                }
            }
            __syncthreads();
            continue;
        }
        // ---- evaluation: shell, contraction, then images (reference summation order per AO) ----
        for (int is = at.sh0; is < at.sh1; ++is) {
            const DevShell sh = p.shells[is];
            const int deg = 2 * sh.l + 1;
            const double fac = sh.l == 0 ? 0.282094791773878143 : (sh.l == 1 ? 0.488602511902919921 : 1.0);
            const double* ex = p.env + sh.exp_off;
            for (int c = 0; c < sh.nctr; ++c) {
                const double* cf = p.env + sh.coef_off + c * sh.nprim;
                double acc[4][7];
                isdf_ao::shell_point_deriv1(sh.l, sh.nprim, ex, cf, fac, sh.rcut, s_img, s_mind, cnt, p.Ls,
                                            at.x, at.y, at.z, px, py, pz, acc);
                if (valid && g >= p.g0 && g < p.g1) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
#pragma unroll
                        for (int d = 0; d < 7; ++d)
                            if (d < deg)
                                p.phi[k * p.cstride + (long long)(sh.ao_off + c * deg + d) * p.ldphi + (g - p.g0)] = acc[k][d];
                }
            }
        }
        __syncthreads();
    }
}


// atm/bas/env -> kernel tables; returns false on an argument the kernel does not support
inline bool build_tables(const int* h_atm, int natm, const int* h_bas, int nbas, const double* h_env, int nenv,
                         const double* h_rcut, std::vector<DevAtom>& atoms, std::vector<DevShell>& shells, int* nao_out)
{
    atoms.assign(natm, DevAtom{});
    shells.assign(nbas, DevShell{});
    for (int ia = 0; ia < natm; ++ia) {
        const int pc = h_atm[ia * 6 + 1];
        atoms[ia] = DevAtom{h_env[pc], h_env[pc + 1], h_env[pc + 2], 0.0, -1, -1};
    }
    int ao_off = 0;
    for (int ib = 0; ib < nbas; ++ib) {
        const int* b = h_bas + ib * 8;
        const int ia = b[0], l = b[1], np_ = b[2], nc = b[3];
        if (!(ia >= 0 && ia < natm && l >= 0 && l <= 3 && nc >= 1 && nc <= MAX_NCTR)) return false;
        if (!(nc * (2 * l + 1) <= MAX_SHELL_AO && b[5] + np_ <= nenv && b[6] + np_ * nc <= nenv)) return false;
        shells[ib] = DevShell{l, np_, nc, ao_off, b[5], b[6], h_rcut[ib]};
        ao_off += nc * (2 * l + 1);
        DevAtom& at = atoms[ia];
        if (at.sh0 < 0) { at.sh0 = ib; at.sh1 = ib + 1; }
        else { if (at.sh1 != ib) return false; at.sh1 = ib + 1; }   // shells of an atom must be contiguous
        at.rcmax = std::max(at.rcmax, h_rcut[ib]);
    }
    for (int ia = 0; ia < natm; ++ia) if (atoms[ia].sh0 < 0) atoms[ia].sh0 = atoms[ia].sh1 = 0;
    *nao_out = ao_off;
    return true;
}

}  // namespace isdf_ao_k
