// AO values and Cartesian gradients on the uniform FFT grid ("next" row 1 of SURVEY 8f: the GGA
// exchange-correlation path needs grad Phi).
//
// Replaces PBCGTOval_sph_deriv1 (pyscf/lib/pbc/grid_ao.c:548-558 -> PBCeval_sph_iter :301-429 with
// GTOshell_eval_grid_cart_deriv1, pyscf/lib/gto/deriv1.c:167-204) for the Gamma point.  Same CTA
// layout and the same image screening as eval_ao.cu (one CTA per reference block of 56 grid points, an
// image contributes to a shell for the whole block iff the block-minimum |r - R - L| < rcut[shell]);
// four outputs per AO -- value, d/dx, d/dy, d/dz -- written as four [nao, ld] planes comp_stride apart.
// One contraction at a time keeps the running sums (4 x 7 doubles) in registers; the radial sums are
// re-evaluated per contraction.  Angular functions and their gradients: ao_core.cuh (CPU-checked).
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include "ao_core.cuh"
#include <vector>
#include <cmath>
#include <algorithm>

namespace {

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

}  // namespace

// C-ABI entry: see include/isdf_b200.h (same arguments as isdf_eval_ao plus the plane stride)
extern "C" int isdf_eval_ao_deriv1(isdf_handle* h,
                            const int* h_atm, int natm, const int* h_bas, int nbas,
                            const double* h_env, int nenv,
                            const double* h_a, const int* h_mesh,
                            const double* h_Ls, int nimg, const double* h_rcut,
                            long long g0, long long g1,
                            double* phi, long long ldphi, long long comp_stride)
{
    ARG_CHECK(h && h_atm && h_bas && h_env && h_a && h_mesh && h_Ls && h_rcut && phi);
    ARG_CHECK(natm > 0 && nbas > 0 && nimg > 0 && nimg <= LIST_CAP);
    const long long ng = (long long)h_mesh[0] * h_mesh[1] * h_mesh[2];
    ARG_CHECK(0 <= g0 && g0 < g1 && g1 <= ng && ldphi >= g1 - g0);
    int nao_total = 0;
    for (int ib = 0; ib < nbas; ++ib) nao_total += h_bas[ib * 8 + 3] * (2 * h_bas[ib * 8 + 1] + 1);
    ARG_CHECK(comp_stride >= (long long)nao_total * ldphi);

    std::vector<DevAtom> atoms(natm);
    std::vector<DevShell> shells(nbas);
    for (int ia = 0; ia < natm; ++ia) {
        const int pc = h_atm[ia * 6 + 1];
        atoms[ia] = DevAtom{h_env[pc], h_env[pc + 1], h_env[pc + 2], 0.0, -1, -1};
    }
    int ao_off = 0;
    for (int ib = 0; ib < nbas; ++ib) {
        const int* b = h_bas + ib * 8;
        const int ia = b[0], l = b[1], np_ = b[2], nc = b[3];
        ARG_CHECK(ia >= 0 && ia < natm && l >= 0 && l <= 3 && nc >= 1 && nc <= MAX_NCTR);
        ARG_CHECK(nc * (2 * l + 1) <= MAX_SHELL_AO && b[5] + np_ <= nenv && b[6] + np_ * nc <= nenv);
        shells[ib] = DevShell{l, np_, nc, ao_off, b[5], b[6], h_rcut[ib]};
        ao_off += nc * (2 * l + 1);
        DevAtom& at = atoms[ia];
        if (at.sh0 < 0) { at.sh0 = ib; at.sh1 = ib + 1; }
        else { ARG_CHECK(at.sh1 == ib); at.sh1 = ib + 1; }   // shells of an atom must be contiguous
        at.rcmax = std::max(at.rcmax, h_rcut[ib]);
    }
    for (int ia = 0; ia < natm; ++ia) if (atoms[ia].sh0 < 0) atoms[ia].sh0 = atoms[ia].sh1 = 0;

    // small tables live in a scratch allocation freed at the end of the call
    const size_t bytes_atoms = align_up(sizeof(DevAtom) * natm, 256);
    const size_t bytes_sh = align_up(sizeof(DevShell) * nbas, 256);
    const size_t bytes_env = align_up(sizeof(double) * nenv, 256);
    const size_t bytes_ls = align_up(sizeof(double) * 3 * nimg, 256);
    char* d_tab = nullptr;
    CUDA_TRY(cudaMallocAsync((void**)&d_tab, bytes_atoms + bytes_sh + bytes_env + bytes_ls, h->stream));
    DevAtom* d_atoms = (DevAtom*)d_tab;
    DevShell* d_sh = (DevShell*)(d_tab + bytes_atoms);
    double* d_env = (double*)(d_tab + bytes_atoms + bytes_sh);
    double* d_ls = (double*)(d_tab + bytes_atoms + bytes_sh + bytes_env);
    CUDA_TRY(cudaMemcpyAsync(d_atoms, atoms.data(), sizeof(DevAtom) * natm, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpyAsync(d_sh, shells.data(), sizeof(DevShell) * nbas, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpyAsync(d_env, h_env, sizeof(double) * nenv, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaMemcpyAsync(d_ls, h_Ls, sizeof(double) * 3 * nimg, cudaMemcpyHostToDevice, h->stream));
    // the source vectors are pageable: make sure the copies are done before they go out of scope
    CUDA_TRY(cudaStreamSynchronize(h->stream));

    AoParams p;
    for (int i = 0; i < 3; ++i) p.mesh[i] = h_mesh[i];
    for (int i = 0; i < 9; ++i) p.a[i] = h_a[i];
    p.g0 = g0; p.g1 = g1; p.ng = ng;
    p.natm = natm; p.nimg = nimg;
    p.atoms = d_atoms; p.shells = d_sh; p.env = d_env; p.Ls = d_ls;
    p.phi = phi; p.ldphi = ldphi; p.cstride = comp_stride;
    const long long b0 = g0 / AO_BLK, b1 = (g1 - 1) / AO_BLK + 1;
    const long long nblk = b1 - b0;
    // split atoms over grid.y until the grid covers the machine a few times
    int ny = 1;
    while (nblk * ny < 8LL * h->num_sms && ny * 2 <= natm) ny *= 2;
    p.ny = ny;
    dim3 grid((unsigned)nblk, (unsigned)ny);
    eval_ao_deriv1_kernel<<<grid, AO_THREADS, 0, h->stream>>>(p);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    CUDA_TRY(cudaFreeAsync(d_tab, h->stream));
    return ISDF_OK;
}
