// Non-local GTH projector part of get_pp (pyscf/pbc/df/fft.py:86-141) in reciprocal space, device side.
//   vppnl = sum_atoms sum_l  Z_a^H h_l Z_a / vol^2 ,   Z_a[proj, nu] = sum_G conj(SI_a(G)) P_proj(G) ftao_nu(G)
// The radial x angular tables P(G) (projectors) and ft(G) (analytic AO transforms) depend on the SPECIES only
// and are prepared on the host per batch of G vectors (pyscf_isdf_b200/pseudo.py, closed forms); what scales
// with the number of atoms -- the structure-factor phases and the [projectors x G] . [G x nao] contraction --
// runs here: this kernel expands a species table to one atom's rows, the contraction is isdf_dgemm.
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <algorithm>

namespace {

// out[(r), g] = tab[r, g] * exp(sign * i G_g . R):  re/im planes, row stride ldo; tab is complex as two planes
__global__ void phase_expand_kernel(const double* __restrict__ tab_re, const double* __restrict__ tab_im, long long ldt,
                                    int nrow, const double* __restrict__ Gv, long long ng, double rx, double ry, double rz,
                                    double sign, double* __restrict__ out_re, double* __restrict__ out_im, long long ldo)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ng) return;
    const double ph = sign * (Gv[3 * g] * rx + Gv[3 * g + 1] * ry + Gv[3 * g + 2] * rz);
    double s, c;
    sincos(ph, &s, &c);
    for (int r = blockIdx.y; r < nrow; r += gridDim.y) {
        const double a = tab_re[(long long)r * ldt + g];
        const double b = tab_im ? tab_im[(long long)r * ldt + g] : 0.0;
        out_re[(long long)r * ldo + g] = a * c - b * s;
        out_im[(long long)r * ldo + g] = a * s + b * c;
    }
}

}  // namespace

extern "C" int isdf_phase_expand(isdf_handle* h, const double* tab_re, const double* tab_im, long long ldt, int nrow,
                                 const double* Gv, long long ng, const double* h_R, double sign,
                                 double* out_re, double* out_im, long long ldo)
{
    ARG_CHECK(h && tab_re && Gv && h_R && out_re && out_im && nrow > 0 && ng > 0 && ldt >= ng && ldo >= ng);
    dim3 grid((unsigned)((ng + 127) / 128), (unsigned)std::min(nrow, 64));
    phase_expand_kernel<<<grid, 128, 0, h->stream>>>(tab_re, tab_im, ldt, nrow, Gv, ng, h_R[0], h_R[1], h_R[2], sign,
                                                    out_re, out_im, ldo);
    CUDA_TRY(cudaGetLastError());
    h->other_launches++;
    return ISDF_OK;
}
