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
#include "eval_ao_deriv1_kernel.cuh"
#include <cmath>

using namespace isdf_ao_k;

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

    std::vector<DevAtom> atoms;
    std::vector<DevShell> shells;
    int nao_total = 0;
    if (!build_tables(h_atm, natm, h_bas, nbas, h_env, nenv, h_rcut, atoms, shells, &nao_total)) {
        isdf_set_error("isdf_eval_ao_deriv1: unsupported basis (l <= 3, nctr <= %d, shells of an atom contiguous)", MAX_NCTR);
        return ISDF_ERR_ARG;
    }
    ARG_CHECK(comp_stride >= (long long)nao_total * ldphi);

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
