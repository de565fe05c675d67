// Exchange-correlation potential matrix of a closed-shell density on the device-resident AO values and
// their gradients ("next" row 1 of SURVEY 8f): NumInt.nr_rks of the reference for one grid slab,
// pyscf/pbc/dft/numint.py:283-381 --
//   T = D Phi;  rho = sum_mu Phi T,  grad rho = 2 sum_mu (grad Phi) T                  (:138-145, hermi = 1)
//   exc, vrho, vsigma = functional(rho, |grad rho|^2)                                  (libxc in the reference)
//   wv = weight (vrho / 2, 2 vsigma grad rho);  aow = sum_c wv_c Phi_c                 (:366, :987-994, :812-815)
//   vmat = Phi aow^T + (Phi aow^T)^T                                                   (:374)
// The two contractions over the AO index and over the grid are the library's FP64 GEMM; everything else
// is streaming work over the grid slab (one thread per grid point, coalesced along the grid index).
// The functional is evaluated with the dual-number routines of xc_core.cuh.
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include "xc_core.cuh"
#include <algorithm>
#include <cctype>
#include <string>

namespace {

constexpr int XC_THREADS = 256;

// rho4[c][g] = f_c sum_mu phi_c[mu,g] T[mu,g],  f = 1, 2, 2, 2
__global__ void __launch_bounds__(XC_THREADS)
xc_rho_kernel(const double* __restrict__ phi4, long long ldphi, long long cstride,
              const double* __restrict__ T, int nao, int ngl, double* __restrict__ rho4, long long ldr)
{
    const int g = blockIdx.x * XC_THREADS + threadIdx.x;
    if (g >= ngl) return;
    isdf_xc::rho_item(g, phi4, ldphi, cstride, T, nao, rho4, ldr);
}

// functional at every grid point; wv4 = weight (vrho/2, 2 vsigma grad rho); per-block partial sums of
// rho weight and rho weight exc, reduced in a fixed tree order (the result does not depend on timing)
__global__ void __launch_bounds__(XC_THREADS)
xc_point_kernel(int code, const double* __restrict__ rho4, long long ldr, int ngl, double weight,
                double* __restrict__ wv4, double* __restrict__ partial)
{
    __shared__ double s_n[XC_THREADS], s_e[XC_THREADS];
    const int g = blockIdx.x * XC_THREADS + threadIdx.x;
    double den = 0.0, ex = 0.0;
    if (g < ngl) isdf_xc::point_item(code, g, rho4, ldr, weight, wv4, &den, &ex);
    s_n[threadIdx.x] = den; s_e[threadIdx.x] = ex;
    __syncthreads();
    for (int o = XC_THREADS / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_n[threadIdx.x] += s_n[threadIdx.x + o]; s_e[threadIdx.x] += s_e[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[2 * blockIdx.x] = s_n[0]; partial[2 * blockIdx.x + 1] = s_e[0]; }
}

// sums[0] = sum of partial[2b], sums[1] = sum of partial[2b+1]: one block, fixed order
__global__ void __launch_bounds__(XC_THREADS)
xc_sum_kernel(const double* __restrict__ partial, int nblk, double* __restrict__ sums)
{
    __shared__ double s_n[XC_THREADS], s_e[XC_THREADS];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblk; i += XC_THREADS) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    s_n[threadIdx.x] = a; s_e[threadIdx.x] = b;
    __syncthreads();
    for (int o = XC_THREADS / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_n[threadIdx.x] += s_n[threadIdx.x + o]; s_e[threadIdx.x] += s_e[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { sums[0] = s_n[0]; sums[1] = s_e[0]; }
}

// aow[mu,g] = sum_c wv4[c][g] phi_c[mu,g]
__global__ void __launch_bounds__(XC_THREADS)
xc_aow_kernel(const double* __restrict__ phi4, long long ldphi, long long cstride,
              const double* __restrict__ wv4, long long ldr, int nao, int ngl, double* __restrict__ aow)
{
    const int g = blockIdx.x * XC_THREADS + threadIdx.x;
    if (g >= ngl) return;
    isdf_xc::aow_item(g, blockIdx.y, gridDim.y, phi4, ldphi, cstride, wv4, ldr, nao, aow);
}

// vmat = mat + mat^T
__global__ void xc_symm_kernel(const double* __restrict__ mat, int n, double* __restrict__ vmat)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j < n) vmat[(long long)i * n + j] = mat[(long long)i * n + j] + mat[(long long)j * n + i];
}

size_t xc_partial_bytes(long long ngl) { return align_up((size_t)((ngl + XC_THREADS - 1) / XC_THREADS) * 2 * sizeof(double), 256); }

}  // namespace

extern "C" int isdf_xc_code(const char* name)
{
    if (!name) return -1;
    std::string s(name);
    for (auto& c : s) c = (char)tolower((unsigned char)c);
    if (s == "lda," || s == "slater,") return isdf_xc::XC_LDA_X;
    if (s == "lda,vwn" || s == "lda,vwn5" || s == "slater,vwn" || s == "svwn") return isdf_xc::XC_LDA_VWN;
    if (s == "blyp" || s == "b88,lyp") return isdf_xc::XC_BLYP;
    if (s == "pbe" || s == "pbe,pbe") return isdf_xc::XC_PBE;
    if (s == "pbe0") return isdf_xc::XC_PBE0;
    if (s == "b3lyp") return isdf_xc::XC_B3LYP;
    return -1;
}

extern "C" double isdf_xc_hybrid_fraction(int code) { return isdf_xc::hybrid_fraction(code); }

extern "C" size_t isdf_xc_ws_bytes(int nao, long long ldphi)
{
    return align_up((size_t)nao * (size_t)ldphi * sizeof(double), 256) + xc_partial_bytes(ldphi)
           + align_up((size_t)nao * nao * sizeof(double), 256) + 512;
}

extern "C" int isdf_xc_density(isdf_handle* h, const double* phi4, long long ldphi, long long comp_stride,
                               int nao, long long ngl, const double* dm, double* rho4, long long ldr,
                               void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi4 && dm && rho4 && ws && nao > 0 && ngl > 0 && ldphi >= ngl && ldr >= ngl);
    ARG_CHECK(ngl < 0x7fffffffLL && comp_stride >= (long long)nao * ldphi);
    if (ws_bytes < isdf_xc_ws_bytes(nao, ldphi)) { isdf_set_error("isdf_xc_density: workspace too small"); return ISDF_ERR_WS; }
    double* T = (double*)align_up((size_t)ws, 256);
    GemmParams p;       // T[mu,g] = sum_nu D[mu,nu] phi[nu,g]
    p.M = nao; p.N = (int)ngl; p.K = nao; p.A = dm; p.lda = nao; p.a_kmajor = 0;
    p.B = phi4; p.ldb = ldphi; p.b_kmajor = 1; p.C = T; p.ldc = ldphi; p.Cin = nullptr; p.ldcin = 0;
    p.E = nullptr; p.lde = 0; p.alpha = 1.0; p.neg = 0; p.epi = EPI_STORE; p.upper_only = 0;
    int rc;
    if ((rc = gemm_launch(h, p))) return rc;
    xc_rho_kernel<<<(unsigned)((ngl + XC_THREADS - 1) / XC_THREADS), XC_THREADS, 0, h->stream>>>(
        phi4, ldphi, comp_stride, T, nao, (int)ngl, rho4, ldr);
    h->other_launches++;
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

extern "C" int isdf_xc_potential(isdf_handle* h, int code, const double* rho4, long long ldr, long long ngl,
                                 double weight, double* wv4, double* d_sums, void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && rho4 && wv4 && d_sums && ws && ngl > 0 && ldr >= ngl && ngl < 0x7fffffffLL);
    ARG_CHECK(code >= isdf_xc::XC_LDA_X && code <= isdf_xc::XC_B3LYP);
    if (ws_bytes < xc_partial_bytes(ngl) + 256) { isdf_set_error("isdf_xc_potential: workspace too small"); return ISDF_ERR_WS; }
    double* partial = (double*)align_up((size_t)ws, 256);
    const int nblk = (int)((ngl + XC_THREADS - 1) / XC_THREADS);
    xc_point_kernel<<<nblk, XC_THREADS, 0, h->stream>>>(code, rho4, ldr, (int)ngl, weight, wv4, partial);
    xc_sum_kernel<<<1, XC_THREADS, 0, h->stream>>>(partial, nblk, d_sums);
    h->other_launches += 2;
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}

extern "C" int isdf_xc_matrix(isdf_handle* h, const double* phi4, long long ldphi, long long comp_stride,
                              int nao, long long ngl, const double* wv4, long long ldr, double* vmat,
                              void* ws, size_t ws_bytes)
{
    ARG_CHECK(h && phi4 && wv4 && vmat && ws && nao > 0 && ngl > 0 && ldphi >= ngl && ldr >= ngl);
    ARG_CHECK(ngl < 0x7fffffffLL && comp_stride >= (long long)nao * ldphi);
    if (ws_bytes < isdf_xc_ws_bytes(nao, ldphi)) { isdf_set_error("isdf_xc_matrix: workspace too small"); return ISDF_ERR_WS; }
    char* base = (char*)align_up((size_t)ws, 256);
    double* aow = (double*)base;
    double* mat = (double*)(base + align_up((size_t)nao * (size_t)ldphi * sizeof(double), 256));
    dim3 g((unsigned)((ngl + XC_THREADS - 1) / XC_THREADS), (unsigned)std::min(nao, 64));
    xc_aow_kernel<<<g, XC_THREADS, 0, h->stream>>>(phi4, ldphi, comp_stride, wv4, ldr, nao, (int)ngl, aow);
    h->other_launches++;
    GemmParams p;       // mat = Phi aow^T
    p.M = nao; p.N = nao; p.K = (int)ngl; p.A = phi4; p.lda = ldphi; p.a_kmajor = 0;
    p.B = aow; p.ldb = ldphi; p.b_kmajor = 0; p.C = mat; p.ldc = nao; p.Cin = nullptr; p.ldcin = 0;
    p.E = nullptr; p.lde = 0; p.alpha = 1.0; p.neg = 0; p.epi = EPI_STORE; p.upper_only = 0;
    int rc;
    if ((rc = gemm_launch(h, p))) return rc;
    xc_symm_kernel<<<dim3((unsigned)((nao + 127) / 128), (unsigned)nao), 128, 0, h->stream>>>(mat, nao, vmat);
    h->other_launches++;
    CUDA_TRY(cudaGetLastError());
    return ISDF_OK;
}
