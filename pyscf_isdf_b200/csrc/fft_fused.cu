// FUSED = true instantiations of the stage-4 FFT kernels (Rader lines through rader_fused_smem,
// fft_core.cuh) and their launchers.  Opt-in: ISDF_FFT_FUSED=1 makes fft.cu build fused plans.
// Separate translation unit so that the default kernels in fft.cu compile exactly as before.
#include "common.cuh"
#include "fft_kernels.cuh"

namespace {

template <typename K>
int set_smem_f(K kern, size_t smem)
{
    return isdf_func_max_smem((const void*)kern, smem);
}

}  // namespace

namespace isdf_fft {

int fused_z_r2c(dim3 grid, size_t smem, cudaStream_t s, const double* in, long long ldin, double2* out,
                long long ldout, FftPlan pl, int nlines, int LB)
{
    int rc;
    if ((rc = set_smem_f(fft_z_r2c_kernel<true>, smem))) return rc;
    fft_z_r2c_kernel<true><<<grid, FFT_THREADS, smem, s>>>(in, ldin, out, ldout, pl, nlines, LB);
    return ISDF_OK;
}

int fused_z_c2r(dim3 grid, size_t smem, cudaStream_t s, const double2* in, long long ldin, double* out,
                long long ldout, FftPlan pl, int nlines, int LB)
{
    int rc;
    if ((rc = set_smem_f(fft_z_c2r_kernel<true>, smem))) return rc;
    fft_z_c2r_kernel<true><<<grid, FFT_THREADS, smem, s>>>(in, ldin, out, ldout, pl, nlines, LB);
    return ISDF_OK;
}

int fused_cols(int mode, dim3 grid, size_t smem, cudaStream_t s, double2* data, long long vec_stride,
               long long outer_stride, FftPlan pl, int ncols, int CH, CoulGeom geo)
{
    int rc;
#define ISDF_FUSED_COLS(M_)                                                                               \
    do {                                                                                                  \
        if ((rc = set_smem_f(fft_cols_kernel<M_, true>, smem))) return rc;                                \
        fft_cols_kernel<M_, true><<<grid, FFT_THREADS, smem, s>>>(data, vec_stride, outer_stride, pl, ncols, CH, geo); \
    } while (0)
    switch (mode) {
    case 0: ISDF_FUSED_COLS(0); break;
    case 1: ISDF_FUSED_COLS(1); break;
    case 2: ISDF_FUSED_COLS(2); break;
    default: ISDF_FUSED_COLS(3); break;
    }
#undef ISDF_FUSED_COLS
    return ISDF_OK;
}

}  // namespace isdf_fft
