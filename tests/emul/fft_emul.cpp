// CPU emulation harness for the shared-memory line FFTs (pyscf_isdf_b200/csrc/fft_core.cuh):
// one CTA = `nthreads` std::threads and a barrier, "shared memory" = two heap buffers.
// Exposes one C function for tests/test_fft_emul_cpu.py.  TEST INFRASTRUCTURE ONLY.
#include "cuda_shim.h"
#include "../../pyscf_isdf_b200/csrc/fft_core.cuh"
#include <thread>
#include <vector>

namespace {

template <int SIGN, bool FUSED>
void run_cta(double2* a, double2* b, const FftPlan& pl, int nlines, int ls, int es, bool line_fast,
             int nthreads, double2** result)
{
    std::barrier<> bar(nthreads);
    emul_barrier = &bar;
    blockDim.x = nthreads;
    std::vector<double2*> res(nthreads, nullptr);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&, t]() {
            threadIdx.x = t;
            res[t] = fft_smem<SIGN, FUSED>(a, b, pl, nlines, ls, es, line_fast);
        });
    for (auto& x : th) x.join();
    *result = res[0];
}

}  // namespace

// in / out: [nlines][n] complex (interleaved re, im).  sign -1 forward, +1 unnormalised inverse.
// line_fast 0: every line contiguous (the z pass: stride lalloc, element stride 1);
// line_fast 1: lines are columns of an [lalloc][nlines] matrix (the y / x passes).
// Returns 0, -1 for an unsupported length; *used_fused = 2 / 3 if the plan took the fused Rader / wide-radix path.
extern "C" int emul_fft(int n, int nlines, int sign, int fused, int line_fast, int nthreads,
                        const double* in, double* out, int* used_fused)
{
    FftPlan pl;
    std::vector<double2> tw;
    std::vector<int> iv;
    AxisOffsets off;
    if (build_axis_tables(n, fused, pl, tw, iv, off)) return -1;
    bind_axis_tables(pl, tw.data(), iv.data(), off);
    *used_fused = pl.rader >= 2 ? pl.rader : 0;
    const int LS = pl.lalloc();
    const int ls = line_fast ? 1 : LS, es = line_fast ? nlines : 1;
    std::vector<double2> A((size_t)LS * nlines, double2{1e300, 1e300}), B((size_t)LS * nlines, double2{-1e300, 1e300});
    for (int l = 0; l < nlines; ++l)
        for (int e = 0; e < n; ++e)
            A[(size_t)l * ls + (size_t)e * es] = double2{in[2 * ((size_t)l * n + e)], in[2 * ((size_t)l * n + e) + 1]};
    double2* r = nullptr;
    fused = pl.rader >= 2;            // the host dispatch of fft.cu: FUSED kernels run fused plans only
    if (sign < 0) {
        if (fused) run_cta<-1, true>(A.data(), B.data(), pl, nlines, ls, es, line_fast != 0, nthreads, &r);
        else run_cta<-1, false>(A.data(), B.data(), pl, nlines, ls, es, line_fast != 0, nthreads, &r);
    } else {
        if (fused) run_cta<+1, true>(A.data(), B.data(), pl, nlines, ls, es, line_fast != 0, nthreads, &r);
        else run_cta<+1, false>(A.data(), B.data(), pl, nlines, ls, es, line_fast != 0, nthreads, &r);
    }
    for (int l = 0; l < nlines; ++l)
        for (int e = 0; e < n; ++e) {
            const double2 v = r[(size_t)l * ls + (size_t)e * es];
            out[2 * ((size_t)l * n + e)] = v.x;
            out[2 * ((size_t)l * n + e) + 1] = v.y;
        }
    return 0;
}
