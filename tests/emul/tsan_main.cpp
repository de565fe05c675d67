// ThreadSanitizer driver for the emulated CTA (tests/test_fft_emul_cpu.py::test_no_data_races):
// every barrier-separated phase of the line FFTs is run by 5 real threads; a missing __syncthreads
// or two work items writing one shared-memory slot shows up as a reported race.
#include <cmath>
#include <cstdio>
#include <vector>
extern "C" int emul_fft(int n, int nlines, int sign, int fused, int line_fast, int nthreads,
                        const double* in, double* out, int* used_fused);
int main()
{
    const int lens[] = {71, 97, 35, 23, 11, 96, 108};
    for (int n : lens)
        for (int fused = 0; fused < 2; ++fused)
            for (int lf = 0; lf < 2; ++lf)
                for (int sign = -1; sign <= 1; sign += 2) {
                    const int nl = 6;
                    std::vector<double> in(2 * nl * n), out(2 * nl * n);
                    for (size_t i = 0; i < in.size(); ++i) in[i] = std::sin(0.37 * i + n);
                    int used = 0;
                    if (emul_fft(n, nl, sign, fused, lf, 5, in.data(), out.data(), &used)) return 2;
                }
    std::printf("tsan run done\n");
    return 0;
}
