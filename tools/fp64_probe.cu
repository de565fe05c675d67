// FP64 pipe probe for B200 (sm_100a): (1) is mma.sync f64 (DMMA) bit-identical to
// an ascending-k FMA chain?  (2) peak DFMA and DMMA throughput.  Results feed
// DESIGN.md (choice of the GEMM inner product) -- development tool, not product.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b, double c0, double c1)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};\n"
                 : "=d"(d0), "=d"(d1) : "d"(a), "d"(b), "d"(c0), "d"(c1));
}

// A[8][4] row-major, B[4][8] (k x n), C[8][8]
__global__ void exact_kernel(const double* A, const double* B, const double* C, double* D, int ntrial)
{
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    for (int tr = blockIdx.x; tr < ntrial; tr += gridDim.x) {
        const double* a = A + tr * 32; const double* b = B + tr * 32; const double* c = C + tr * 64;
        double d0, d1;
        dmma884(d0, d1, a[g * 4 + t], b[t * 8 + g], c[g * 8 + 2 * t], c[g * 8 + 2 * t + 1]);
        D[tr * 64 + g * 8 + 2 * t] = d0;
        D[tr * 64 + g * 8 + 2 * t + 1] = d1;
    }
}

template <int ILP>
__global__ void dfma_tput(double* out, int iters)
{
    double acc[ILP];
    double a = 1.0 + threadIdx.x * 1e-9, b = 0.999999;
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = fma(a, acc[i], b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void dmma_tput(double* out, int iters)
{
    double c0[ILP], c1[ILP];
    double a = 1.0 + threadIdx.x * 1e-9, b = 1e-3;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c0[i] = i; c1[i] = -i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) dmma884(c0[i], c1[i], a, b, c0[i], c1[i]);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

static double rnd(int wide)
{
    double m = (double)rand() / RAND_MAX * 2.0 - 1.0;
    int e = wide ? (rand() % 60) - 30 : 0;
    return ldexp(m, e);
}

int main()
{
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    printf("device %s sm_%d%d SMs %d clock %d kHz\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount, prop.clockRate);
    // ---- exactness ----
    for (int wide = 0; wide < 2; ++wide) {
        const int nt = 20000;
        std::vector<double> A(nt * 32), B(nt * 32), C(nt * 64), D(nt * 64);
        srand(1234 + wide);
        for (auto& x : A) x = rnd(wide);
        for (auto& x : B) x = rnd(wide);
        for (auto& x : C) x = rnd(wide);
        double *dA, *dB, *dC, *dD;
        cudaMalloc(&dA, A.size() * 8); cudaMalloc(&dB, B.size() * 8); cudaMalloc(&dC, C.size() * 8); cudaMalloc(&dD, D.size() * 8);
        cudaMemcpy(dA, A.data(), A.size() * 8, cudaMemcpyHostToDevice);
        cudaMemcpy(dB, B.data(), B.size() * 8, cudaMemcpyHostToDevice);
        cudaMemcpy(dC, C.data(), C.size() * 8, cudaMemcpyHostToDevice);
        exact_kernel<<<256, 32>>>(dA, dB, dC, dD, nt);
        cudaMemcpy(D.data(), dD, D.size() * 8, cudaMemcpyDeviceToHost);
        long asc = 0, desc = 0, pair = 0, total = 0;
        for (int tr = 0; tr < nt; ++tr)
            for (int m = 0; m < 8; ++m)
                for (int n = 0; n < 8; ++n) {
                    const double* a = &A[tr * 32 + m * 4]; const double* b = &B[tr * 32];
                    double c = C[tr * 64 + m * 8 + n], d = D[tr * 64 + m * 8 + n];
                    double r1 = c; for (int k = 0; k < 4; ++k) r1 = fma(a[k], b[k * 8 + n], r1);
                    double r2 = c; for (int k = 3; k >= 0; --k) r2 = fma(a[k], b[k * 8 + n], r2);
                    double r3 = fma(a[1], b[8 + n], a[0] * b[n]) + fma(a[3], b[24 + n], a[2] * b[16 + n]) + c;
                    total++;
                    asc += (memcmp(&r1, &d, 8) == 0); desc += (memcmp(&r2, &d, 8) == 0); pair += (memcmp(&r3, &d, 8) == 0);
                }
        printf("DMMA m8n8k4 exactness (wide=%d): ascending-chain match %ld/%ld, descending %ld, pairwise %ld\n",
               wide, asc, total, desc, pair);
        cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFree(dD);
    }
    // ---- throughput ----
    double* out; cudaMalloc(&out, 148 * 8 * 1024 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int bpsm = 1; bpsm <= 2; ++bpsm) {
        for (int rep = 0; rep < 2; ++rep) {
            dfma_tput<16><<<prop.multiProcessorCount * bpsm, 512>>>(out, iters);
            cudaEventRecord(e0);
            dfma_tput<16><<<prop.multiProcessorCount * bpsm, 512>>>(out, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fl = 2.0 * 16 * iters * 512.0 * prop.multiProcessorCount * bpsm;
            if (rep) printf("DFMA  %d CTA/SM x 512 thr: %.2f TFLOP/s (%.3f ms)\n", bpsm, fl / ms / 1e9, ms);
            dmma_tput<8><<<prop.multiProcessorCount * bpsm, 512>>>(out, iters);
            cudaEventRecord(e0);
            dmma_tput<8><<<prop.multiProcessorCount * bpsm, 512>>>(out, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
            fl = 2.0 * 8 * 8 * 4 * 8.0 * iters * 16.0 * prop.multiProcessorCount * bpsm;   // 8 mma/iter/warp, 16 warps
            if (rep) printf("DMMA  %d CTA/SM x 512 thr: %.2f TFLOP/s (%.3f ms)\n", bpsm, fl / ms / 1e9, ms);
        }
    }
    // DMMA rate vs resident warps and independent accumulators (GEMM-like: 32 accumulators)
    for (int thr = 64; thr <= 1024; thr *= 2) {
        float ms;
        dmma_tput<32><<<prop.multiProcessorCount, thr>>>(out, 4000);
        cudaEventRecord(e0);
        dmma_tput<32><<<prop.multiProcessorCount, thr>>>(out, 4000);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 256 * 32.0 * 4000 * (thr / 32) * prop.multiProcessorCount;
        printf("DMMA  ILP32, %4d thr/SM (%2d warps): %.2f TFLOP/s\n", thr, thr / 32, fl / ms / 1e9);
    }
    printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
