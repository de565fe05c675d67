// Handle management, error reporting and the generic GEMM entry point.
#include "common.cuh"
#include "../../include/isdf_b200.h"
#include <cstdarg>
#include <map>
#include <mutex>
#include <utility>

static thread_local char g_err[1024] = "";

void isdf_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* isdf_last_error(void) { return g_err; }

// ---- per-(function, device) kernel attributes and per-device set-up, safe for several threads ----
namespace {
std::mutex g_setup_mutex;
std::map<std::pair<const void*, int>, size_t> g_smem_set;                       // (func, device) -> bytes
std::map<std::pair<std::pair<const void*, int>, int>, cudaError_t> g_attr_set;  // ((func, attr), device)
std::map<std::pair<const void*, int>, bool> g_once_done;                        // (key, device)
}  // namespace

int isdf_func_max_smem(const void* func, size_t bytes)
{
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_setup_mutex);
    size_t& have = g_smem_set[std::make_pair(func, dev)];
    if (bytes > have) {
        CUDA_TRY(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        have = bytes;
    }
    return ISDF_OK;
}

cudaError_t isdf_func_attr_once(const void* func, int attr, int value)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(g_setup_mutex);
    auto key = std::make_pair(std::make_pair(func, attr), dev);
    auto it = g_attr_set.find(key);
    if (it != g_attr_set.end()) return it->second;
    e = cudaFuncSetAttribute(func, (cudaFuncAttribute)attr, value);
    if (e != cudaSuccess) (void)cudaGetLastError();
    g_attr_set[key] = e;
    return e;
}

int isdf_once_per_device(const void* key, int (*fn)(void*), void* arg)
{
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_setup_mutex);
    bool& done = g_once_done[std::make_pair(key, dev)];
    if (!done) {
        const int rc = fn(arg);
        if (rc) return rc;
        done = true;
    }
    return ISDF_OK;
}

extern "C" int isdf_create(int device, isdf_handle** out)
{
    ARG_CHECK(out != nullptr);
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    ARG_CHECK(device >= 0 && device < ndev);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        isdf_set_error("isdf_b200 is built for sm_100a only; device %d is sm_%d%d",
                       device, prop.major, prop.minor);
        return ISDF_ERR_ARG;
    }
    isdf_handle* h = new isdf_handle();
    memset(h, 0, sizeof(*h));
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->stream = 0;
    for (int i = 0; i < NSTAGE_TIMERS; ++i) {
        CUDA_TRY(cudaEventCreate(&h->ev0[i]));
        CUDA_TRY(cudaEventCreate(&h->ev1[i]));
    }
    CUDA_TRY(cudaMalloc(&h->d_small, 1 << 20));
    CUDA_TRY(cudaMallocHost(&h->h_pinned, 1 << 16));
    *out = h;
    return ISDF_OK;
}

extern "C" int isdf_destroy(isdf_handle* h)
{
    if (!h) return ISDF_OK;
    cudaSetDevice(h->device);
    for (int i = 0; i < NSTAGE_TIMERS; ++i) {
        cudaEventDestroy(h->ev0[i]);
        cudaEventDestroy(h->ev1[i]);
    }
    if (h->fft_cache && h->fft_cache_free) h->fft_cache_free(h->fft_cache);
    cudaFree(h->d_small);
    cudaFreeHost(h->h_pinned);
    for (int i = 0; i < 2 * h->prof_cap; ++i) cudaEventDestroy(h->prof_ev[i]);
    delete[] h->prof_ev;
    delete h;
    return ISDF_OK;
}

extern "C" int isdf_set_stream(isdf_handle* h, void* stream)
{
    ARG_CHECK(h != nullptr);
    h->stream = (cudaStream_t)stream;
    return ISDF_OK;
}

extern "C" int isdf_num_sms(isdf_handle* h) { return h ? h->num_sms : 0; }

extern "C" long long isdf_launch_count(isdf_handle* h, int which)
{
    if (!h) return 0;
    return which == 0 ? h->gemm_launches : h->other_launches;
}

extern "C" int isdf_stage_time_ms(isdf_handle* h, int stage, float* out)
{
    ARG_CHECK(h && out && stage >= 0 && stage < NSTAGE_TIMERS);
    if (!h->ev_used[stage]) { *out = -1.f; return ISDF_OK; }
    CUDA_TRY(cudaEventSynchronize(h->ev1[stage]));
    CUDA_TRY(cudaEventElapsedTime(out, h->ev0[stage], h->ev1[stage]));
    return ISDF_OK;
}

extern "C" int isdf_set_gemm_tile(isdf_handle* h, int which)
{
    ARG_CHECK(h != nullptr && (which == 0 || (which >= 2 && which <= 11)));
    h->gemm_force = which;
    return ISDF_OK;
}

// Per-launch GEMM timing for bench.py's roofline: enable -> every GEMM launch is
// bracketed by an event pair (up to 65536 launches); read -> total algorithmic
// flops, summed kernel time and launch count since enable.
extern "C" int isdf_gemm_profile(isdf_handle* h, int enable)
{
    ARG_CHECK(h != nullptr);
    if (enable && h->prof_cap == 0) {
        const int cap = 1 << 16;
        h->prof_ev = new cudaEvent_t[2 * cap];
        for (int i = 0; i < 2 * cap; ++i) CUDA_TRY(cudaEventCreate(&h->prof_ev[i]));
        h->prof_cap = cap;
    }
    if (enable) { h->prof_n = 0; h->prof_flops = 0.0; }
    h->prof_on = enable ? 1 : 0;
    return ISDF_OK;
}

extern "C" int isdf_gemm_profile_read(isdf_handle* h, double* h_flops, double* h_ms, long long* h_count)
{
    ARG_CHECK(h && h_flops && h_ms && h_count);
    double ms = 0.0;
    for (int i = 0; i < h->prof_n; ++i) {
        float t = 0.f;
        CUDA_TRY(cudaEventSynchronize(h->prof_ev[2 * i + 1]));
        CUDA_TRY(cudaEventElapsedTime(&t, h->prof_ev[2 * i], h->prof_ev[2 * i + 1]));
        ms += t;
    }
    *h_flops = h->prof_flops; *h_ms = ms; *h_count = h->prof_n;
    return ISDF_OK;
}

extern "C" int isdf_dgemm(isdf_handle* h, int M, int N, int K,
                          const double* A, long long lda, int a_kmajor,
                          const double* B, long long ldb, int b_kmajor,
                          double* C, long long ldc,
                          const double* Cin, long long ldcin,
                          const double* E, long long lde,
                          double alpha, int neg, int epi, int upper_only)
{
    ARG_CHECK(h != nullptr);
    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.A = A; p.lda = lda; p.a_kmajor = a_kmajor;
    p.B = B; p.ldb = ldb; p.b_kmajor = b_kmajor;
    p.C = C; p.ldc = ldc; p.Cin = Cin; p.ldcin = ldcin;
    p.E = E; p.lde = lde; p.alpha = alpha; p.neg = neg; p.epi = epi;
    p.upper_only = upper_only;
    return gemm_launch(h, p);
}
