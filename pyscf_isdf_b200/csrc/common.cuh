// Shared helpers for the ISDF B200 library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#define ISDF_OK 0
#define ISDF_ERR_ARG (-1)
#define ISDF_ERR_CUDA (-2)
#define ISDF_ERR_WS (-3)
#define ISDF_ERR_NUMERIC (-4)

void isdf_set_error(const char* fmt, ...);

// Kernel attributes are per function AND per device, and several handles (threads, devices) may
// share one process: these two are the only places the library sets them.
//  isdf_func_max_smem: raises cudaFuncAttributeMaxDynamicSharedMemorySize of `func` on the current
//    device to at least `bytes`; never lowers it (a launch in flight on another thread keeps its limit).
//  isdf_func_attr_once: sets `attr` = `value` once per (function, device); returns the first call's result.
int isdf_func_max_smem(const void* func, size_t bytes);
cudaError_t isdf_func_attr_once(const void* func, int attr, int value);
// runs `fn(arg)` once per (key, current device) under the library's set-up lock; a failing fn is retried
// by the next call.  For constant-memory tables and similar per-device set-up.
int isdf_once_per_device(const void* key, int (*fn)(void*), void* arg);

#define CUDA_TRY(expr)                                                        \
    do {                                                                      \
        cudaError_t _e = (expr);                                              \
        if (_e != cudaSuccess) {                                              \
            isdf_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,       \
                           cudaGetErrorString(_e));                           \
            return ISDF_ERR_CUDA;                                             \
        }                                                                     \
    } while (0)

#define ARG_CHECK(cond)                                                       \
    do {                                                                      \
        if (!(cond)) {                                                        \
            isdf_set_error("%s:%d bad argument: %s", __FILE__, __LINE__, #cond); \
            return ISDF_ERR_ARG;                                              \
        }                                                                     \
    } while (0)

#define NSTAGE_TIMERS 16

struct isdf_handle {
    int device;
    int num_sms;
    cudaStream_t stream;
    cudaEvent_t ev0[NSTAGE_TIMERS], ev1[NSTAGE_TIMERS];
    bool ev_used[NSTAGE_TIMERS];
    // scratch owned by the handle (tiny: scalars, histograms, pinned mailbox)
    void* d_small;       // 1 MiB device scratch for scalars / histograms
    void* h_pinned;      // 64 KiB pinned host mailbox
    long long gemm_launches;   // kernel-launch counter (bench reports it)
    long long other_launches;
    // optional per-launch GEMM timing (bench.py roofline): event pairs + flop sum
    int prof_on, prof_n, prof_cap;
    cudaEvent_t* prof_ev;      // 2 * prof_cap events
    double prof_flops;
    int gemm_force;            // 0 auto, 2..5: CTA tile / ring depth (tuning knob, see isdf_set_gemm_tile)
    // tables of the stage-4 transforms (twiddles, DFT matrices, Coulomb weights), kept between calls for the
    // last mesh so that the per-batch calls of a build do not allocate, copy and synchronise each time;
    // owned by fft.cu, released through fft_cache_free by isdf_destroy
    void* fft_cache;
    void (*fft_cache_free)(void*);
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bump allocator over a caller-provided workspace
struct WsAlloc {
    char* base; size_t size; size_t off;
    WsAlloc(void* p, size_t n) : base((char*)p), size(n), off(0) {}
    template <typename T> T* get(size_t count) {
        off = align_up(off, 256);
        T* r = (T*)(base + off);
        off += count * sizeof(T);
        return r;
    }
    bool ok() const { return off <= size; }
};

// ---------------------------------------------------------------- GEMM ----
// C[m,n] (row-major, ldc) = epi( Cin[m,n] (optional) (+/-) sum_k opA(k,m) * opB(k,n) )
// The k-sum of every output element is ONE fused-multiply-add chain in
// ascending k starting from Cin (or 0): this fixed order is what makes the
// interpolation-point selection bit-reproducible against the C oracle.
enum { EPI_STORE = 0, EPI_SQUARE = 1, EPI_MUL = 2 };

struct GemmParams {
    int M, N, K;
    const double* A; long long lda; int a_kmajor;   // kmajor: A[k*lda+m] else A[m*lda+k]
    const double* B; long long ldb; int b_kmajor;   // kmajor: B[k*ldb+n] else B[n*ldb+k]
    double* C; long long ldc;
    const double* Cin; long long ldcin;             // nullable accumulator init
    const double* E; long long lde;                 // EPI_MUL operand
    double alpha;                                   // EPI_STORE: C = alpha*acc
    int neg;                                        // acc -= a*b instead of +=
    int epi;
    int upper_only;                                 // skip tiles strictly below the diagonal
};

int gemm_launch(isdf_handle* h, const GemmParams& p);
// gemm_tma.cu: the same product with TMA-fed operand tiles (form 7..10 = CTA tile, see there)
bool gemm_tma_applicable(const GemmParams& p);
int gemm_tma_launch(isdf_handle* h, const GemmParams& p, int form);
