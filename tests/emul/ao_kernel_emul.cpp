// CPU harness that runs eval_ao_deriv1_kernel (pyscf_isdf_b200/csrc/eval_ao_deriv1_kernel.cuh) itself:
// one CTA at a time, its 64 threads as real threads, __syncthreads as a barrier, the two warps' shuffles
// and ballots as exchanges through a per-warp buffer.  TEST INFRASTRUCTURE ONLY (tests/test_ao_emul_cpu.py):
// checks the image screening, the list compaction, the slab offsets and the output indexing of the kernel
// against the oracle; it says nothing about speed.
#include <barrier>
#include <cmath>
#include <thread>
#include <vector>
#include <algorithm>
using std::exp; using std::sqrt; using std::fmin; using std::fmax; using std::min;

#define __host__
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __shared__ static
#define __launch_bounds__(x)

struct EmulDim3 { int x = 0, y = 0, z = 0; };
static thread_local EmulDim3 threadIdx;
static EmulDim3 blockIdx;
static std::barrier<>* cta_bar = nullptr;
static std::barrier<>* warp_bar[2] = {nullptr, nullptr};
static double xch_d[2][32];
static unsigned xch_b[2][32];

static inline void __syncthreads() { cta_bar->arrive_and_wait(); }
static inline double __shfl_xor_sync(unsigned, double v, int o)
{
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    xch_d[w][l] = v;
    warp_bar[w]->arrive_and_wait();
    const double r = xch_d[w][l ^ o];
    warp_bar[w]->arrive_and_wait();
    return r;
}
static inline unsigned __ballot_sync(unsigned, bool pred)
{
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    xch_b[w][l] = pred ? 1u : 0u;
    warp_bar[w]->arrive_and_wait();
    unsigned r = 0;
    for (int i = 0; i < 32; ++i) r |= xch_b[w][i] << i;
    warp_bar[w]->arrive_and_wait();
    return r;
}
static inline int __popc(unsigned v) { return __builtin_popcount(v); }

#include "../../pyscf_isdf_b200/csrc/eval_ao_deriv1_kernel.cuh"

// same arguments as the C-ABI entry isdf_eval_ao_deriv1 (host pointers throughout); ny = grid.y
extern "C" int emul_eval_ao_deriv1(const int* atm, int natm, const int* bas, int nbas, const double* env, int nenv,
                                   const double* a, const int* mesh, const double* Ls, int nimg, const double* rcut,
                                   long long g0, long long g1, double* phi, long long ldphi, long long comp_stride, int ny)
{
    using namespace isdf_ao_k;
    std::vector<DevAtom> atoms;
    std::vector<DevShell> shells;
    int nao = 0;
    if (!build_tables(atm, natm, bas, nbas, env, nenv, rcut, atoms, shells, &nao)) return 1;
    if (nimg > LIST_CAP) return 2;
    AoParams p;
    for (int i = 0; i < 3; ++i) p.mesh[i] = mesh[i];
    for (int i = 0; i < 9; ++i) p.a[i] = a[i];
    p.g0 = g0; p.g1 = g1; p.ng = (long long)mesh[0] * mesh[1] * mesh[2];
    p.natm = natm; p.nimg = nimg; p.ny = ny;
    p.atoms = atoms.data(); p.shells = shells.data(); p.env = env; p.Ls = Ls;
    p.phi = phi; p.ldphi = ldphi; p.cstride = comp_stride;
    const long long b0 = g0 / AO_BLK, b1 = (g1 - 1) / AO_BLK + 1;
    std::barrier<> cb(AO_THREADS), w0(32), w1(32);
    cta_bar = &cb; warp_bar[0] = &w0; warp_bar[1] = &w1;
    for (int by = 0; by < ny; ++by)
        for (long long bx = 0; bx < b1 - b0; ++bx) {
            blockIdx.x = (int)bx; blockIdx.y = by;
            std::vector<std::thread> th;
            for (int t = 0; t < AO_THREADS; ++t)
                th.emplace_back([t, p]() { threadIdx.x = t; eval_ao_deriv1_kernel(p); });
            for (auto& x : th) x.join();
        }
    return 0;
}
