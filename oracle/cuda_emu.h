/*
 * cuda_emu.h -- TEST INFRASTRUCTURE (oracle/).  Just enough of the CUDA execution model to run a
 * kernel of ffmpeg_hybrid_b200/csrc/*.cuh on the host, unchanged, inside the CPU test-suite:
 * every CUDA thread of a CTA is an OS thread, __syncthreads() / __syncwarp(mask) are real barriers,
 * __shared__ variables are function-local statics (one CTA runs at a time), cp.async is a memcpy.
 * CTAs run one after the other in block-index order, so kernels whose CTAs wait for LATER CTAs cannot be run this way.
 *
 * Used by oracle/emu_inter.cc to pin the shared-memory inter-prediction kernel (k_inter_sb.cuh)
 * against the oracle restatement before it ever runs on a GPU.  Only tests/ execute it.
 */
#ifndef VP9_CUDA_EMU_H
#define VP9_CUDA_EMU_H

#include <atomic>
#include <climits>
#include <linux/futex.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#define VP9_EMU 1
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__ static
#define __restrict__
#define __launch_bounds__(...)
#define __grid_constant__
#define __align__(n) __attribute__((aligned(n)))

struct uint2 { unsigned x, y; };
struct uint3 { unsigned x, y, z; };
struct uint4 { unsigned x, y, z, w; };
struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return uint4{ a, b, c, d }; }
static inline uint2 make_uint2(unsigned a, unsigned b) { return uint2{ a, b }; }

namespace emu {

/* reusable barrier for a fixed number of threads: a counter and a generation word the waiters sleep on
 * (futex) -- one system call per waiter and phase; a mutex + condition variable cost several, contended */
struct Barrier {
    std::atomic<int> count{ 0 };
    std::atomic<unsigned> gen{ 0 };
    int need;
    explicit Barrier(int n) : need(n) {}
    void wait()
    {
        const unsigned g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == need) {
            count.store(0, std::memory_order_relaxed);
            gen.store(g + 1, std::memory_order_release);
            syscall(SYS_futex, (unsigned *)&gen, FUTEX_WAKE_PRIVATE, INT_MAX, nullptr, nullptr, 0);
        } else {
            while (gen.load(std::memory_order_acquire) == g)
                syscall(SYS_futex, (unsigned *)&gen, FUTEX_WAIT_PRIVATE, g, nullptr, nullptr, 0);
        }
    }
};

struct Cta {
    int nthreads = 0;
    Barrier *all = nullptr;
    std::mutex mu;
    std::map<uint64_t, Barrier *> warp_bar;     /* (warp, mask) -> barrier of popcount(mask) threads */
    std::vector<uint32_t> shfl;                 /* one exchange slot per thread */
    Barrier *warp_barrier(int warp, unsigned mask)
    {
        std::lock_guard<std::mutex> lk(mu);
        Barrier *&b = warp_bar[(uint64_t)warp << 32 | mask];
        if (!b) b = new Barrier(__builtin_popcount(mask));
        return b;
    }
    ~Cta() { for (auto &kv : warp_bar) delete kv.second; delete all; }
};

extern Cta *g_cta;
extern unsigned char *g_dyn_smem;

} /* namespace emu */

extern thread_local uint3 threadIdx, blockIdx;
extern thread_local dim3 blockDim, gridDim;

#ifdef VP9_EMU_IMPL
thread_local uint3 threadIdx, blockIdx;
thread_local dim3 blockDim, gridDim;
emu::Cta *emu::g_cta;
unsigned char *emu::g_dyn_smem;
#endif

static inline void __syncthreads() { emu::g_cta->all->wait(); }
static inline void __syncwarp(unsigned mask = 0xffffffffu)
{
    const int warp = threadIdx.x >> 5;
    /* a partial last warp: only the threads that exist can arrive */
    const int first = warp * 32, live = emu::g_cta->nthreads - first;
    if (live < 32) mask &= (1u << live) - 1;
    emu::g_cta->warp_barrier(warp, mask)->wait();
}
static inline int __shfl_sync(unsigned mask, int v, int src)
{
    emu::Cta *c = emu::g_cta;
    c->shfl[threadIdx.x] = (uint32_t)v;
    __syncwarp(mask);
    const int r = (int)c->shfl[(threadIdx.x & ~31u) + (src & 31)];
    __syncwarp(mask);
    return r;
}
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned atomicAdd(unsigned *p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline int atomicMax(int *p, int v)
{
    int old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
static inline int atomicExch(int *p, int v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
static inline void __threadfence_system() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline void __threadfence_block() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline long long clock64() { return 0; }
static inline void __nanosleep(unsigned) { std::this_thread::yield(); }
template <class T> static inline T __ldcg(const T *p) { return *p; }
static inline int __shfl_up_sync(unsigned mask, int v, int delta)
{
    emu::Cta *c = emu::g_cta;
    c->shfl[threadIdx.x] = (uint32_t)v;
    __syncwarp(mask);
    const int lane = threadIdx.x & 31;
    const int r = lane >= delta ? (int)c->shfl[threadIdx.x - delta] : v;
    __syncwarp(mask);
    return r;
}
static inline int __reduce_add_sync(unsigned mask, int v)
{
    emu::Cta *c = emu::g_cta;
    c->shfl[threadIdx.x] = (uint32_t)v;
    __syncwarp(mask);
    int s = 0;
    for (int l = 0; l < 32; l++)
        if (mask >> l & 1) s += (int)c->shfl[(threadIdx.x & ~31u) + l];
    __syncwarp(mask);
    return s;
}
static inline int __all_sync(unsigned mask, int pred)
{
    emu::Cta *c = emu::g_cta;
    c->shfl[threadIdx.x] = (uint32_t)(pred != 0);
    __syncwarp(mask);
    int r = 1;
    for (int l = 0; l < 32; l++)
        if ((mask >> l & 1) && !c->shfl[(threadIdx.x & ~31u) + l]) r = 0;
    __syncwarp(mask);
    return r;
}
static inline int atomicCAS(int *p, int cmp, int v)
{
    __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;
}
static inline int __dp2a_lo(int a, int b, int c) { return c + (int16_t)(a & 0xffff) * (int8_t)(b & 0xff) + (int16_t)(a >> 16) * (int8_t)(b >> 8 & 0xff); }
static inline int __dp2a_hi(int a, int b, int c) { return c + (int16_t)(a & 0xffff) * (int8_t)(b >> 16 & 0xff) + (int16_t)(a >> 16) * (int8_t)(b >> 24 & 0xff); }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) { s &= 31; return s ? lo >> s | hi << (32 - s) : lo; }
static inline unsigned __byte_perm(unsigned x, unsigned y, unsigned sel)
{
    const uint64_t v = (uint64_t)y << 32 | x;
    unsigned r = 0;
    for (int i = 0; i < 4; i++)
        r |= (unsigned)(v >> (8 * (sel >> (4 * i) & 7)) & 0xff) << (8 * i);
    return r;
}
static inline int __reduce_max_sync(unsigned mask, int v)
{
    emu::Cta *c = emu::g_cta;
    c->shfl[threadIdx.x] = (uint32_t)v;
    __syncwarp(mask);
    int s = v;
    for (int l = 0; l < 32; l++)
        if ((mask >> l & 1) && (int)c->shfl[(threadIdx.x & ~31u) + l] > s) s = (int)c->shfl[(threadIdx.x & ~31u) + l];
    __syncwarp(mask);
    return s;
}
static inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
namespace emu {
static inline void named_barrier(int id, int n)
{
    Barrier *b;
    {
        std::lock_guard<std::mutex> lk(g_cta->mu);
        Barrier *&q = g_cta->warp_bar[(uint64_t)0xffff0000u << 32 | (unsigned)n << 8 | (unsigned)id]   /* the Cta outlives one CTA: key by size too */;
        if (!q) q = new Barrier(n);
        b = q;
    }
    b->wait();
}
}

namespace emu {

/* run `body` as a grid of CTAs of `threads` OS threads each, one CTA after the other */
static inline void launch(dim3 grid, int threads, size_t dyn_smem, const std::function<void()> &body)
{
    std::vector<unsigned char> smem(dyn_smem + 64);
    g_dyn_smem = (unsigned char *)(((uintptr_t)smem.data() + 63) & ~(uintptr_t)63);
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                Cta cta;
                cta.nthreads = threads;
                cta.all = new Barrier(threads);
                cta.shfl.assign(threads, 0);
                g_cta = &cta;
                std::vector<std::thread> th;
                for (int t = 0; t < threads; t++)
                    th.emplace_back([=, &body] {
                        threadIdx = uint3{ (unsigned)t, 0, 0 };
                        blockIdx = uint3{ bx, by, bz };
                        blockDim = dim3(threads);
                        gridDim = grid;
                        body();
                    });
                for (auto &t : th) t.join();
                g_cta = nullptr;
            }
}

} /* namespace emu */

#define VP9_DYN_SMEM(name) unsigned char *name = emu::g_dyn_smem
#endif
