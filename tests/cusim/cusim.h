// cusim.h — TEST INFRASTRUCTURE: a tiny CPU emulator of the CUDA execution model.
//
// Lets `pytest -m "not gpu"` run the *actual* kernel sources of modern-rzip_b200/csrc on the CPU
// (there is no GPU in the build container): tests/cusim/Makefile compiles the .cu files as C++
// with `-include cusim.h -DMRZ_CUSIM` into tests/cusim/libmrz_sim.so, which exports the same C-ABI
// as libmrz_gpu.so.  The product never builds, links or loads any of this; libmrz_gpu.so has no
// CPU path.
//
// Model: one block at a time; every CUDA thread is a fiber (own stack, hand-written context
// switch); __syncthreads and the *_sync warp collectives are rendezvous points.  Fibers are only
// switched at those points, so atomics are plain read-modify-writes.  The runtime API subset the
// host code uses (malloc/memcpy/memset/streams/events) is mapped onto libc.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <functional>

// ---------------------------------------------------------------- keywords
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __shared__ static
#define __constant__ static
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct uint3_ { unsigned x, y, z; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};
struct uint4 { unsigned x, y, z, w; };
struct alignas(16) longlong2 { long long x, y; };
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { return uint4{a, b, c, d}; }

// ---------------------------------------------------------------- runtime API subset
typedef int cudaError_t;
typedef void *cudaStream_t;
struct cusim_event { double t; };
typedef cusim_event *cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2 };
enum { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };
enum { cudaStreamNonBlocking = 1, cudaHostAllocDefault = 0 };
static inline const char *cudaGetErrorString(cudaError_t) { return "cusim error"; }
static inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaGetLastError() { return 0; }
enum { cudaDevAttrMultiProcessorCount = 16 };
static inline cudaError_t cudaDeviceGetAttribute(int *v, int, int) { *v = 4; return 0; }
template <class T> static inline cudaError_t cudaMalloc(T **p, size_t n) {
    *p = (T *)aligned_alloc(256, (n + 255) / 256 * 256 + 256);
    if (*p && getenv("CUSIM_POISON")) memset((void *)*p, 0xA5, (n + 255) / 256 * 256 + 256);   // recycled device memory is not zero
    return *p ? 0 : cudaErrorMemoryAllocation;
}
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaHostAlloc(void **p, size_t n, int) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
static inline cudaError_t cudaFreeHost(void *p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, int) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, int, cudaStream_t) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, int) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new cusim_event{0}; return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) {
    timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    e->t = ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
    return 0;
}
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return 0; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return 0; }
#define cudaMemcpyToSymbol(sym, src, n) (memcpy(&(sym), (src), (n)), 0)
#define cudaGetSymbolAddress(pp, sym) (*(pp) = (void *)(sym), 0)

// ---------------------------------------------------------------- fibers
namespace cusim {
struct Fiber {
    void *sp;            // saved stack pointer
    char *stack;
    int status;          // 0 runnable, 1 waiting at block barrier, 2 waiting at warp rendezvous, 3 done, 4 grid barrier
    int block;           // index of its block (cooperative launches)
    unsigned bar_gen;    // generation waited for
    unsigned warp_gen;
    volatile unsigned *wait_addr; unsigned wait_old;   // status 5: waiting for *wait_addr != wait_old
    uint3_ tid;
    int lane, warp;
    int line;            // source line of the last barrier / warp collective this fiber entered (deadlock reports)
};
struct WarpSlot {
    uint64_t in[32], out[32];
    int pin[32];
    unsigned arrived, gen, out_ballot;
};
struct Block {
    Fiber *fib;
    WarpSlot *warps;
    int nthreads, nlive, bar_arrived;
    unsigned bar_gen;
    uint3_ bid;
    dim3 bdim, gdim;
    const std::function<void()> *body;
    void *smem;          // dynamic shared memory of a cooperative launch (one per block)
    int cluster, crank, csize;   // thread-block cluster (launch_cluster): id, rank in cluster, CTAs per cluster
};
extern thread_local Fiber *cur;
extern thread_local Block *blkp;     // block of the running fiber
void yield_to_scheduler();
void launch(dim3 grid, dim3 block, const std::function<void()> &body);
// all blocks resident at once (cudaLaunchCooperativeKernel): grid-wide barriers become possible
void launch_coop(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()> &body);
void sync_grid();                    // called by ONE thread of every block
// thread-block clusters: consecutive groups of `cl` blocks, all resident; barrier.cluster == sync_cluster()
void launch_cluster(dim3 grid, dim3 block, int cl, size_t smem_bytes, const std::function<void()> &body);
void sync_cluster();                 // called by ONE thread of every block of the cluster
void *map_shared_rank(void *p, int rank);   // address of the same dynamic-shared-memory object in the block of `rank` of this cluster
static inline int cluster_rank() { return blkp->crank; }
static inline int cluster_size() { return blkp->csize; }
// spin-wait of a cooperative kernel: the fiber sleeps until *addr != old (polled by the scheduler, no busy switching)
unsigned wait_change(volatile unsigned *addr, unsigned old);
static inline void *dyn_smem() { return blkp->smem; }

static inline void sync_block() {
    Fiber *f = cur;
    Block &b = *blkp;
    if (++b.bar_arrived == b.nlive) { b.bar_arrived = 0; b.bar_gen++; return; }
    f->status = 1; f->bar_gen = b.bar_gen;
    yield_to_scheduler();
}
// Rendezvous of the lanes in `mask`; every lane contributes (v, p) and then sees all lanes' values.
static inline WarpSlot &warp_exchange(unsigned mask, uint64_t v, int p) {
    Fiber *f = cur;
    WarpSlot &w = blkp->warps[f->warp];
    w.in[f->lane] = v; w.pin[f->lane] = p;
    w.arrived |= 1u << f->lane;
    if ((w.arrived & mask) == mask) {
        unsigned b = 0;
        for (int l = 0; l < 32; l++) {
            if (!((mask >> l) & 1u)) continue;
            w.out[l] = w.in[l];
            if (w.pin[l]) b |= 1u << l;
        }
        w.out_ballot = b;
        w.arrived &= ~mask;
        w.gen++;
        return w;
    }
    f->status = 2; f->warp_gen = w.gen;
    yield_to_scheduler();
    return w;
}
}  // namespace cusim

#define threadIdx (cusim::cur->tid)
#define blockIdx (cusim::blkp->bid)
#define blockDim (cusim::blkp->bdim)
#define gridDim (cusim::blkp->gdim)

static inline void cusim__syncthreads() { cusim::sync_block(); }
static inline void cusim__syncwarp(unsigned mask = 0xffffffffu) { cusim::warp_exchange(mask, 0, 0); }
static inline unsigned cusim__ballot_sync(unsigned mask, int pred) { return cusim::warp_exchange(mask, 0, pred).out_ballot & mask; }
static inline int __any_sync(unsigned mask, int pred) { return cusim__ballot_sync(mask, pred) != 0; }
static inline int __all_sync(unsigned mask, int pred) { return cusim__ballot_sync(mask, pred) == mask; }

template <class T> static inline uint64_t cusim_bits(T v) { uint64_t b = 0; memcpy(&b, &v, sizeof v); return b; }
template <class T> static inline T cusim_val(uint64_t b) { T v; memcpy(&v, &b, sizeof v); return v; }
template <class T> static inline T cusim__shfl_sync(unsigned mask, T v, int src, int width = 32) {
    int lane = cusim::cur->lane;
    int s = (lane & ~(width - 1)) | (src & (width - 1));
    return cusim_val<T>(cusim::warp_exchange(mask, cusim_bits(v), 0).out[s]);
}
template <class T> static inline T cusim__shfl_up_sync(unsigned mask, T v, unsigned d, int width = 32) {
    int lane = cusim::cur->lane;
    cusim::WarpSlot &w = cusim::warp_exchange(mask, cusim_bits(v), 0);
    int s = lane - (int)d;
    return (s < (lane & ~(width - 1))) ? v : cusim_val<T>(w.out[s]);
}
template <class T> static inline T cusim__shfl_down_sync(unsigned mask, T v, unsigned d, int width = 32) {
    int lane = cusim::cur->lane;
    cusim::WarpSlot &w = cusim::warp_exchange(mask, cusim_bits(v), 0);
    int s = lane + (int)d;
    return (s > (lane | (width - 1))) ? v : cusim_val<T>(w.out[s]);
}
template <class T> static inline T cusim__shfl_xor_sync(unsigned mask, T v, int x, int width = 32) {
    int lane = cusim::cur->lane;
    (void)width;
    return cusim_val<T>(cusim::warp_exchange(mask, cusim_bits(v), 0).out[lane ^ x]);
}

// every rendezvous records its source line in the fiber, so that a deadlock report can say who waits where
#define __syncthreads() (cusim::cur->line = __LINE__, cusim__syncthreads())
#define __syncwarp(...) (cusim::cur->line = __LINE__, cusim__syncwarp(__VA_ARGS__))
#define __ballot_sync(...) (cusim::cur->line = __LINE__, cusim__ballot_sync(__VA_ARGS__))
#define __shfl_sync(...) (cusim::cur->line = __LINE__, cusim__shfl_sync(__VA_ARGS__))
#define __shfl_up_sync(...) (cusim::cur->line = __LINE__, cusim__shfl_up_sync(__VA_ARGS__))
#define __shfl_down_sync(...) (cusim::cur->line = __LINE__, cusim__shfl_down_sync(__VA_ARGS__))
#define __shfl_xor_sync(...) (cusim::cur->line = __LINE__, cusim__shfl_xor_sync(__VA_ARGS__))

// ---------------------------------------------------------------- intrinsics
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
template <class T> static inline T __ldg(const T *p) { return *p; }
template <class T> static inline T __ldcg(const T *p) { return *(const volatile T *)p; }
static inline uint4 __ldcg(const uint4 *p) { uint4 v; const volatile unsigned *q = (const volatile unsigned *)p; v.x = q[0]; v.y = q[1]; v.z = q[2]; v.w = q[3]; return v; }
static inline longlong2 __ldcg(const longlong2 *p) { longlong2 v; v.x = ((const volatile long long *)p)[0]; v.y = ((const volatile long long *)p)[1]; return v; }
template <class T> static inline void __stcg(T *p, T v) { *p = v; }
static inline long long clock64() { return (long long)__builtin_ia32_rdtsc(); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __nanosleep(unsigned) {}

template <class T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicMin(T *p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T> static inline T atomicMax(T *p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T> static inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; return o; }
template <class T> static inline T atomicAnd(T *p, T v) { T o = *p; *p = o & v; return o; }
template <class T> static inline T atomicExch(T *p, T v) { T o = *p; *p = v; return o; }
template <class T> static inline T atomicCAS(T *p, T c, T v) { T o = *p; if (o == c) *p = v; return o; }

#define MRZ_LAUNCH(kernel, grid, block, stream, ...) \
    cusim::launch(dim3(grid), dim3(block), [=]() { kernel(__VA_ARGS__); })
#define MRZ_LAUNCH_COOP(kernel, grid, block, smem, stream, ...) \
    cusim::launch_coop(dim3(grid), dim3(block), (smem), [=]() { kernel(__VA_ARGS__); })
#define MRZ_LAUNCH_CLUSTER(kernel, grid, cl, block, smem, stream, ...) \
    cusim::launch_cluster(dim3(grid), dim3(block), (cl), (smem), [=]() { kernel(__VA_ARGS__); })
