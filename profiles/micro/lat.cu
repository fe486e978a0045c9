// Microbenchmark (B200): latency of the primitives the cluster automaton is built from.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lat lat.cu && ./lat
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
typedef long long i64;

__global__ void k_chase(const unsigned *a, int steps, int mode, i64 *out, unsigned *sink) {
    unsigned idx = threadIdx.x;
    i64 t0 = clock64();
    for (int s = 0; s < steps; s++) {
        if (mode == 0) idx = *(volatile const unsigned *)(a + idx);
        else if (mode == 1) idx = __ldcg(a + idx);
        else if (mode == 2) idx = __ldg(a + idx);
        else idx = atomicAdd((unsigned *)a + idx, 0u);
    }
    i64 t1 = clock64();
    if (threadIdx.x == 0) out[0] = (t1 - t0) / steps;
    *sink = idx;
}

__global__ void k_cluster(int iters, i64 *out, unsigned *sink) {
    cg::cluster_group cl = cg::this_cluster();
    __shared__ unsigned sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (i * 37 + 11) & 1023;
    cl.sync();
    unsigned r = cl.block_rank(), n = cl.num_blocks();
    // barrier cost
    i64 t0 = clock64();
    for (int i = 0; i < iters; i++)
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    i64 t1 = clock64();
    // remote smem pointer chase (rank r reads rank (r+1)%n)
    volatile unsigned *rem = cl.map_shared_rank(sm, (r + 1) % n);
    unsigned idx = threadIdx.x & 1023;
    i64 t2 = clock64();
    for (int i = 0; i < iters; i++) idx = rem[idx];
    i64 t3 = clock64();
    // local smem pointer chase through a generic volatile pointer
    volatile unsigned *loc = sm;
    for (int i = 0; i < iters; i++) idx = loc[idx];
    i64 t4 = clock64();
    // __syncthreads
    for (int i = 0; i < iters; i++) __syncthreads();
    i64 t5 = clock64();
    cl.sync();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        out[0] = (t1 - t0) / iters; out[1] = (t3 - t2) / iters; out[2] = (t4 - t3) / iters; out[3] = (t5 - t4) / iters;
    }
    *sink = idx;
}

// store by rank 0, barrier, load by the others: time from before the store to after the load (one hop of shared state)
__global__ void k_hop(unsigned *g, int iters, int mode, i64 *out) {
    cg::cluster_group cl = cg::this_cluster();
    __shared__ unsigned sm[32];
    unsigned r = cl.block_rank();
    volatile unsigned *m0 = cl.map_shared_rank(sm, 0);
    unsigned acc = 0;
    cl.sync();
    i64 t0 = clock64();
    for (int i = 0; i < iters; i++) {
        if (r == 0 && threadIdx.x == 0) { if (mode == 0) *(volatile unsigned *)g = i; else sm[0] = i; }
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        acc += mode == 0 ? *(volatile unsigned *)g : m0[0];
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    i64 t1 = clock64();
    if (blockIdx.x == 1 && threadIdx.x == 0) { out[0] = (t1 - t0) / iters; out[1] = acc; }
}

int main() {
    unsigned *a, *sink; i64 *out;
    const int N = 1 << 22;   // 16 MiB: L2 resident
    cudaMalloc(&a, (size_t)N * 4 * 64); cudaMalloc(&sink, 64); cudaMalloc(&out, 64);
    unsigned *h = (unsigned *)malloc((size_t)N * 4);
    for (int i = 0; i < N; i++) h[i] = (unsigned)(((long long)i * 2654435761LL + 12345) % N);
    cudaMemcpy(a, h, (size_t)N * 4, cudaMemcpyHostToDevice);
    i64 ho[4];
    const char *nm[4] = {"ld.volatile", "ld.cg", "ld.nc(__ldg)", "atomicAdd0"};
    for (int rep = 0; rep < 2; rep++)
        for (int mode = 0; mode < 4; mode++) {
            k_chase<<<1, 32>>>(a, 2000, mode, out, sink);
            cudaMemcpy(ho, out, 8, cudaMemcpyDeviceToHost);
            if (rep) printf("pointer chase, 16 MiB (L2), 1 warp: %-14s %lld cycles/load\n", nm[mode], ho[0]);
        }
    // DRAM: 1 GiB footprint
    {
        const size_t M = (size_t)N * 64;
        unsigned *hb = (unsigned *)malloc(M * 4);
        for (size_t i = 0; i < M; i++) hb[i] = (unsigned)((i * 2654435761ULL + 12345) % M);
        cudaMemcpy(a, hb, M * 4, cudaMemcpyHostToDevice);
        k_chase<<<1, 32>>>(a, 2000, 1, out, sink);
        cudaMemcpy(ho, out, 8, cudaMemcpyDeviceToHost);
        printf("pointer chase, 1 GiB (DRAM), 1 warp: ld.cg %lld cycles/load\n", ho[0]);
        free(hb);
    }
    cudaFuncSetAttribute(k_cluster, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaFuncSetAttribute(k_hop, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cl = 1; cl <= 16; cl *= 2)
        for (int T = 128; T <= 512; T *= 2) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cl); cfg.blockDim = dim3(T);
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaError_t e = cudaLaunchKernelEx(&cfg, k_cluster, 2000, out, sink);
            cudaDeviceSynchronize();
            cudaMemcpy(ho, out, 32, cudaMemcpyDeviceToHost);
            printf("cluster %2d x %3d threads: barrier.cluster %lld  remote-smem load %lld  local-smem(generic volatile) %lld  __syncthreads %lld  (%s)\n",
                   cl, T, ho[0], ho[1], ho[2], ho[3], cudaGetErrorString(e));
            if (cl > 1)
                for (int mode = 0; mode < 2; mode++) {
                    e = cudaLaunchKernelEx(&cfg, k_hop, a, 2000, mode, out);
                    cudaDeviceSynchronize();
                    cudaMemcpy(ho, out, 16, cudaMemcpyDeviceToHost);
                    printf("      hop (store, barrier, load, barrier) via %s: %lld cycles\n", mode ? "rank-0 smem" : "global", ho[0]);
                }
        }
    return 0;
}
