// Cost of sparse-lane 64-bit CAS atomics on an L2-resident table: ops/s as a function of the
// fraction of active lanes per warp instruction.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o microbench3 microbench3.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int MODE>
__global__ void k_sparse(unsigned long long* tbl, unsigned mask, int n, unsigned seed, unsigned thr, unsigned long long* sink)
{
    unsigned long long acc = 0;
    const int base = (blockIdx.x * blockDim.x + threadIdx.x);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int i = base + k * gridDim.x * blockDim.x;
        const unsigned h = hash32(i * 2654435761u + seed);
        const bool on = (hash32(h ^ 0x5bd1e995u) & 1023u) < thr;
        if (i < n && on) {
            if (MODE == 0) acc += atomicCAS(tbl + (h & mask), 0xFFFFFFFFFFFFFFFFull, ((unsigned long long)h << 32) | i);
            if (MODE == 1) acc += atomicMin((unsigned*)tbl + (h & mask), (unsigned)i);
            if (MODE == 2) acc += __ldcg(tbl + (h & mask));
        }
    }
    if (acc == 0x123456789ull) *sink = acc;
}

template <int MODE>
void run(unsigned long long* tbl, unsigned mask, unsigned long long* sink, const char* name)
{
    const int n = 32 << 20;
    const int threads = 256, blocks = n / 4 / threads;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    printf("%-14s", name);
    for (unsigned thr : {1024u, 768u, 512u, 368u, 256u, 128u, 64u, 32u}) {
        CK(cudaMemset(tbl, 0xFF, ((size_t)mask + 1) * 8));
        k_sparse<MODE><<<blocks, threads>>>(tbl, mask, n, 1, thr, sink);
        cudaEventRecord(a);
        for (int r = 0; r < 5; ++r) k_sparse<MODE><<<blocks, threads>>>(tbl, mask, n, r * 7919 + 3, thr, sink);
        cudaEventRecord(b); CK(cudaEventSynchronize(b));
        float ms; cudaEventElapsedTime(&ms, a, b); ms /= 5;
        const double ops = (double)n * thr / 1024.0, instr = (double)n / 32;
        printf(" | act %4.1f%%: %6.1f Gop/s %5.1f cyc/instr/SM", thr / 10.24, ops / ms / 1e6, ms * 1e-3 * 1.9e9 * 148 / instr);
    }
    printf("\n");
}

int main()
{
    unsigned long long* sink; CK(cudaMalloc(&sink, 8));
    const size_t slots = 1ull << 23;   // 64 MB
    unsigned long long* tbl; CK(cudaMalloc(&tbl, slots * 8));
    run<0>(tbl, slots - 1, sink, "atom.cas.u64");
    run<1>(tbl, slots - 1, sink, "atom.min.u32");
    run<2>(tbl, slots - 1, sink, "ld.cg 8B");
    return 0;
}
