// Microbenchmarks that size the design: random L2/DRAM loads and atomics on B200.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o microbench microbench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

template <int MODE, int ILP>
__global__ void k_rand(unsigned long long* tbl, unsigned mask, int n, unsigned seed, unsigned long long* sink)
{
    unsigned long long acc = 0;
    const int base = (blockIdx.x * blockDim.x + threadIdx.x) * ILP;
    unsigned long long v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) {
        const int i = base + k;
        const unsigned h = hash32(i * 2654435761u + seed) & mask;
        if (i < n) {
            if (MODE == 0) v[k] = __ldcg(tbl + h);                                        // random 8B load
            if (MODE == 1) v[k] = atomicMin(tbl + h, ((unsigned long long)h << 32) | i);  // atom.min.u64 w/ return
            if (MODE == 2) v[k] = atomicCAS(tbl + h, 0xFFFFFFFFFFFFFFFFull, ((unsigned long long)h << 32) | i);
            if (MODE == 3) { atomicMin(tbl + h, ((unsigned long long)h << 32) | i); v[k] = 0; }   // red (no return)
            if (MODE == 4) v[k] = atomicMin((unsigned*)tbl + h, (unsigned)i);             // 32-bit atom w/ return
            if (MODE == 5) { atomicMin((unsigned*)tbl + h, (unsigned)i); v[k] = 0; }      // 32-bit red
            if (MODE == 6) { tbl[h] = i; v[k] = 0; }                                       // random 8B store
            if (MODE == 7) {                                                               // load then dependent atomic
                unsigned long long e = __ldcg(tbl + h);
                v[k] = (unsigned)e > (unsigned)i ? atomicMin(tbl + h, ((unsigned long long)(e >> 32) << 32) | i) : e;
            }
        } else v[k] = 0;
    }
#pragma unroll
    for (int k = 0; k < ILP; ++k) acc += v[k];
    if (acc == 0x123456789ull) *sink = acc;
}

__global__ void k_stream(const float4* __restrict__ in, int n, float* sink)
{
    float a = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 v = __ldg(in + i);
        a += v.x + v.y + v.z + v.w;
    }
    if (a == 1.2345f) *sink = a;
}

template <int MODE, int ILP>
float run(unsigned long long* tbl, unsigned mask, int n, unsigned long long* sink, int reps)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int threads = 256, blocks = (n / ILP + threads - 1) / threads;
    k_rand<MODE, ILP><<<blocks, threads>>>(tbl, mask, n, 1, sink);
    cudaEventRecord(a);
    for (int r = 0; r < reps; ++r) k_rand<MODE, ILP><<<blocks, threads>>>(tbl, mask, n, r * 7919 + 3, sink);
    cudaEventRecord(b);
    CK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main()
{
    const int n = 8 << 20;  // 8M operations per launch
    unsigned long long* sink; CK(cudaMalloc(&sink, 8));
    const char* names[] = {"ld.cg 8B", "atom.min.u64", "atom.cas.u64", "red.min.u64", "atom.min.u32", "red.min.u32", "st 8B", "ld+dep atom.min.u64"};
    for (int lg = 17; lg <= 28; lg += (lg < 24 ? 3 : 2)) {   // 1 MB .. 2 GB tables of 8B entries
        const size_t slots = 1ull << lg;
        unsigned long long* tbl; CK(cudaMalloc(&tbl, slots * 8));
        CK(cudaMemset(tbl, 0xFF, slots * 8));
        printf("table %7.1f MB:", slots * 8 / 1048576.0);
        float t;
        t = run<0, 1>(tbl, slots - 1, n, sink, 5); printf("  %s %.1f Gop/s |", names[0], n / t / 1e6);
        t = run<0, 4>(tbl, slots - 1, n, sink, 5); printf(" ilp4 %.1f |", n / t / 1e6);
        t = run<6, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[6], n / t / 1e6);
        CK(cudaMemset(tbl, 0xFF, slots * 8));
        t = run<1, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[1], n / t / 1e6);
        t = run<1, 4>(tbl, slots - 1, n, sink, 5); printf(" ilp4 %.1f |", n / t / 1e6);
        t = run<3, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[3], n / t / 1e6);
        t = run<2, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[2], n / t / 1e6);
        t = run<4, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[4], n / t / 1e6);
        t = run<5, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[5], n / t / 1e6);
        t = run<7, 1>(tbl, slots - 1, n, sink, 5); printf(" %s %.1f |", names[7], n / t / 1e6);
        t = run<7, 4>(tbl, slots - 1, n, sink, 5); printf(" ilp4 %.1f\n", n / t / 1e6);
        cudaFree(tbl);
    }
    // streaming read reference
    const int m = 64 << 20;  // 1 GB of float4
    float4* buf; CK(cudaMalloc(&buf, (size_t)m * 16)); CK(cudaMemset(buf, 0, (size_t)m * 16));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k_stream<<<148 * 16, 256>>>(buf, m, (float*)sink);
    cudaEventRecord(a);
    for (int r = 0; r < 5; ++r) k_stream<<<148 * 16, 256>>>(buf, m, (float*)sink);
    cudaEventRecord(b); CK(cudaEventSynchronize(b));
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("stream read 1 GB: %.0f GB/s\n", (double)m * 16 * 5 / ms / 1e6);
    // launch overhead: empty kernels back to back
    cudaEventRecord(a);
    for (int r = 0; r < 1000; ++r) k_stream<<<1, 32>>>(buf, 0, (float*)sink);
    cudaEventRecord(b); CK(cudaEventSynchronize(b));
    cudaEventElapsedTime(&ms, a, b);
    printf("empty kernel launch cadence: %.2f us\n", ms);
    return 0;
}
