// Insert-pattern microbenchmark: ldcg probe + CAS(EMPTY -> key|idx) into per-frame tables.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)
__device__ __forceinline__ unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// n ops; op i targets frame (i / per_frame), cell = hash(i) % cells_per_frame; table per frame = slots entries
template <int MODE>
__global__ void k_ins(unsigned long long* tbl, int slots_lg, int per_frame, int cells, int n, unsigned seed, int* sink)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int f = i / per_frame;
    const unsigned cell = hash32(i * 2654435761u + seed) % cells;
    unsigned long long* t = tbl + ((size_t)f << slots_lg);
    unsigned h = (cell * 0x9E3779B1u) >> (32 - slots_lg);
    const unsigned mask = (1u << slots_lg) - 1;
    const unsigned long long mine = ((unsigned long long)cell << 32) | (unsigned)(i - f * per_frame);
    int won = 0;
    if (MODE == 0) {            // load then CAS/min (product protocol)
        unsigned long long e = __ldcg(t + h);
        while (true) {
            if ((unsigned)(e >> 32) == cell) { if ((unsigned)e > (unsigned)mine) { won = (unsigned)atomicMin(t + h, mine) > (unsigned)mine; } break; }
            if (e == ~0ull) { unsigned long long o = atomicCAS(t + h, ~0ull, mine); if (o == ~0ull) { won = 1; break; } e = o; continue; }
            h = (h + 1) & mask; e = __ldcg(t + h);
        }
    } else if (MODE == 1) {     // CAS first
        while (true) {
            unsigned long long o = atomicCAS(t + h, ~0ull, mine);
            if (o == ~0ull) { won = 1; break; }
            if ((unsigned)(o >> 32) == cell) { if ((unsigned)o > (unsigned)mine) won = (unsigned)atomicMin(t + h, mine) > (unsigned)mine; break; }
            h = (h + 1) & mask;
        }
    } else if (MODE == 2) {     // load only
        unsigned long long e = __ldcg(t + h); won = (e == 12345ull);
    }
    if (won && i == 0x7fffffff) *sink = 1;
}
__global__ void k_flush(float4* p, size_t n) { for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = make_float4(1, 2, 3, 4); }

int main()
{
    int* sink; CK(cudaMalloc(&sink, 4));
    const int per_frame = 60000, cells = 42000, slots_lg = 17;
    float4* fl; const size_t fln = 32 << 20; CK(cudaMalloc(&fl, fln * 16));   // 512 MB flush buffer
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int frames : {4, 16, 64}) {
        const int n = frames * per_frame;
        unsigned long long* tbl; CK(cudaMalloc(&tbl, ((size_t)frames << slots_lg) * 8));
        for (int cold = 0; cold < 2; ++cold) for (int mode = 0; mode < 3; ++mode) {
            float best = 1e9, sum = 0;
            for (int r = 0; r < 6; ++r) {
                CK(cudaMemset(tbl, 0xFF, ((size_t)frames << slots_lg) * 8));
                if (cold) k_flush<<<148 * 8, 256>>>(fl, fln);
                cudaEventRecord(a);
                if (mode == 0) k_ins<0><<<(n + 255) / 256, 256>>>(tbl, slots_lg, per_frame, cells, n, r, sink);
                if (mode == 1) k_ins<1><<<(n + 255) / 256, 256>>>(tbl, slots_lg, per_frame, cells, n, r, sink);
                if (mode == 2) k_ins<2><<<(n + 255) / 256, 256>>>(tbl, slots_lg, per_frame, cells, n, r, sink);
                cudaEventRecord(b); CK(cudaEventSynchronize(b));
                float ms; cudaEventElapsedTime(&ms, a, b);
                if (r > 0) { best = ms < best ? ms : best; sum += ms; }
            }
            printf("frames %2d n=%7d %s mode %d (%s): best %.1f us avg %.1f us -> %.1f Gop/s\n", frames, n, cold ? "cold" : "warm",
                   mode, mode == 0 ? "ld+cas" : mode == 1 ? "cas-first" : "ld only", best * 1e3, sum / 5 * 1e3, n / best / 1e6);
        }
        cudaFree(tbl);
    }
    return 0;
}
