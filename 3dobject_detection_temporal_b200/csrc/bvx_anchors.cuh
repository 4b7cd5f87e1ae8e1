// bvx_anchors.cuh -- anchors-area from voxel coordinates.
//
// Reference (second/core/box_np_ops.py:776-781, :784-810; used as in
// second/kittiviewer/viewer.py:395-401):
//   dense = sparse_sum_for_anchors_mask(coors, (ny, nx))   2-D histogram of (y, x)
//   dense = dense.cumsum(0).cumsum(1)                      summed-area table
//   area  = fused_get_anchors_area(dense, anchors_bv, voxel_size, range, grid)
//   mask  = area > anchor_area_threshold
//
// Every value is a voxel count (< 2^24), so the reference's float32 sums are exact
// and the table can be held in integers: 16-bit cells packed two per 32-bit word
// when a frame holds at most 65535 voxels and nx is even (the common case), 32-bit
// cells otherwise.  Both layouts are arrays of 32-bit words, and a word-wise add is
// a cell-wise add for both (a packed half never carries: no table entry exceeds the
// frame's voxel count).
//
//   k_aa_hist     one thread per voxel: atomicAdd into the zeroed word map
//   k_aa_rowscan  one warp per row: inclusive prefix along x, 4 words per lane
//   k_aa_colscan  one block per (16-word column strip, frame): the whole strip sits
//                 in shared memory, 32 row segments scanned in parallel, one DRAM
//                 read + one DRAM write per word
//   k_aa_query    one thread per (anchor, frame): four corner reads
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvx {

constexpr int kAaStripWords = 16;   // column strip width of k_aa_colscan (64 bytes per row)
constexpr int kAaColThreads = 512;  // 16 lanes x 32 row segments
constexpr int kAaSegs = kAaColThreads / kAaStripWords;

struct AnchorsMap {
    unsigned int* words;  // [batch][ny][wn]
    int batch, ny, nx;
    int wn;               // words per row: nx / 2 (packed) or nx
    int packed;           // 1: two 16-bit cells per word
};

__device__ __forceinline__ unsigned int aa_cell(const AnchorsMap& mp, int b, int y, int x)
{
    const size_t row = ((size_t)b * mp.ny + y) * mp.wn;
    if (mp.packed) return (__ldg(mp.words + row + (x >> 1)) >> ((x & 1) * 16)) & 0xFFFFu;
    return __ldg(mp.words + row + x);
}

// coors rows are (z, y, x) [cols == 3, single frame] or (b, z, y, x) [cols == 4].
__global__ void __launch_bounds__(256) k_aa_hist(const int* __restrict__ coors, int m, int cols,
                                                 const AnchorsMap mp)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= m) return;
    int b = 0, y, x;
    if (cols == 4) {
        const int4 q = __ldg(reinterpret_cast<const int4*>(coors) + i);
        b = q.x; y = q.z; x = q.w;
    } else {
        y = __ldg(coors + 3 * (size_t)i + 1);
        x = __ldg(coors + 3 * (size_t)i + 2);
    }
    if ((unsigned)b >= (unsigned)mp.batch || (unsigned)y >= (unsigned)mp.ny || (unsigned)x >= (unsigned)mp.nx)
        return;
    const size_t row = ((size_t)b * mp.ny + y) * mp.wn;
    if (mp.packed) atomicAdd(mp.words + row + (x >> 1), 1u << ((x & 1) * 16));
    else atomicAdd(mp.words + row + x, 1u);
}

// inclusive prefix of one word given the running sum of everything to its left
__device__ __forceinline__ unsigned int aa_word_prefix(unsigned int w, unsigned int& run, int packed)
{
    if (packed) {
        const unsigned int a = run + (w & 0xFFFFu);
        run = a + (w >> 16);
        return a | (run << 16);
    }
    run += w;
    return run;
}

__device__ __forceinline__ unsigned int aa_word_total(unsigned int w, int packed)
{
    return packed ? (w & 0xFFFFu) + (w >> 16) : w;
}

// One warp per row.  V = 4: rows are 16-byte aligned and wn % 4 == 0 (uint4 accesses).
template <int V>
__global__ void __launch_bounds__(256) k_aa_rowscan(const AnchorsMap mp)
{
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= (long long)mp.batch * mp.ny) return;
    unsigned int* r = mp.words + (size_t)row * mp.wn;
    unsigned int carry = 0;
    for (int w0 = 0; w0 < mp.wn; w0 += 32 * V) {
        const int w = w0 + lane * V;
        unsigned int v[V];
        if (V == 4) {
            uint4 q = make_uint4(0u, 0u, 0u, 0u);
            if (w < mp.wn) q = *reinterpret_cast<const uint4*>(r + w);
            v[0] = q.x; v[1 % V] = q.y; v[2 % V] = q.z; v[3 % V] = q.w;
        } else {
#pragma unroll
            for (int k = 0; k < V; ++k) v[k] = (w + k < mp.wn) ? r[w + k] : 0u;
        }
        unsigned int tot = 0;
#pragma unroll
        for (int k = 0; k < V; ++k) tot += aa_word_total(v[k], mp.packed);
        unsigned int inc = tot;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        unsigned int run = carry + inc - tot;
        carry += __shfl_sync(0xffffffffu, inc, 31);
#pragma unroll
        for (int k = 0; k < V; ++k) v[k] = aa_word_prefix(v[k], run, mp.packed);
        if (V == 4) {
            if (w < mp.wn) *reinterpret_cast<uint4*>(r + w) = make_uint4(v[0], v[1 % V], v[2 % V], v[3 % V]);
        } else {
#pragma unroll
            for (int k = 0; k < V; ++k)
                if (w + k < mp.wn) r[w + k] = v[k];
        }
    }
}

// grid (strips, batch), 512 threads, dynamic smem = ny * 16 words.  Thread t owns
// column t % 16 of row segment t / 16.
__global__ void __launch_bounds__(kAaColThreads) k_aa_colscan(const AnchorsMap mp)
{
    extern __shared__ unsigned int aa_s[];
    __shared__ unsigned int s_tot[kAaSegs][kAaStripWords], s_off[kAaSegs][kAaStripWords];
    const int c0 = blockIdx.x * kAaStripWords;
    const int col = threadIdx.x % kAaStripWords, seg = threadIdx.x / kAaStripWords;
    const bool live = c0 + col < mp.wn;
    unsigned int* g = mp.words + (size_t)blockIdx.y * mp.ny * mp.wn + c0 + col;
    // load: thread (col, seg) reads rows seg, seg + 32, ... (two rows per warp instruction)
    for (int y = seg; y < mp.ny; y += kAaSegs)
        aa_s[y * kAaStripWords + col] = live ? __ldcs(g + (size_t)y * mp.wn) : 0u;
    __syncthreads();
    const int per = (mp.ny + kAaSegs - 1) / kAaSegs;
    const int y0 = seg * per, y1 = min(mp.ny, y0 + per);
    unsigned int acc = 0;
    for (int y = y0; y < y1; ++y) {
        acc += aa_s[y * kAaStripWords + col];
        aa_s[y * kAaStripWords + col] = acc;
    }
    s_tot[seg][col] = acc;
    __syncthreads();
    unsigned int above = 0;
    for (int s = 0; s < seg; ++s) above += s_tot[s][col];
    s_off[seg][col] = above;
    __syncthreads();
    // store: row y gets the total of the segments above its own
    for (int y = seg; y < mp.ny; y += kAaSegs)
        if (live) g[(size_t)y * mp.wn] = aa_s[y * kAaStripWords + col] + s_off[y / per][col];
}

// Any ny (strip does not fit in shared memory): one thread per word column.
__global__ void __launch_bounds__(256) k_aa_colscan_serial(const AnchorsMap mp)
{
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= mp.wn) return;
    unsigned int* g = mp.words + (size_t)blockIdx.y * mp.ny * mp.wn + c;
    unsigned int acc = 0;
    for (int y = 0; y < mp.ny; ++y) {
        acc += g[(size_t)y * mp.wn];
        g[(size_t)y * mp.wn] = acc;
    }
}

struct AnchorsQuery {
    float stride[2];   // voxel_size x, y
    float offset[2];   // point_cloud_range x0, y0
    int gx, gy;        // grid_size x, y
};

// anchors_bv rows are (x0, y0, x1, y1); out[b][i] = ID - IB - IC + IA.
// The reference clamps only (x0, y0) from below and (x1, y1) from above; an anchor
// wholly outside the range indexes out of bounds there.  Here all four are clamped
// into the map.
__global__ void __launch_bounds__(256) k_aa_query(const AnchorsMap mp, const float* __restrict__ anchors_bv,
                                                  int n, const AnchorsQuery q, float* __restrict__ out)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float4 a = __ldg(reinterpret_cast<const float4*>(anchors_bv) + i);
    const float av[4] = {a.x, a.y, a.z, a.w};
    int c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float f = floorf(__fdiv_rn(__fsub_rn(av[k], q.offset[k & 1]), q.stride[k & 1]));
        const int hi = ((k & 1) ? q.gy : q.gx) - 1;
        c[k] = f < 0.0f ? 0 : (f > (float)hi ? hi : (int)f);   // NaN -> 0
    }
    const int b = blockIdx.y;
    const int ID = (int)aa_cell(mp, b, c[3], c[2]);
    const int IA = (int)aa_cell(mp, b, c[1], c[0]);
    const int IB = (int)aa_cell(mp, b, c[3], c[0]);
    const int IC = (int)aa_cell(mp, b, c[1], c[2]);
    out[(size_t)b * n + i] = (float)(ID - IB - IC + IA);
}

// The reference's two functions under their own names (float maps, caller does the cumsums).
__global__ void __launch_bounds__(256) k_aa_hist_f32(const int* __restrict__ coors, int m, int cols, int yoff,
                                                     int ny, int nx, float* __restrict__ ret)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= m) return;
    const int y = __ldg(coors + (size_t)cols * i + yoff), x = __ldg(coors + (size_t)cols * i + yoff + 1);
    if ((unsigned)y >= (unsigned)ny || (unsigned)x >= (unsigned)nx) return;
    atomicAdd(ret + (size_t)y * nx + x, 1.0f);   // counts: exact in fp32, order-free
}

__global__ void __launch_bounds__(256) k_aa_query_f32(const float* __restrict__ dense, int ny, int nx,
                                                      const float* __restrict__ anchors_bv, int n,
                                                      const AnchorsQuery q, float* __restrict__ out)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float4 a = __ldg(reinterpret_cast<const float4*>(anchors_bv) + i);
    const float av[4] = {a.x, a.y, a.z, a.w};
    int c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float f = floorf(__fdiv_rn(__fsub_rn(av[k], q.offset[k & 1]), q.stride[k & 1]));
        const int hi = min(((k & 1) ? q.gy : q.gx), ((k & 1) ? ny : nx)) - 1;
        c[k] = f < 0.0f ? 0 : (f > (float)hi ? hi : (int)f);
    }
    const float ID = __ldg(dense + (size_t)c[3] * nx + c[2]);
    const float IA = __ldg(dense + (size_t)c[1] * nx + c[0]);
    const float IB = __ldg(dense + (size_t)c[3] * nx + c[0]);
    const float IC = __ldg(dense + (size_t)c[1] * nx + c[2]);
    out[i] = __fadd_rn(__fsub_rn(__fsub_rn(ID, IB), IC), IA);
}

}  // namespace bvx
