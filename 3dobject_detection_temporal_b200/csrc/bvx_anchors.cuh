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
// The histogram is sparse (<= max_voxels cells of a map of millions), so the dense
// map is never zero-filled, read back or scanned in global memory.  The map is cut
// into bands of H rows; the voxels are binned by (frame, band) with a counting sort,
// and each band is built, scanned both ways and written exactly once from shared
// memory.  The table is BANDED -- T(y, x) = band-local table + prefix of the bands
// above, kept as a second, small table -- so no pass depends on another band:
//   S(y, x) = band_table[y][x] + above[y / H][x]
//
//   k_aa_count / k_aa_offsets / k_aa_bin   counting sort of the voxels by (frame, band)
//   k_aa_band     one block per (band, frame): histogram in shared memory, prefix
//                 along x, prefix along y, one coalesced write of the band
//   k_aa_above    exclusive prefix of the band totals over the bands of a frame
//   k_aa_query    one thread per (anchor, frame): four corners, two reads each
// DRAM traffic: one write of the table (2 bytes per map cell) + the corner reads.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvx {

constexpr int kAaMaxBandRows = 32;
constexpr int kAaYShift = 27;            // binned voxel = (row inside the band << 27) | x

struct AnchorsMap {
    unsigned int* words;   // banded table [batch][ny][wn]
    unsigned int* above;   // [batch][nb][wn]: totals of the bands above (band totals before k_aa_above)
    int* cnt;              // [batch * nb] voxels per (frame, band)
    int* fill;             // [batch * nb] bin cursors
    int* ticket;           // [1] blocks of k_aa_offsets finished (zeroed with cnt and fill)
    int* off;              // [batch * nb] exclusive prefix of cnt inside the frame
    int* frame_base;       // [batch + 1] exclusive prefix of the frames' voxel counts
    unsigned int* binned;  // [m] voxels grouped by (frame, band)
    int batch, ny, nx;
    int wn;                // words per row: nx / 2 (packed) or nx
    int packed;            // 1: two 16-bit cells per word
    int h;                 // band height (power of two), hshift = log2(h)
    int hshift;
    int nb;                // bands per frame
};

__device__ __forceinline__ unsigned int aa_word_cell(unsigned int w, int x, int packed)
{
    return packed ? (w >> ((x & 1) * 16)) & 0xFFFFu : w;
}

__device__ __forceinline__ unsigned int aa_cell(const AnchorsMap& mp, int b, int y, int x)
{
    const int wx = mp.packed ? x >> 1 : x;
    const unsigned int t = __ldg(mp.words + ((size_t)b * mp.ny + y) * mp.wn + wx);
    const unsigned int a = __ldg(mp.above + ((size_t)b * mp.nb + (y >> mp.hshift)) * mp.wn + wx);
    return aa_word_cell(t, x, mp.packed) + aa_word_cell(a, x, mp.packed);
}

// coors rows are (z, y, x) [cols == 3, single frame] or (b, z, y, x) [cols == 4].
// Rows outside the map or the batch are ignored (like the BEV scatter).
__device__ __forceinline__ bool aa_voxel(const int* __restrict__ coors, int i, int cols, const AnchorsMap& mp,
                                         int& b, int& y, int& x)
{
    b = 0;
    if (cols == 4) {
        const int4 q = __ldg(reinterpret_cast<const int4*>(coors) + i);
        b = q.x; y = q.z; x = q.w;
    } else {
        y = __ldg(coors + 3 * (size_t)i + 1);
        x = __ldg(coors + 3 * (size_t)i + 2);
    }
    return (unsigned)b < (unsigned)mp.batch && (unsigned)y < (unsigned)mp.ny && (unsigned)x < (unsigned)mp.nx;
}

// Counting sort, both passes: a block takes 1024 consecutive voxels (almost always of one
// frame) and bins them in shared memory first, so the global counters see one atomic per
// (block, non-empty band) instead of one per voxel.  Voxels of another frame than the block's
// first one (a block straddling a frame boundary), or maps with more than 1024 bands, go to
// the global counters directly.
constexpr int kAaSortPts = 4;
constexpr int kAaSmemBins = 1024;

__global__ void __launch_bounds__(256) k_aa_count(const int* __restrict__ coors, int m, int cols,
                                                  const AnchorsMap mp)
{
    __shared__ int s_bin[kAaSmemBins];
    const int i0 = blockIdx.x * (256 * kAaSortPts);
    const bool staged = mp.nb <= kAaSmemBins;
    const int f0 = cols == 4 ? __ldg(coors + 4 * (size_t)i0) : 0;
    if (staged) {
        for (int k = threadIdx.x; k < mp.nb; k += 256) s_bin[k] = 0;
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kAaSortPts; ++k) {
        const int i = i0 + k * 256 + threadIdx.x;
        int b, y, x;
        if (i >= m || !aa_voxel(coors, i, cols, mp, b, y, x)) continue;
        const int band = y >> mp.hshift;
        if (staged && b == f0) atomicAdd(s_bin + band, 1);
        else atomicAdd(mp.cnt + b * mp.nb + band, 1);
    }
    if (!staged || (unsigned)f0 >= (unsigned)mp.batch) return;
    __syncthreads();
    for (int k = threadIdx.x; k < mp.nb; k += 256)
        if (s_bin[k]) atomicAdd(mp.cnt + f0 * mp.nb + k, s_bin[k]);
}

// grid (batch), 256 threads: off = exclusive prefix of cnt inside each frame; the last block
// to finish turns the frame totals into frame_base.
__global__ void __launch_bounds__(256) k_aa_offsets(const AnchorsMap mp)
{
    __shared__ int sm[kWarps + 1];
    __shared__ int s_last;
    const int f = blockIdx.x;
    int carry = 0;
    for (int k0 = 0; k0 < mp.nb; k0 += 256) {
        const int k = k0 + threadIdx.x;
        const int v = k < mp.nb ? mp.cnt[f * mp.nb + k] : 0;
        int tot;
        const int ex = block_exclusive_scan(v, &tot, sm);
        if (k < mp.nb) mp.off[f * mp.nb + k] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) {
        mp.frame_base[f + 1] = carry;          // frame total for now
        __threadfence();
        s_last = (atomicAdd(mp.ticket, 1) == mp.batch - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    carry = 0;
    for (int f0 = 0; f0 < mp.batch; f0 += 256) {
        const int ff = f0 + threadIdx.x;
        const int v = ff < mp.batch ? __ldcg(mp.frame_base + ff + 1) : 0;
        int tot;
        const int ex = block_exclusive_scan(v, &tot, sm);
        __syncthreads();
        if (ff < mp.batch) mp.frame_base[ff + 1] = carry + ex + v;
        carry += tot;
    }
    if (threadIdx.x == 0) mp.frame_base[0] = 0;
}

__global__ void __launch_bounds__(256) k_aa_bin(const int* __restrict__ coors, int m, int cols,
                                                const AnchorsMap mp)
{
    __shared__ int s_bin[kAaSmemBins];
    const int i0 = blockIdx.x * (256 * kAaSortPts);
    const bool staged = mp.nb <= kAaSmemBins;
    const int f0 = cols == 4 ? __ldg(coors + 4 * (size_t)i0) : 0;
    if (staged) {
        for (int k = threadIdx.x; k < mp.nb; k += 256) s_bin[k] = 0;
        __syncthreads();
    }
    int key[kAaSortPts], rank[kAaSortPts];
    unsigned int val[kAaSortPts];
#pragma unroll
    for (int k = 0; k < kAaSortPts; ++k) {
        const int i = i0 + k * 256 + threadIdx.x;
        int b, y, x;
        key[k] = -1;
        rank[k] = -1;                                   // -1: global path
        if (i >= m || !aa_voxel(coors, i, cols, mp, b, y, x)) continue;
        const int band = y >> mp.hshift;
        key[k] = b * mp.nb + band;
        val[k] = ((unsigned)(y & (mp.h - 1)) << kAaYShift) | (unsigned)x;
        if (staged && b == f0) rank[k] = atomicAdd(s_bin + band, 1);
    }
    if (staged) {
        __syncthreads();
        if ((unsigned)f0 < (unsigned)mp.batch)
            for (int k = threadIdx.x; k < mp.nb; k += 256) {
                const int c = s_bin[k];
                if (c) s_bin[k] = atomicAdd(mp.fill + f0 * mp.nb + k, c);   // the block's base in the bin
            }
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kAaSortPts; ++k) {
        if (key[k] < 0) continue;
        const int f = key[k] / mp.nb;
        const int in_bin = rank[k] >= 0 ? s_bin[key[k] - f * mp.nb] + rank[k] : atomicAdd(mp.fill + key[k], 1);
        mp.binned[mp.frame_base[f] + mp.off[key[k]] + in_bin] = val[k];
    }
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

// grid (nb, batch), 256 threads, dynamic smem = h * wn words.
// V = 4: wn % 4 == 0 (16-byte row stores).
template <int V>
__global__ void __launch_bounds__(256) k_aa_band(const AnchorsMap mp)
{
    extern __shared__ __align__(16) unsigned int aa_s[];
    const int band = blockIdx.x, f = blockIdx.y;
    const int y0 = band << mp.hshift;
    const int rows = min(mp.h, mp.ny - y0);
    const int wn = mp.wn;
    const int key = f * mp.nb + band;
    const int j0 = mp.frame_base[f] + mp.off[key], j1 = j0 + mp.cnt[key];
    unsigned int* g = mp.words + ((size_t)f * mp.ny + y0) * wn;
    unsigned int* tot = mp.above + (size_t)key * wn;
    if (j0 == j1) {                      // empty band: zeros
        for (int w = threadIdx.x * V; w < wn; w += 256 * V) {
            for (int r = 0; r < rows; ++r) {
                if (V == 4) *reinterpret_cast<uint4*>(g + (size_t)r * wn + w) = make_uint4(0u, 0u, 0u, 0u);
                else g[(size_t)r * wn + w] = 0u;
            }
            if (V == 4) *reinterpret_cast<uint4*>(tot + w) = make_uint4(0u, 0u, 0u, 0u);
            else tot[w] = 0u;
        }
        return;
    }
    for (int k = threadIdx.x; k < rows * wn; k += 256) aa_s[k] = 0u;   // (wn odd: rows are not 16-byte aligned)
    __syncthreads();
    for (int j = j0 + threadIdx.x; j < j1; j += 256) {
        const unsigned int e = __ldg(mp.binned + j);
        const int yl = e >> kAaYShift, x = e & ((1u << kAaYShift) - 1u);
        if (mp.packed) atomicAdd(aa_s + yl * wn + (x >> 1), 1u << ((x & 1) * 16));
        else atomicAdd(aa_s + yl * wn + x, 1u);
    }
    __syncthreads();
    // prefix along x: one warp per row, each lane a contiguous run of words (odd length: no
    // bank conflicts), one warp scan of the lane totals per row
    const int lane = threadIdx.x & 31;
    const int run_len = ((wn + 31) / 32) | 1;
    for (int r = threadIdx.x >> 5; r < rows; r += 8) {
        unsigned int* row = aa_s + r * wn;
        const int w0 = min(wn, lane * run_len), w1 = min(wn, w0 + run_len);
        unsigned int t = 0;
        for (int w = w0; w < w1; ++w) t += aa_word_total(row[w], mp.packed);
        unsigned int inc = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int u = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += u;
        }
        unsigned int run = inc - t;
        for (int w = w0; w < w1; ++w) row[w] = aa_word_prefix(row[w], run, mp.packed);
    }
    __syncthreads();
    // prefix along y inside the band, written straight to the table (coalesced rows)
    for (int w = threadIdx.x * V; w < wn; w += 256 * V) {
        unsigned int acc[V];
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = 0u;
        for (int r = 0; r < rows; ++r) {
            if (V == 4) {
                const uint4 q = *reinterpret_cast<const uint4*>(aa_s + r * wn + w);
                acc[0] += q.x; acc[1 % V] += q.y; acc[2 % V] += q.z; acc[3 % V] += q.w;
                *reinterpret_cast<uint4*>(g + (size_t)r * wn + w) = make_uint4(acc[0], acc[1 % V], acc[2 % V], acc[3 % V]);
            } else {
                acc[0] += aa_s[r * wn + w];
                g[(size_t)r * wn + w] = acc[0];
            }
        }
        if (V == 4) *reinterpret_cast<uint4*>(tot + w) = make_uint4(acc[0], acc[1 % V], acc[2 % V], acc[3 % V]);
        else tot[w] = acc[0];
    }
}

// above[f][band][w] <- sum of the totals of bands 0 .. band-1 (in place), 16 bands' loads in
// flight per thread.
__global__ void __launch_bounds__(256) k_aa_above(const AnchorsMap mp)
{
    const int w = blockIdx.x * 256 + threadIdx.x;
    if (w >= mp.wn) return;
    unsigned int* a = mp.above + (size_t)blockIdx.y * mp.nb * mp.wn + w;
    unsigned int run = 0;
    for (int b0 = 0; b0 < mp.nb; b0 += 16) {
        unsigned int v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = b0 + k < mp.nb ? __ldcg(a + (size_t)(b0 + k) * mp.wn) : 0u;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (b0 + k < mp.nb) a[(size_t)(b0 + k) * mp.wn] = run;
            run += v[k];
        }
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
