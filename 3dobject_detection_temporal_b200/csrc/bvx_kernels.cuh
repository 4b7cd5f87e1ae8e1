// bvx_kernels.cuh -- sm_100a kernels of the voxelization hot path.
//
// Pipeline (one pass of VoxelGenerator.generate over B frames; DESIGN.md has the
// data layout and the per-kernel roofline):
//   k_insert  per point : cell key -> table slot, first occurrence by 64-bit
//                         atomicMin/atomicCAS on {cell:32 | point index:32}
//   k_verify  per cand. : candidate bit -> true first-occurrence bit, word popcounts
//   k_rank    per frame : exclusive scan of the popcounts -> first-come voxel
//                         ranks, M = min(total, cap), break cut-off index i*
//   k_slot    per point : voxel id from the first-occurrence rank; deterministic
//                         slot in input order (sorted atomicMin chain, or
//                         count + bucket for large T)
//   k_overflow          : (large T only) T-smallest selection for crowded voxels
//   k_gather  per voxel : shared-memory staged row gather, zero padding,
//                         num_points, optional fused mean VFE; the extra blocks
//                         of the same launch reset the touched table entries
// Everything is fp32 / int32; there is no contraction anywhere, hence no tensor
// cores: the bound is HBM / L2 bandwidth.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvx {

constexpr unsigned long long kEmpty64 = 0xFFFFFFFFFFFFFFFFull;
constexpr int kEmptyIdx = 0x7FFFFFFF;
constexpr int kChainMaxT = 8;      // T <= 8: sorted atomicMin chain; else count+bucket
constexpr int kBlock = 256;
constexpr int kRankThreads = 1024;

struct GridParams {
    float lo[3];
    float vs[3];
    float gsf[3];  // grid size as float, x y z
    int gs[3];     // x y z
};

struct FrameDesc {
    unsigned long long tbl_off;  // entry offset of this frame's table region
    unsigned int tbl_mask;       // slots - 1 (hashed)
    int tbl_shift;               // 32 - log2(slots) (hashed); -1 => dense (slot = cell)
    int n;                       // points in the frame (all segments)
    int pt_off;                  // base into slot[]
    int word_off;                // base into cand/first/wordrank
    int nwords;
    int seg_begin;
    int seg_count;
};

struct SegDesc {
    const float* pts;
    int n;
    int start;          // first frame-local point index of the segment
    int has_transform;
    int has_dt;
    float R[9];
    float t[3];
    float dt;
};

struct FrameState {
    int total;   // distinct in-range cells
    int m;       // voxels kept = min(total, cap)
    int istar;   // break cut-off: points with index >= istar are dropped
    int pad;
};

struct Pipe {
    GridParams g;
    const FrameDesc* frames;
    const SegDesc* segs;
    FrameState* state;
    unsigned long long* table;
    int* slot;
    unsigned int* cand;
    unsigned int* first;
    int* wordrank;
    int* bucket;
    int* cnt;
    int2* ovf;
    int* counters;       // [0] frames-done ticket, [1] overflow count
    int* frame_voxel_offsets;
    int nframes;
    int f_in;
    int f_out;
    int T;
    int cap;             // per-frame max_voxels of this call
    int overflow_continue;
    int coor_cols;       // 3 (single frame) or 4 (batched, batch index first)
    int ovf_cap;
    int vec4;            // f_in == 4 and every segment pointer 16-byte aligned
    // outputs
    float* voxels;
    int* coors;
    int* num_points;
    float* mean_out;
    int mean_c;
};

__device__ __forceinline__ const SegDesc& find_seg(const Pipe& p, const FrameDesc& fr, int i)
{
    int s = fr.seg_begin;
    const int e = fr.seg_begin + fr.seg_count - 1;
    while (s < e && i >= p.segs[s].start + p.segs[s].n) ++s;
    return p.segs[s];
}

// x' = ((r00*x + r01*y) + r02*z) + tx  -- fixed association, no FMA (A.4)
__device__ __forceinline__ void xform(const SegDesc& sg, float& x, float& y, float& z)
{
    if (sg.has_transform) {
        const float x0 = x, y0 = y, z0 = z;
        x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(sg.R[0], x0), __fmul_rn(sg.R[1], y0)), __fmul_rn(sg.R[2], z0)), sg.t[0]);
        y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(sg.R[3], x0), __fmul_rn(sg.R[4], y0)), __fmul_rn(sg.R[5], z0)), sg.t[1]);
        z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(sg.R[6], x0), __fmul_rn(sg.R[7], y0)), __fmul_rn(sg.R[8], z0)), sg.t[2]);
    }
}

// floor((p - lo) / size) in fp32 with IEEE subtract/divide; -1 if outside / NaN.
// Linear cell index in the reference's zyx grid order.
__device__ __forceinline__ int cell_of(const GridParams& g, float x, float y, float z)
{
    const float cx = floorf(__fdiv_rn(__fsub_rn(x, g.lo[0]), g.vs[0]));
    const float cy = floorf(__fdiv_rn(__fsub_rn(y, g.lo[1]), g.vs[1]));
    const float cz = floorf(__fdiv_rn(__fsub_rn(z, g.lo[2]), g.vs[2]));
    const bool ok = (cx >= 0.0f) && (cx < g.gsf[0]) && (cy >= 0.0f) && (cy < g.gsf[1]) &&
                    (cz >= 0.0f) && (cz < g.gsf[2]);
    if (!ok) return -1;
    return ((int)cz * g.gs[1] + (int)cy) * g.gs[0] + (int)cx;
}

__device__ __forceinline__ unsigned int hash_cell(unsigned int cell, int shift)
{
    return (cell * 0x9E3779B1u) >> shift;
}

// ---------------------------------------------------------------- k_insert
// grid (ceil(max_n / kBlock), B); one point per thread; a warp covers one
// 32-point mask word.
__global__ void __launch_bounds__(kBlock) k_insert(const Pipe p)
{
    const FrameDesc fr = p.frames[blockIdx.y];
    const int i = blockIdx.x * kBlock + threadIdx.x;
    if (blockIdx.x * kBlock >= fr.n) return;
    bool won = false;
    int myslot = -1;
    if (i < fr.n) {
        const SegDesc& sg = find_seg(p, fr, i);
        const float* q = sg.pts + (size_t)(i - sg.start) * p.f_in;
        float x, y, z;
        if (p.vec4) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(q));
            x = v.x; y = v.y; z = v.z;
        } else {
            x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
        }
        xform(sg, x, y, z);
        const int cell = cell_of(p.g, x, y, z);
        if (cell >= 0) {
            unsigned long long* tbl = p.table + fr.tbl_off;
            unsigned int h = fr.tbl_shift < 0 ? (unsigned)cell : hash_cell((unsigned)cell, fr.tbl_shift);
            const unsigned long long mine = ((unsigned long long)(unsigned)cell << 32) | (unsigned)i;
            while (true) {
                unsigned long long e = __ldcg(tbl + h);
                bool done = false;
                while (true) {
                    if ((unsigned)(e >> 32) == (unsigned)cell) {
                        if ((unsigned)e > (unsigned)i) {
                            const unsigned long long old = atomicMin(tbl + h, mine);
                            won = (unsigned)old > (unsigned)i;
                        }
                        done = true;
                        break;
                    }
                    if (e == kEmpty64) {
                        const unsigned long long old = atomicCAS(tbl + h, kEmpty64, mine);
                        if (old == kEmpty64) { won = true; done = true; break; }
                        e = old;
                        continue;
                    }
                    break;
                }
                if (done) break;
                h = (h + 1) & fr.tbl_mask;
            }
            myslot = (int)h;
        }
        p.slot[fr.pt_off + i] = myslot;
    }
    const unsigned int b = __ballot_sync(0xffffffffu, won);
    if ((threadIdx.x & 31) == 0 && (i >> 5) < fr.nwords) p.cand[fr.word_off + (i >> 5)] = b;
}

// ---------------------------------------------------------------- k_verify
// grid (ceil(max_words / kBlock), B); one mask word per thread.
__global__ void __launch_bounds__(kBlock) k_verify(const Pipe p)
{
    const FrameDesc fr = p.frames[blockIdx.y];
    const int w = blockIdx.x * kBlock + threadIdx.x;
    if (w >= fr.nwords) return;
    unsigned int m = p.cand[fr.word_off + w];
    unsigned int out = 0;
    const unsigned long long* tbl = p.table + fr.tbl_off;
    while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        const int i = w * 32 + b;
        const int s = p.slot[fr.pt_off + i];
        const unsigned long long e = __ldcg(tbl + s);
        if ((unsigned)e == (unsigned)i) out |= 1u << b;
    }
    p.first[fr.word_off + w] = out;
    p.wordrank[fr.word_off + w] = __popc(out);
}

// ---------------------------------------------------------------- k_rank
__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* smem /*[32]*/)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) smem[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int s = lane < (blockDim.x >> 5) ? smem[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += t;
        }
        smem[lane] = s;  // inclusive over warps
    }
    __syncthreads();
    const int base = wid ? smem[wid - 1] : 0;
    *total = smem[(blockDim.x >> 5) - 1];
    __syncthreads();
    return base + inc - v;
}

// One block per frame.  Scans the per-word first-occurrence counts in place
// (wordrank becomes the exclusive prefix), derives M and the break cut-off; the
// last block to finish prefix-sums M over the frames of the call.
__global__ void __launch_bounds__(kRankThreads) k_rank(const Pipe p)
{
    __shared__ int sm[32];
    __shared__ int s_last;
    const int f = blockIdx.x;
    const FrameDesc fr = p.frames[f];
    int* wr = p.wordrank + fr.word_off;
    const unsigned int* fm = p.first + fr.word_off;
    const int per = (fr.nwords + kRankThreads - 1) / kRankThreads;
    const int w0 = min(threadIdx.x * per, fr.nwords), w1 = min(w0 + per, fr.nwords);
    int local = 0;
    for (int w = w0; w < w1; ++w) local += wr[w];
    int total;
    const int excl = block_exclusive_scan(local, &total, sm);
    const int cap = p.cap;
    const bool cut = (!p.overflow_continue) && total > cap;
    if (threadIdx.x == 0) {
        FrameState st;
        st.total = total;
        st.m = min(total, cap);
        st.istar = fr.n;  // overwritten below by the owner of the cut word
        st.pad = 0;
        p.state[f] = st;
    }
    __syncthreads();
    int run = excl;
    for (int w = w0; w < w1; ++w) {
        const int c = wr[w];
        wr[w] = run;
        if (cut && run <= cap && cap < run + c) {
            // (cap - run)-th set bit (0-based) of first[w] is the point that would open voxel #cap
            unsigned int m = fm[w];
            for (int k = cap - run; k > 0; --k) m &= m - 1;
            p.state[f].istar = w * 32 + (__ffs(m) - 1);
        }
        run += c;
    }
    // ---- last block: exclusive prefix of M over frames
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&p.counters[0], 1) == p.nframes - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    int carry = 0;
    for (int base = 0; base < p.nframes; base += kRankThreads) {
        const int ff = base + threadIdx.x;
        const int m = ff < p.nframes ? ((volatile FrameState*)p.state)[ff].m : 0;
        int tot;
        const int ex = block_exclusive_scan(m, &tot, sm);
        if (ff < p.nframes) p.frame_voxel_offsets[ff] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) {
        p.frame_voxel_offsets[p.nframes] = carry;
        p.counters[0] = 0;
        p.counters[1] = 0;
    }
}

// ---------------------------------------------------------------- k_slot
// grid (ceil(max_n / kBlock), B); one point per thread.
__global__ void __launch_bounds__(kBlock) k_slot(const Pipe p)
{
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    const FrameState st = p.state[f];
    const int i = blockIdx.x * kBlock + threadIdx.x;
    if (i >= fr.n || i >= st.istar) return;
    const int s = p.slot[fr.pt_off + i];
    if (s < 0) return;
    const unsigned long long e = __ldcg(p.table + fr.tbl_off + s);
    const unsigned int fidx = (unsigned)e;
    const int w = fidx >> 5;
    const int rank = p.wordrank[fr.word_off + w] +
                     __popc(p.first[fr.word_off + w] & ((1u << (fidx & 31)) - 1u));
    if (rank >= p.cap) return;  // continue mode: voxel beyond the cap
    const size_t g = (size_t)f * p.cap + rank;
    if (fidx == (unsigned)i) {
        const int cell = (int)(e >> 32);
        const int gx = p.g.gs[0], gy = p.g.gs[1];
        const int x = cell % gx, r = cell / gx;
        const int y = r % gy, z = r / gy;
        int* c = p.coors + (size_t)(p.frame_voxel_offsets[f] + rank) * p.coor_cols;
        if (p.coor_cols == 4) { *reinterpret_cast<int4*>(c) = make_int4(f, z, y, x); }
        else { c[0] = z; c[1] = y; c[2] = x; }
    }
    int* b = p.bucket + g * p.T;
    if (p.T <= kChainMaxT) {
        // sorted insertion: slot k ends up holding the (k+1)-th smallest index offered
        int cur = i;
        for (int k = 0; k < p.T; ++k) {
            const int old = atomicMin(b + k, cur);
            if (old == kEmptyIdx) break;
            cur = max(old, cur);
        }
    } else {
        const int c = atomicAdd(p.cnt + g, 1);
        if (c < p.T) {
            b[c] = i;
        } else {
            const int o = atomicAdd(&p.counters[1], 1);
            if (o < p.ovf_cap) p.ovf[o] = make_int2((int)g, i);
        }
    }
}

// ---------------------------------------------------------------- k_overflow
// Large-T path: a crowded voxel's bucket holds T arbitrary members; every other
// member is offered here and the bucket keeps the T smallest indices.
__global__ void __launch_bounds__(kBlock) k_overflow(const Pipe p)
{
    const int n = min(((volatile int*)p.counters)[1], p.ovf_cap);
    for (int j = blockIdx.x * kBlock + threadIdx.x; j < n; j += gridDim.x * kBlock) {
        const int2 it = p.ovf[j];
        int* b = p.bucket + (size_t)it.x * p.T;
        int mx = 0;
        for (int k = 0; k < p.T; ++k) mx = max(mx, __ldcg(b + k));
        if (it.y > mx) continue;  // entries only decrease: cannot be among the T smallest
        int cur = it.y;
        for (int k = 0; k < p.T; ++k) {
            const int old = atomicMin(b + k, cur);
            cur = max(old, cur);
        }
    }
}

// ---------------------------------------------------------------- k_gather
__device__ __forceinline__ void fetch_row(const Pipe& p, const FrameDesc& fr, int i, float* dst)
{
    const SegDesc& sg = find_seg(p, fr, i);
    const float* q = sg.pts + (size_t)(i - sg.start) * p.f_in;
    float x, y, z;
    if (p.vec4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(q));
        x = v.x; y = v.y; z = v.z;
        dst[3] = v.w;
    } else {
        x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
        for (int k = 3; k < p.f_in; ++k) dst[k] = __ldg(q + k);
    }
    xform(sg, x, y, z);
    dst[0] = x; dst[1] = y; dst[2] = z;
    if (sg.has_dt) dst[p.f_in] = sg.dt;
}

// grid (tiles + reset_blocks, B).  Blocks [0, tiles): a tile of vt voxels (vt*T
// bucket entries) staged through shared memory.  Blocks [tiles, ...): reset the
// table entries the frame touched (every candidate's slot), so the next call
// starts from an empty table without a memset.
// dynamic smem: round4(E) ints + E*f_out floats, E = vt*T.
__global__ void __launch_bounds__(kBlock) k_gather(const Pipe p, int tiles, int vt)
{
    extern __shared__ int smem_i[];
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    if ((int)blockIdx.x >= tiles) {
        const int w = (blockIdx.x - tiles) * kBlock + threadIdx.x;
        if (w >= fr.nwords) return;
        unsigned int m = p.cand[fr.word_off + w];
        unsigned long long* tbl = p.table + fr.tbl_off;
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            tbl[p.slot[fr.pt_off + w * 32 + b]] = kEmpty64;
        }
        return;
    }
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * vt;
    if (v0 >= mf) return;
    const int T = p.T, F = p.f_out;
    const int nv = min(vt, mf - v0);
    const int E = nv * T;
    int* sidx = smem_i;
    float* srow = reinterpret_cast<float*>(smem_i + ((vt * T + 3) & ~3));
    int* gb = p.bucket + ((size_t)f * p.cap + v0) * T;
    for (int e = threadIdx.x; e < E; e += kBlock) {
        sidx[e] = gb[e];
        gb[e] = kEmptyIdx;
    }
    if (T > kChainMaxT)
        for (int v = threadIdx.x; v < nv; v += kBlock) p.cnt[(size_t)f * p.cap + v0 + v] = 0;
    __syncthreads();
    const size_t gout = (size_t)p.frame_voxel_offsets[f] + v0;
    for (int e = threadIdx.x; e < E; e += kBlock) {
        const int v = e / T, s = e - v * T;
        const int my = sidx[e];
        const int* vb = sidx + v * T;
        int nvalid = 0, r = 0;
        for (int k = 0; k < T; ++k) {
            const int o = vb[k];
            nvalid += (o != kEmptyIdx);
            r += (o < my);
        }
        if (my != kEmptyIdx) fetch_row(p, fr, my, srow + (size_t)(v * T + r) * F);
        if (s >= nvalid) {
            float* z = srow + (size_t)(v * T + s) * F;
            for (int k = 0; k < F; ++k) z[k] = 0.0f;
        }
        if (s == 0) p.num_points[gout + v] = nvalid;
    }
    __syncthreads();
    float* out = p.voxels + gout * T * F;
    const int tot = E * F;
    if ((tot & 3) == 0 && ((((size_t)gout * T * F) & 3) == 0)) {
        float4* o4 = reinterpret_cast<float4*>(out);
        const float4* s4 = reinterpret_cast<const float4*>(srow);
        for (int k = threadIdx.x; k < (tot >> 2); k += kBlock) o4[k] = s4[k];
    } else {
        for (int k = threadIdx.x; k < tot; k += kBlock) out[k] = srow[k];
    }
    if (p.mean_out) {
        const int C = p.mean_c;
        for (int k = threadIdx.x; k < nv * C; k += kBlock) {
            const int v = k / C, c = k - v * C;
            const int* vb = sidx + v * T;
            int nvalid = 0;
            for (int j = 0; j < T; ++j) nvalid += (vb[j] != kEmptyIdx);
            float s = 0.0f;
            for (int j = 0; j < T; ++j) s = __fadd_rn(s, srow[(size_t)(v * T + j) * F + c]);
            p.mean_out[(gout + v) * C + c] = __fdiv_rn(s, (float)nvalid);
        }
    }
}

// ---------------------------------------------------------------- mean VFE
// out[v, c] = sum_t voxels[v, t, c] / num[v]; tile of vt voxels staged in smem so
// the [M, T, F] read is fully coalesced.  simple != 0: [ |xy|, z, refl ].
__global__ void __launch_bounds__(kBlock) k_mean_vfe(const float* __restrict__ voxels,
                                                     const int* __restrict__ num, int m, int T,
                                                     int F, int C, int simple, int vt,
                                                     float* __restrict__ out)
{
    extern __shared__ float smem_f[];
    float* smean = smem_f + (size_t)vt * T * F;
    const int v0 = blockIdx.x * vt;
    if (v0 >= m) return;
    const int nv = min(vt, m - v0);
    const size_t base = (size_t)v0 * T * F;
    const int tot = nv * T * F;
    if ((tot & 3) == 0 && (base & 3) == 0) {
        const float4* g4 = reinterpret_cast<const float4*>(voxels + base);
        float4* s4 = reinterpret_cast<float4*>(smem_f);
        for (int k = threadIdx.x; k < (tot >> 2); k += kBlock) s4[k] = __ldg(g4 + k);
    } else {
        for (int k = threadIdx.x; k < tot; k += kBlock) smem_f[k] = __ldg(voxels + base + k);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < nv * C; k += kBlock) {
        const int v = k / C, c = k - v * C;
        float s = 0.0f;
        for (int j = 0; j < T; ++j) s = __fadd_rn(s, smem_f[(size_t)(v * T + j) * F + c]);
        const float q = __fdiv_rn(s, (float)num[v0 + v]);
        if (simple) smean[k] = q; else out[(size_t)(v0 + v) * C + c] = q;
    }
    if (simple) {
        __syncthreads();
        for (int v = threadIdx.x; v < nv; v += kBlock) {
            const float* q = smean + v * C;
            float* o = out + (size_t)(v0 + v) * 3;
            o[0] = sqrtf(__fadd_rn(__fmul_rn(q[0], q[0]), __fmul_rn(q[1], q[1])));
            o[1] = q[2];
            o[2] = q[3];
        }
    }
}

// ---------------------------------------------------------------- BEV scatter
__global__ void __launch_bounds__(kBlock) k_fill_zero(float4* __restrict__ p, size_t n4,
                                                      float* __restrict__ tail, int ntail)
{
    const size_t stride = (size_t)gridDim.x * kBlock;
    for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n4; i += stride)
        p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (blockIdx.x == 0 && (int)threadIdx.x < ntail) tail[threadIdx.x] = 0.f;
}

// one thread per (voxel, channel): canvas[b, c*D + z, y, x] = feats[i, c]
__global__ void __launch_bounds__(kBlock) k_scatter(const float* __restrict__ feats,
                                                    const int4* __restrict__ coors, int m, int c,
                                                    int b, int d, int ny, int nx,
                                                    float* __restrict__ canvas)
{
    const long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (t >= (long long)m * c) return;
    const int i = (int)(t / c), ch = (int)(t - (long long)i * c);
    const int4 q = __ldg(coors + i);
    if ((unsigned)q.x >= (unsigned)b || (unsigned)q.y >= (unsigned)d ||
        (unsigned)q.z >= (unsigned)ny || (unsigned)q.w >= (unsigned)nx)
        return;
    const size_t o = ((((size_t)q.x * c + ch) * d + q.y) * ny + q.z) * nx + q.w;
    canvas[o] = __ldg(feats + t);
}

}  // namespace bvx
