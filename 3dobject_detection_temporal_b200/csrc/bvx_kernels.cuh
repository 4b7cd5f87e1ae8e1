// bvx_kernels.cuh -- sm_100a kernels of the voxelization hot path.
//
// Pipeline (one pass of VoxelGenerator.generate over B frames; DESIGN.md has the
// data layout and the per-kernel roofline).  Work is cut into chunks of 1024
// consecutive points (256 threads x 4 points, all four loads in flight at once):
//   k_insert  per chunk : cell key -> table slot; first occurrence by 64-bit
//                         atomicCAS/atomicMin on {cell:32 | point index:32};
//                         candidate bit per point ("held the minimum when I left")
//   k_verify  per chunk : candidate -> true first-occurrence bits, per-word and
//                         per-chunk counts; the LAST chunk of a frame scans the
//                         chunk counts (first-come voxel ranks, M = min(total,
//                         cap), break cut-off i*), the last frame prefix-sums M
//   k_slot    per chunk : voxel id = rank of the voxel's first point; slot inside
//                         the voxel in input order (sorted atomicMin chain for
//                         T <= 8, count + bucket for larger T)
//   k_overflow          : (T > 8) T-smallest selection for crowded voxels
//   k_gather  per voxel : rows -> voxels (zero padded), coordinates from the first
//                         point, num_points, optional fused mean VFE; extra blocks
//                         of the same launch reset the table entries the frame
//                         touched, so the next call needs no memset
// Everything is fp32 / int32; nothing here is a contraction, hence no tensor
// cores: the bounds are HBM bandwidth and L2 random-access / atomic throughput.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvx {

constexpr unsigned long long kEmpty64 = 0xFFFFFFFFFFFFFFFFull;
constexpr int kEmptyIdx = 0x7FFFFFFF;
constexpr int kChainMaxT = 8;        // T <= 8: sorted atomicMin chain; else count+bucket
constexpr int kBlock = 256;
constexpr int kPts = 4;              // points per thread
constexpr int kChunk = kBlock * kPts;  // 1024 points per chunk
constexpr int kChunkWords = kChunk / 32;
constexpr int kWarps = kBlock / 32;

struct GridParams {
    float lo[3];
    float vs[3];
    float gsf[3];  // grid size as float, x y z
    int gs[3];     // x y z
};

struct FrameDesc {
    const float* pts0;           // points of the frame when it is ONE plain segment, else null
    unsigned long long tbl_off;  // entry offset of this frame's table region
    unsigned int tbl_mask;       // slots - 1 (hashed)
    int tbl_shift;               // 32 - log2(slots) (hashed); -1 => dense (slot = cell)
    int n;                       // points in the frame (all segments)
    int pt_off;                  // base into slot[]
    int word_off;                // base into cand/first/wordlocal
    int nwords;
    int chunk_off;               // base into chunkbase[]
    int nchunks;
    int seg_begin;
    int seg_count;
};

struct SegDesc {
    const float* pts;
    const int* perm;    // optional: row j of the segment is pts[perm[j]] (shuffle fused into the load)
    int n;
    int start;          // first frame-local point index of the segment
    int has_transform;
    int has_dt;
    float R[9];
    float t[3];
    float dt;
    int n_planes;       // crop: a point with a*x + b*y + c*z + d >= 0 for any plane is dropped
    double planes[8][4];
};

struct FrameState {
    int total;        // distinct in-range cells among the inserted points
    int m;            // voxels kept = min(total, cap)
    int istar;        // break cut-off: points with index >= istar are dropped
    int done_verify;  // chunks of k_verify finished (last-block detection); reset by the scanner
    unsigned long long ticket;  // k_insert: bits 0-30 = next chunk of the frame to hand out (in index order),
                                // bit 31 = table known to be empty, high 32 bits = cells claimed so far --
                                // one atomic returns all three
    int n_eff;        // k_insert: start of the first chunk skipped by the break early exit
    int n_done;       // points that went through k_insert (prefix), published by the scan
    int cleared;      // k_insert: table slices cleared so far (hashed grid); reset by k_slot
    int rows;         // k_gather_small, compact-row mode: kept points of the frame (reset by the frame scan)
};

struct Pipe {
    GridParams g;
    const FrameDesc* frames;
    const SegDesc* segs;
    FrameState* state;
    unsigned long long* table;
    int* slot;           // per point: index of the first point of its cell as seen when it was inserted (-1: out of range)
    unsigned int* cand;
    unsigned int* lost;  // per point bit: was the cell's first point for a while, then displaced by an earlier one
    unsigned int* first;
    int* wordlocal;      // exclusive prefix of first-occurrence counts inside the word's chunk
    int* chunkbase;      // per chunk: count, then exclusive prefix over the frame's chunks
    int* bucket;
    int* cnt;
    int2* ovf;
    int* counters;       // [0] frames-done ticket, [1] overflow count, [2] compact rows handed out
    int* frame_voxel_offsets;
    int* voxel_num;      // optional [1]: total number of voxels of the call
    int nframes;
    int f_in;
    int f_out;
    int T;
    int cap;             // per-frame max_voxels of this call
    int overflow_continue;
    int coor_cols;       // 3 (single frame) or 4 (batched, batch index first)
    int frame_base;      // batch index of frame 0 of this call (sub-batches of a pipelined host call)
    int ovf_cap;
    int vec4;            // f_in == 4 and every segment pointer 16-byte aligned
    int clear_rows;      // hashed tables: the first clear_rows block rows of k_insert clear the tables (0: dense grid)
    // outputs
    float4* crow;        // optional, specialised gather only: the kept rows of the call, compact (no padding): a tile
                         // of 256 voxels owns the range starting at tile_base[frame * tiles + tile] (ranges are
                         // handed out in arrival order from counters[2]), voxel after voxel, slots in order
    int* tile_base;
    int* idx_out;        // optional [sum M, T]: frame-local point index of every kept (voxel, slot), -1 = empty.
                         // With it, voxels may be null: the host form assembles the rows from its own copy
                         // of the points (the padded rows are 4-5x their information and PCIe is the bound)
    float* voxels;
    int* coors;
    int* num_points;
    float* mean_out;
    int mean_c;
};

// Programmatic dependent launch: every pipeline kernel lets its successor start being scheduled as
// soon as all of its own blocks are resident (the successor's blocks then fill the SM slots the last
// wave frees), and itself waits here until its predecessor has completed and flushed.  Both
// instructions are no-ops for a launch without the attribute.
__device__ __forceinline__ void pdl_prologue()
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

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

// Convex crop on the raw coordinates, the reference's _points_in_convex_polygon_3d_jit
// (second/core/geometry.py:122-141): float64, sign = x*a + y*b + z*c + d summed left to right,
// outside iff sign >= 0 for some surface.
__device__ __forceinline__ bool cropped_out(const SegDesc& sg, float x, float y, float z)
{
    bool out = false;
    for (int k = 0; k < sg.n_planes; ++k) {
        const double s = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)x, sg.planes[k][0]),
                                                       __dmul_rn((double)y, sg.planes[k][1])),
                                             __dmul_rn((double)z, sg.planes[k][2])),
                                   sg.planes[k][3]);
        out |= s >= 0.0;
    }
    return out;
}

// xyz of frame-local point i (sweep transform applied).  CROP: also apply the frame's crop planes
// (only k_insert needs that: every later kernel only looks at points that got a cell).
template <bool CROP = false>
__device__ __forceinline__ void load_xyz(const Pipe& p, const FrameDesc& fr, int i, float& x, float& y,
                                         float& z)
{
    if (fr.pts0) {
        if (p.vec4) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(fr.pts0) + i);  // streamed once
            x = v.x; y = v.y; z = v.z;
        } else {
            const float* q = fr.pts0 + (size_t)i * p.f_in;
            x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
        }
        return;
    }
    const SegDesc& sg = find_seg(p, fr, i);
    const int j = sg.perm ? __ldg(sg.perm + (i - sg.start)) : i - sg.start;
    const float* q = sg.pts + (size_t)j * p.f_in;
    if (p.vec4) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(q));
        x = v.x; y = v.y; z = v.z;
    } else {
        x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
    }
    if (CROP && sg.n_planes && cropped_out(sg, x, y, z)) { x = __int_as_float(0x7FC00000); return; }  // NaN: no cell
    xform(sg, x, y, z);
}

// full output row (f_out floats) of frame-local point i
__device__ __forceinline__ void fetch_row(const Pipe& p, const FrameDesc& fr, int i, float* dst)
{
    const float* q;
    const SegDesc* sg = nullptr;
    if (fr.pts0) {
        q = fr.pts0 + (size_t)i * p.f_in;
    } else {
        sg = &find_seg(p, fr, i);
        const int j = sg->perm ? __ldg(sg->perm + (i - sg->start)) : i - sg->start;
        q = sg->pts + (size_t)j * p.f_in;
    }
    float x, y, z;
    if (p.vec4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(q));
        x = v.x; y = v.y; z = v.z;
        dst[3] = v.w;
    } else {
        x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
        for (int k = 3; k < p.f_in; ++k) dst[k] = __ldg(q + k);
    }
    if (sg) {
        xform(*sg, x, y, z);
        if (sg->has_dt) dst[p.f_in] = sg->dt;
    }
    dst[0] = x; dst[1] = y; dst[2] = z;
}

// ---------------------------------------------------------------- k_insert
// grid (B, clear rows + chunk rows): x = frame, so blocks are dispatched row by row across all
// frames: first the rows that clear the tables, then the (persistent) chunk rows.  A block takes the frame's next chunk in index order
// from a ticket counter.  Thread t of chunk c owns points c*1024 + k*256 + t,
// k = 0..3: every warp-wide access is coalesced and a warp covers exactly one
// 32-point mask word per k.
//
// Break early exit (the reference's serial loop stops at the point that would open
// voxel #cap): the ticket counter carries, in its high half, the number of distinct
// cells claimed so far (every warp adds its new cells when it is done).  A block
// gets its ticket and that count from ONE atomic.  If the count already exceeds cap,
// every cell it counts was claimed by a chunk with a smaller ticket, so the cut-off
// index i* lies before this chunk and the chunk is skipped (nothing it holds can be
// kept).  The count only grows, so skipping is monotone in the ticket and the
// inserted points always form a prefix [0, n_eff).
//
// Table protocol, CAS first, in ROUNDS (one atomic per point in flight, 4 per
// thread):
//   atomicCAS(empty -> {cell, i}):  was empty   -> new cell, I am a candidate
//                                   same cell   -> atomicMin if my index is smaller
//                                   other cell  -> probe the next slot
// A point whose atomicMin displaces the entry of a later point marks that point in the
// `lost` bitmap, so after the kernel   first = cand & ~lost   holds exactly the points that
// are still table entries -- no pass over the table is needed to find them.  Every point
// also records the first point of its cell as it saw it (slot[]); k_slot trusts that
// record iff the recorded point is still a first point.
template <bool CROP>
__global__ void __launch_bounds__(kBlock, CROP ? 4 : 8) k_insert(const Pipe p)
{
    __shared__ int s_chunk, s_ready;
    pdl_prologue();
    const int f = blockIdx.x;
    const FrameDesc fr = p.frames[f];
    FrameState* st = p.state + f;
    if ((int)blockIdx.y < p.clear_rows) {
        // Clear rows: the frame's table is emptied here, right before it is used, with normal
        // stores -- whole sectors, so the lines are allocated in the L2 without a DRAM fill and
        // the atomics below find them there.  These blocks never wait, so the insert blocks of
        // the frame (later in dispatch order) can safely wait for them.
        if (fr.n > 0) {
            const unsigned int n2 = (fr.tbl_mask + 1u) >> 1;        // uint4 = two slots
            const unsigned int per = (n2 + p.clear_rows - 1) / p.clear_rows;
            const unsigned int lo = blockIdx.y * per, hi = min(n2, lo + per);
            uint4* t4 = reinterpret_cast<uint4*>(p.table + fr.tbl_off);
            const uint4 e = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
            for (unsigned int j = lo + threadIdx.x; j < hi; j += kBlock) t4[j] = e;
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) atomicAdd(&st->cleared, 1);
        }
        return;
    }
    if ((int)(blockIdx.y - p.clear_rows) * kChunk >= fr.n) return;
    const bool brk = !p.overflow_continue;
    unsigned long long* tbl = p.table + fr.tbl_off;
    // Chunk rows are persistent: a block keeps taking tickets of its frame until the frame is
    // exhausted or cut, so the launch carries (rows x frames) blocks, not one block per chunk.
  for (;;) {
    if (threadIdx.x == 0) {
        const unsigned long long t = atomicAdd(&st->ticket, 1ull);
        int c = (int)((unsigned int)t & 0x7FFFFFFFu);
        s_ready = (int)(((unsigned int)t >> 31) | (p.clear_rows == 0));   // bit 31: the table is known to be empty
        if (brk && (int)(t >> 32) > p.cap && c * kChunk < fr.n) {
            atomicMin(&st->n_eff, c * kChunk);
            c = 0x7FFFFFF;                             // skipped
        }
        s_chunk = c;
    }
    __syncthreads();
    const long long basel = (long long)s_chunk * kChunk;
    const int ready = s_ready;
    if (basel >= fr.n) return;
    const int base = (int)basel;
    int cell[kPts];
    unsigned int h[kPts];
    if (!CROP && fr.pts0 && p.vec4) {
        // plain frame of float4 points: all four loads are issued before the first one is used
        // (one DRAM round trip per chunk, not four)
        const float4* q = reinterpret_cast<const float4*>(fr.pts0) + base + threadIdx.x;
        float4 v[kPts];
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (base + k * kBlock + (int)threadIdx.x < fr.n) v[k] = __ldcs(q + k * kBlock);   // streamed once
        }
#pragma unroll
        for (int k = 0; k < kPts; ++k)
            cell[k] = (base + k * kBlock + (int)threadIdx.x < fr.n) ? cell_of(p.g, v[k].x, v[k].y, v[k].z) : -1;
    } else {
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const int i = base + k * kBlock + threadIdx.x;
            cell[k] = -1;
            if (i < fr.n) {
                float x, y, z;
                load_xyz<CROP>(p, fr, i, x, y, z);
                cell[k] = cell_of(p.g, x, y, z);
            }
        }
    }
    if (!ready) {
        // The table must be empty before the first atomic (the point loads above did not need it).
        // The first block of a frame to see all clear rows done says so in bit 31 of the ticket
        // word; blocks that get the bit with their ticket skip this wait and its barrier.  (The clear
        // stores were fenced to the L2 before `cleared` was bumped, and the table is only ever
        // touched by atomics, which act at the L2.)
        if (threadIdx.x == 0) {
            while (*(volatile int*)&st->cleared < p.clear_rows) { }
            __threadfence();
            atomicOr(&st->ticket, 0x80000000ull);
        }
        __syncthreads();
    }
    unsigned int active = 0, needmin = 0, wonmask = 0;
    int nnew = 0;
#pragma unroll
    for (int k = 0; k < kPts; ++k) {
        h[k] = 0;
        if (cell[k] >= 0) {
            h[k] = fr.tbl_shift < 0 ? (unsigned)cell[k] : hash_cell((unsigned)cell[k], fr.tbl_shift);
            active |= 1u << k;
        }
    }
    // (a straight-line, select-based first round in front of this loop measured slower: 62.4 vs 59.3 us)
    while (__any_sync(0xffffffffu, active)) {          // warp-uniform rounds
        unsigned long long r[kPts];
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const unsigned long long mine = ((unsigned long long)(unsigned)cell[k] << 32) |
                                            (unsigned)(base + k * kBlock + threadIdx.x);
            r[k] = 0;
            if (active & (1u << k))
                r[k] = (needmin & (1u << k)) ? atomicMin(tbl + h[k], mine)
                                             : atomicCAS(tbl + h[k], kEmpty64, mine);
        }
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const unsigned int bit = 1u << k;
            if (!(active & bit)) continue;
            const unsigned int i = (unsigned)(base + k * kBlock + threadIdx.x);
            // a point that is done keeps, in h[k], the first point of its cell as it saw it
            if (needmin & bit) {
                const unsigned int o = (unsigned)r[k];
                if (o > i) {
                    wonmask |= bit;
                    atomicOr(p.lost + fr.word_off + (o >> 5), 1u << (o & 31));   // o is no longer first
                    h[k] = i;
                } else {
                    h[k] = o;
                }
                active &= ~bit;
            } else if (r[k] == kEmpty64) {
                wonmask |= bit; ++nnew; active &= ~bit;
                h[k] = i;
            } else if ((unsigned)(r[k] >> 32) == (unsigned)cell[k]) {
                if ((unsigned)r[k] > i) needmin |= bit; else { active &= ~bit; h[k] = (unsigned)r[k]; }
            } else {
                h[k] = (h[k] + 1) & fr.tbl_mask;
            }
        }
    }
    if (brk) {
        // publish the cells this warp claimed (per warp, as soon as it is done: measured better
        // than one add per block); later tickets see them in their ticket word
        nnew = __reduce_add_sync(0xffffffffu, nnew);
        if (nnew && (threadIdx.x & 31) == 0) atomicAdd(&st->ticket, (unsigned long long)nnew << 32);
    }
#pragma unroll
    for (int k = 0; k < kPts; ++k) {
        const int i = base + k * kBlock + threadIdx.x;
        if (i < fr.n) p.slot[fr.pt_off + i] = cell[k] >= 0 ? (int)h[k] : -1;
        const unsigned int b = __ballot_sync(0xffffffffu, (wonmask >> k) & 1u);
        const int w = i >> 5;
        if ((threadIdx.x & 31) == 0 && w < fr.nwords) p.cand[fr.word_off + w] = b;
    }
    __syncthreads();        // s_chunk / s_ready are rewritten by the next round
  }
}

// ---------------------------------------------------------------- k_verify
__device__ __forceinline__ int warp_exclusive_scan(int v, int lane, int* total)
{
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    *total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* smem /*[kWarps+1]*/)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int wt;
    const int ex = warp_exclusive_scan(v, lane, &wt);
    __syncthreads();
    if (lane == 31) smem[wid] = wt;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < kWarps; ++k) {
        const int s = smem[k];
        if (k < wid) base += s;
        tot += s;
    }
    *total = tot;
    return base + ex;
}

// Frame-level part, run by the last chunk of the frame to finish k_verify: scan the
// chunk counts in place (chunkbase becomes the exclusive prefix), publish total / M /
// i*, and -- for the last frame of the call -- the exclusive prefix of M over frames.
__device__ void frame_scan(const Pipe& p, const FrameDesc& fr, int f, int eff, int* sm)
{
    const int nchunks = (eff + kChunk - 1) / kChunk, nwords = (eff + 31) / 32;
    int* cb = p.chunkbase + fr.chunk_off;
    const int cap = p.cap;
    int carry = 0;
    int cut_chunk = -1, cut_base = 0;
    for (int b0 = 0; b0 < nchunks; b0 += kBlock) {
        const int c = b0 + threadIdx.x;
        const int v = c < nchunks ? __ldcg(cb + c) : 0;
        int tot;
        const int ex = carry + block_exclusive_scan(v, &tot, sm);
        if (c < nchunks) {
            cb[c] = ex;
            if (ex <= cap && cap < ex + v) { cut_chunk = c; cut_base = ex; }  // at most one thread
        }
        carry += tot;
    }
    const int total = carry;
    const bool cut = (!p.overflow_continue) && total > cap;
    if (threadIdx.x == 0) {
        p.state[f].total = total;
        p.state[f].m = min(total, cap);
        p.state[f].done_verify = 0;            // ready for the next call
        p.state[f].rows = 0;
        p.state[f].n_done = eff;
        if (!cut) p.state[f].istar = eff;
    }
    {
        // The point that would open voxel #cap: the (cap - prefix)-th first occurrence of the cut
        // chunk.  At most one thread found that chunk; broadcast it, then warp 0 searches the
        // chunk's 32 words with one load per lane.
        __shared__ int s_cut[2];
        if (threadIdx.x == 0) s_cut[0] = -1;
        __syncthreads();
        if (cut_chunk >= 0) { s_cut[0] = cut_chunk; s_cut[1] = cut_base; }
        __syncthreads();
        if (cut && s_cut[0] >= 0 && threadIdx.x < 32) {
            const int cc = s_cut[0];
            int need = cap - s_cut[1];
            const int w0 = fr.word_off + cc * kChunkWords;
            const int nw = min(kChunkWords, nwords - cc * kChunkWords);
            const int lane = threadIdx.x;
            // word w holds the target iff wordlocal[w] <= need < wordlocal[w + 1]; 32 words, one per lane
            const int lo = lane < nw ? __ldcg(p.wordlocal + w0 + lane) : 0x7FFFFFFF;
            const unsigned int le = __ballot_sync(0xffffffffu, lo <= need);
            const int w = 31 - __clz(le);                  // last word whose prefix is <= need (word 0 always is)
            if (lane == w) {
                need -= lo;
                unsigned int m = __ldcg(p.first + w0 + w);
                for (; need > 0; --need) m &= m - 1;
                p.state[f].istar = (cc * kChunkWords + w) * 32 + (__ffs(m) - 1);
            }
        }
    }
    __threadfence();
    __syncthreads();
    __shared__ int s_lastframe;
    if (threadIdx.x == 0) s_lastframe = (atomicAdd(&p.counters[0], 1) == p.nframes - 1);
    __syncthreads();
    if (!s_lastframe) return;
    __threadfence();
    carry = 0;
    for (int b0 = 0; b0 < p.nframes; b0 += kBlock) {
        const int ff = b0 + threadIdx.x;
        const int m = ff < p.nframes ? __ldcg(&p.state[ff].m) : 0;
        int tot;
        const int ex = block_exclusive_scan(m, &tot, sm);
        if (ff < p.nframes) p.frame_voxel_offsets[ff] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) {
        p.frame_voxel_offsets[p.nframes] = carry;
        if (p.voxel_num) *p.voxel_num = carry;         // single-frame form: M for the caller
        p.counters[0] = 0;
        p.counters[1] = 0;
        p.counters[2] = 0;
    }
}

// grid (ceil(max chunks per frame / 8), B): one warp per chunk, one lane per 32-point word.
// first = cand & ~lost; per-word and per-chunk counts; the block that finishes a frame's last
// chunk runs the frame-level scan.  Clears the lost words it read (clean for the next call).
__global__ void __launch_bounds__(kBlock, 8) k_verify(const Pipe p)
{
    __shared__ int sm[kWarps + 1];
    __shared__ int s_last;
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    const int eff = min(fr.n, __ldcg(&p.state[f].n_eff));   // prefix that went through k_insert
    const int nchunks = (eff + kChunk - 1) / kChunk;
    const int c0 = blockIdx.x * kWarps;
    {
        // Every block of the frame also empties its share of the frame's bucket rows (T <= 8:
        // the sorted chain needs them empty) or voxel counters (larger T), right before k_slot
        // fills them: whole sectors, normal stores, so k_slot's atomics find the lines in the L2.
        const bool chain = p.T <= kChainMaxT;
        int* base = chain ? p.bucket + (size_t)f * p.cap * p.T : p.cnt + (size_t)f * p.cap;
        const int rows = min(p.cap, fr.n);                  // a frame of n points opens at most n voxels
        const int total = chain ? rows * p.T : rows;
        const int per = (total + gridDim.x - 1) / gridDim.x;
        const int lo = blockIdx.x * per, hi = min(total, lo + per);
        const int fill = chain ? kEmptyIdx : 0;
        if (hi > lo) {
            // scalar head up to a 16-byte boundary, 16-byte stores, scalar tail
            int* b0 = base + lo;
            const int n = hi - lo;
            const int head = min(n, (int)(((16u - (unsigned)((uintptr_t)b0 & 15u)) & 15u) >> 2));
            if ((int)threadIdx.x < head) b0[threadIdx.x] = fill;
            int4* b4 = reinterpret_cast<int4*>(b0 + head);
            const int n4 = (n - head) >> 2;
            const int4 f4 = make_int4(fill, fill, fill, fill);
            for (int j = threadIdx.x; j < n4; j += kBlock) b4[j] = f4;
            const int t0 = head + n4 * 4;
            if ((int)threadIdx.x < n - t0) b0[t0 + threadIdx.x] = fill;
        }
    }
    if (c0 >= nchunks && !(eff == 0 && blockIdx.x == 0)) return;
    const int nwords = (eff + 31) / 32;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int c = c0 + wid;
    if (c < nchunks) {
        const int w = c * kChunkWords + lane;
        unsigned int fw = 0;
        if (w < nwords) {
            const unsigned int lw = __ldcg(p.lost + fr.word_off + w);
            fw = __ldcg(p.cand + fr.word_off + w) & ~lw;
            if (lw) p.lost[fr.word_off + w] = 0u;
            // points past the inserted prefix carry no bits: cand words are written whole by k_insert
        }
        int tot;
        const int ex = warp_exclusive_scan(__popc(fw), lane, &tot);
        if (w < nwords) {
            p.first[fr.word_off + w] = fw;
            p.wordlocal[fr.word_off + w] = ex;
        }
        if (lane == 0) p.chunkbase[fr.chunk_off + c] = tot;
    }
    // last chunk of the frame -> frame-level scan
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int mine = eff == 0 ? 1 : min(kWarps, nchunks - c0);
        s_last = (atomicAdd(&p.state[f].done_verify, mine) + mine == max(nchunks, 1));
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    frame_scan(p, fr, f, eff, sm);
}

// ---------------------------------------------------------------- k_slot
// Rare path of k_slot: the first point a point recorded at insert time was displaced later
// (chunks of one cell finishing out of order): find the cell's entry in the table.
__device__ __forceinline__ unsigned int refind_first(const Pipe& p, const FrameDesc& fr, int i)
{
    float x, y, z;
    load_xyz(p, fr, i, x, y, z);
    const unsigned int cell = (unsigned)cell_of(p.g, x, y, z);
    const unsigned long long* tbl = p.table + fr.tbl_off;
    unsigned int hh = fr.tbl_shift < 0 ? cell : hash_cell(cell, fr.tbl_shift);
    unsigned long long e = __ldcg(tbl + hh);
    while ((unsigned)(e >> 32) != cell) {
        hh = (hh + 1) & fr.tbl_mask;
        e = __ldcg(tbl + hh);
    }
    return (unsigned)e;
}

// grid (gx, B): block x of frame f walks chunks x, x+gx, ... below the break
// cut-off, so the launch carries no blocks that only exit.  Same round structure as
// k_insert: per round one atomic per point in flight.
__global__ void __launch_bounds__(kBlock, 8) k_slot(const Pipe p)
{
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    const int lim = min(p.state[f].n_done, p.state[f].istar);
    if (blockIdx.x == 0 && threadIdx.x == 0) {   // k_insert's per-frame counters, for the next call
        p.state[f].ticket = 0ull;
        p.state[f].cleared = 0;
        p.state[f].n_eff = 0x7FFFFFFF;
    }
    const int T = p.T;
    for (int base = blockIdx.x * kChunk; base < lim; base += gridDim.x * kChunk) {
        int sl[kPts];
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const int i = base + k * kBlock + threadIdx.x;
            sl[k] = i < lim ? __ldg(p.slot + fr.pt_off + i) : -1;   // first point of my cell, as seen at insert
        }
        // voxel number = rank of the recorded first point among the frame's first points: three
        // small lookups, issued unconditionally (index 0 for out-of-range points) for all four points
        unsigned int fw[kPts];
        int rank[kPts];
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const unsigned int fidx = (unsigned)max(sl[k], 0);
            const int w = fr.word_off + (int)(fidx >> 5);
            fw[k] = __ldg(p.first + w);
            rank[k] = __ldg(p.chunkbase + fr.chunk_off + (int)(fidx / kChunk)) + __ldg(p.wordlocal + w);
        }
        unsigned int stale = 0;
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const unsigned int fidx = (unsigned)max(sl[k], 0);
            const unsigned int bit = (fw[k] >> (fidx & 31)) & 1u;
            if (sl[k] >= 0 && !bit) stale |= 1u << k;
            rank[k] = sl[k] >= 0 ? rank[k] + __popc(fw[k] & ((1u << (fidx & 31)) - 1u)) : 0x7FFFFFFF;
        }
        if (stale) {
            // rare: the recorded point lost its place later; look the cell up in the table
#pragma unroll
            for (int k = 0; k < kPts; ++k) {
                if (!(stale & (1u << k))) continue;
                const unsigned int fidx = refind_first(p, fr, base + k * kBlock + threadIdx.x);
                sl[k] = (int)fidx;
                const int w = fr.word_off + (int)(fidx >> 5);
                rank[k] = __ldg(p.chunkbase + fr.chunk_off + (int)(fidx / kChunk)) + __ldg(p.wordlocal + w) +
                          __popc(__ldg(p.first + w) & ((1u << (fidx & 31)) - 1u));
            }
        }
        // rank >= cap: out of range, cut off, or (continue mode) a voxel beyond the cap
        if (T <= kChainMaxT) {
            // Slot 0 belongs to the voxel's first point (the smallest index by definition): a plain
            // store into the row k_verify just emptied, no round trip.  The other points go through a
            // sorted insertion over slots 1 .. T-1: slot j ends up holding the (j+1)-th smallest index.
            int cur[kPts];
            int* b[kPts];
            unsigned int active = 0;
#pragma unroll
            for (int k = 0; k < kPts; ++k) {
                cur[k] = base + k * kBlock + threadIdx.x;
                b[k] = p.bucket + ((size_t)f * p.cap + (rank[k] < p.cap ? rank[k] : 0)) * T;
                if (rank[k] < p.cap) {
                    if (sl[k] == cur[k]) b[k][0] = cur[k];
                    else active |= 1u << k;
                }
            }
            for (int j = 1; j < T && active; ++j) {
                int old[kPts];
#pragma unroll
                for (int k = 0; k < kPts; ++k)
                    old[k] = (active & (1u << k)) ? atomicMin(b[k] + j, cur[k]) : 0;
#pragma unroll
                for (int k = 0; k < kPts; ++k) {
                    if (active & (1u << k)) {
                        if (old[k] == kEmptyIdx) active &= ~(1u << k);
                        else cur[k] = max(old[k], cur[k]);
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < kPts; ++k) {
                if (rank[k] >= p.cap) continue;
                const int i = base + k * kBlock + threadIdx.x;
                const size_t g = (size_t)f * p.cap + rank[k];
                const int c = atomicAdd(p.cnt + g, 1);
                if (c < T) {
                    p.bucket[g * T + c] = i;
                } else {
                    const int o = atomicAdd(&p.counters[1], 1);
                    if (o < p.ovf_cap) p.ovf[o] = make_int2((int)g, i);
                }
            }
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
// Dense-grid table reset, shared by both gather kernels: blocks past the voxel tiles walk
// the frame's chunks and clear the cell of every candidate (a superset of the entries the
// frame created), so the next call starts from an empty grid without a memset.
__device__ __forceinline__ void reset_chunk(const Pipe& p, const FrameDesc& fr, int chunk)
{
    const int base = chunk * kChunk;
    const int eff = p.state[blockIdx.y].n_done;
    if (base >= eff) return;
    const int nwords = (eff + 31) / 32;
    const int lane = threadIdx.x & 31;
    unsigned long long* tbl = p.table + fr.tbl_off;
#pragma unroll
    for (int k = 0; k < kPts; ++k) {
        const int i = base + k * kBlock + threadIdx.x;
        const int w = i >> 5;
        const unsigned int cw = w < nwords ? __ldg(p.cand + fr.word_off + w) : 0u;
        if ((cw >> lane) & 1u) {
            float x, y, z;
            load_xyz(p, fr, i, x, y, z);
            tbl[cell_of(p.g, x, y, z)] = kEmpty64;      // dense grid: slot = cell
        }
    }
}

__device__ __forceinline__ void write_coor(const Pipe& p, int f, size_t gv, int cell)
{
    const int gx = p.g.gs[0], gy = p.g.gs[1];
    const int x = cell % gx, r = cell / gx;
    const int y = r % gy, z = r / gy;
    if (p.coor_cols == 4) {
        reinterpret_cast<int4*>(p.coors)[gv] = make_int4(f + p.frame_base, z, y, x);
    } else {
        int* c = p.coors + gv * 3;
        c[0] = z; c[1] = y; c[2] = x;
    }
}

// Fast path: T <= 8 (bucket already sorted by the atomicMin chain), 4 floats per
// point, 16-byte aligned rows.  One thread per voxel: T independent row loads in
// flight, mean and coordinates from registers; the voxel rows go through a
// per-warp shared-memory transpose so the [M, T, 4] store is fully coalesced.
// grid (tiles + reset chunks, B); dynamic smem = kWarps * 32 * T * 16 bytes.
template <int T>
__global__ void __launch_bounds__(kBlock) k_gather_small(const Pipe p, int tiles)
{
    extern __shared__ float4 smem_v[];
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    if ((int)blockIdx.x >= tiles) { reset_chunk(p, fr, blockIdx.x - tiles); return; }
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * kBlock;
    if (v0 >= mf) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int v = v0 + threadIdx.x;
    const bool live = v < mf;
    const size_t gbase = (size_t)p.frame_voxel_offsets[f];
    int idx[T];
    int* gb = p.bucket + ((size_t)f * p.cap + v) * T;
#pragma unroll
    for (int s = 0; s < T; ++s) {
        idx[s] = live ? gb[s] : kEmptyIdx;
    }
    float4 row[T];
    int nvalid = 0;
#pragma unroll
    for (int s = 0; s < T; ++s) {
        row[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (idx[s] != kEmptyIdx) {
            row[s] = __ldg(reinterpret_cast<const float4*>(fr.pts0) + idx[s]);
            ++nvalid;
        }
    }
    if (p.voxels) {
        // per-warp transpose: voxel-major float4 index lane*T + s, read back linearly
        float4* sw = smem_v + wid * (32 * T);
#pragma unroll
        for (int s = 0; s < T; ++s) {
            int q = lane * T + s;
            if ((T & 1) == 0) q ^= (q >> 3) & 7;
            sw[q] = row[s];
        }
        __syncwarp();
        const int wv0 = v0 + wid * 32;                    // first voxel of this warp
        const int wn = min(32, mf - wv0);                 // live voxels of this warp (may be <= 0)
        float4* out = reinterpret_cast<float4*>(p.voxels) + (gbase + wv0) * T;
#pragma unroll
        for (int s = 0; s < T; ++s) {
            const int j = s * 32 + lane;
            int q = j;
            if ((T & 1) == 0) q ^= (q >> 3) & 7;
            if (j < wn * T) __stcs(out + j, sw[q]);
        }
    }
    if (p.crow) {
        // compact rows for the host form: the tile reserves a range for its kept points, each voxel
        // writes its rows (slots 0 .. nvalid-1 of the sorted bucket) at its prefix inside the range
        __shared__ int s_scan[kWarps + 1];
        __shared__ int s_base;
        int tot;
        const int ex = block_exclusive_scan(nvalid, &tot, s_scan);
        if (threadIdx.x == 0) {
            s_base = atomicAdd(&p.counters[2], tot);
            p.tile_base[f * tiles + blockIdx.x] = s_base;
            atomicAdd(&p.state[f].rows, tot);
        }
        __syncthreads();
        float4* dst = p.crow + s_base + ex;
#pragma unroll
        for (int s = 0; s < T; ++s)
            if (s < nvalid) __stcs(dst + s, row[s]);
    }
    if (!live) return;
    const size_t gv = gbase + v;
    if (p.idx_out) {
#pragma unroll
        for (int s = 0; s < T; ++s) p.idx_out[gv * T + s] = idx[s] == kEmptyIdx ? -1 : idx[s];
    }
    p.num_points[gv] = nvalid;
    write_coor(p, f, gv, cell_of(p.g, row[0].x, row[0].y, row[0].z));
    if (p.mean_out) {
        float sx = 0.f, sy = 0.f, sz = 0.f, swt = 0.f;
#pragma unroll
        for (int s = 0; s < T; ++s) {
            sx = __fadd_rn(sx, row[s].x); sy = __fadd_rn(sy, row[s].y);
            sz = __fadd_rn(sz, row[s].z); swt = __fadd_rn(swt, row[s].w);
        }
        const float nf = (float)nvalid;
        const float m4[4] = {__fdiv_rn(sx, nf), __fdiv_rn(sy, nf), __fdiv_rn(sz, nf), __fdiv_rn(swt, nf)};
        if (p.mean_c == 4) {
            reinterpret_cast<float4*>(p.mean_out)[gv] = make_float4(m4[0], m4[1], m4[2], m4[3]);
        } else {
            for (int c = 0; c < p.mean_c; ++c) p.mean_out[gv * p.mean_c + c] = m4[c];
        }
    }
}

// General path (any T, any feature count, sweeps): a tile of vt voxels (vt*T bucket
// entries) staged through shared memory; entries are rank-sorted in shared memory
// (the large-T bucket is unordered).
// grid (tiles + reset chunks, B); dynamic smem: round4(E) + round4(vt) ints + E*f_out floats, E = vt*T.
__global__ void __launch_bounds__(kBlock) k_gather(const Pipe p, int tiles, int vt)
{
    extern __shared__ int smem_i[];
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    if ((int)blockIdx.x >= tiles) { reset_chunk(p, fr, blockIdx.x - tiles); return; }
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * vt;
    if (v0 >= mf) return;
    const int T = p.T, F = p.f_out;
    const int nv = min(vt, mf - v0);
    const int E = nv * T;
    int* sidx = smem_i;
    int* scnt = smem_i + ((vt * T + 3) & ~3);                  // live entries per voxel
    float* srow = reinterpret_cast<float*>(scnt + ((vt + 3) & ~3));
    int* gb = p.bucket + ((size_t)f * p.cap + v0) * T;
    for (int e = threadIdx.x; e < E; e += kBlock) sidx[e] = gb[e];
    if (T > kChainMaxT) {
        // count + bucket path: the counter says how many points the voxel saw; the live
        // entries are the first min(count, T) positions of its bucket, in arrival order
        // (what lies behind them is left over from earlier calls)
        for (int v = threadIdx.x; v < nv; v += kBlock)
            scnt[v] = min(p.cnt[(size_t)f * p.cap + v0 + v], T);
        __syncthreads();
    } else {
        __syncthreads();
        for (int v = threadIdx.x; v < nv; v += kBlock) {
            int c = 0;
            for (int k = 0; k < T; ++k) c += (sidx[v * T + k] != kEmptyIdx);
            scnt[v] = c;
        }
        __syncthreads();
    }
    const size_t gout = (size_t)p.frame_voxel_offsets[f] + v0;
    for (int e = threadIdx.x; e < E; e += kBlock) {
        const int v = e / T, s = e - v * T;
        const int my = sidx[e];
        const int nvalid = scnt[v];
        if (s < nvalid) {               // (T <= 8: the chain fills slots 0 .. nvalid-1 in order)
            const int* vb = sidx + v * T;
            int r = 0;
            for (int k = 0; k < nvalid; ++k) r += (vb[k] < my);   // slot = rank of the index in the voxel
            float* dst = srow + (size_t)(v * T + r) * F;
            fetch_row(p, fr, my, dst);
            if (p.idx_out) p.idx_out[(gout + v) * T + r] = my;
            if (r == 0) write_coor(p, f, gout + v, cell_of(p.g, dst[0], dst[1], dst[2]));
        }
        if (s >= nvalid) {
            float* z = srow + (size_t)(v * T + s) * F;
            for (int k = 0; k < F; ++k) z[k] = 0.0f;
            if (p.idx_out) p.idx_out[(gout + v) * T + s] = -1;
        }
        if (s == 0) p.num_points[gout + v] = nvalid;
    }
    __syncthreads();
    float* out = p.voxels + gout * T * F;
    const int tot = E * F;
    if (!p.voxels) {
        // rows are assembled by the host from idx_out
    } else if ((tot & 3) == 0 && ((((size_t)gout * T * F) & 3) == 0)) {
        float4* o4 = reinterpret_cast<float4*>(out);
        const float4* s4 = reinterpret_cast<const float4*>(srow);
        for (int k = threadIdx.x; k < (tot >> 2); k += kBlock) __stcs(o4 + k, s4[k]);
    } else {
        for (int k = threadIdx.x; k < tot; k += kBlock) out[k] = srow[k];
    }
    if (p.mean_out) {
        const int C = p.mean_c;
        for (int k = threadIdx.x; k < nv * C; k += kBlock) {
            const int v = k / C, c = k - v * C;
            const int nvalid = scnt[v];
            float s = 0.0f;
            for (int j = 0; j < T; ++j) s = __fadd_rn(s, srow[(size_t)(v * T + j) * F + c]);
            p.mean_out[(gout + v) * C + c] = __fdiv_rn(s, (float)nvalid);
        }
    }
}

// ---------------------------------------------------------------- descriptor upload
// One block copies the frame and segment descriptors of a call from the handle's pinned host
// buffer (read over PCIe by the SM, no copy engine involved) into device memory.
__global__ void __launch_bounds__(kBlock) k_upload(const int* __restrict__ src_a, int* __restrict__ dst_a, int na,
                                                   const int* __restrict__ src_b, int* __restrict__ dst_b, int nb)
{
    for (int i = threadIdx.x; i < na; i += kBlock) dst_a[i] = src_a[i];
    for (int i = threadIdx.x; i < nb; i += kBlock) dst_b[i] = src_b[i];
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
// m_dev (optional): the number of voxels lives on the device (frame_voxel_offsets[B] of the call that
// produced them); m is then the capacity the launch was sized for.
__global__ void __launch_bounds__(kBlock) k_scatter(const float* __restrict__ feats,
                                                    const int4* __restrict__ coors, int m, int c,
                                                    int b, int d, int ny, int nx,
                                                    float* __restrict__ canvas, const int* __restrict__ m_dev)
{
    const long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (m_dev) m = min(m, __ldg(m_dev));
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
