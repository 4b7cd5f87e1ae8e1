// bvx_kernels.cuh -- sm_100a kernels of the voxelization hot path.
//
// One pass of VoxelGenerator.generate over B frames is TWO kernels (DESIGN.md has
// the data layout and the per-kernel roofline):
//
//   k_chain   persistent blocks take (chunk, frame) tickets; a chunk is kCChunk
//             consecutive points of one frame.  Everything that does not depend on
//             earlier points -- load, cell key, de-duplication of the chunk's
//             cells and the input-order rank of every point among the chunk's
//             points of the same cell -- is done in shared memory first.  Then the
//             block waits for the frame's previous chunk (a release/acquire
//             counter: the frame's chunks form a CHAIN, which is what makes the
//             reference's serial first-come numbering exact in one pass), resolves
//             one representative per distinct cell against the frame's table in
//             L2 (atomicCAS: empty -> new voxel, else the voxel id comes back),
//             numbers the new voxels by a block scan in input order starting at
//             the frame's running voxel count, finds the break cut-off i* / the
//             cap exactly, reserves the points' slots with one atomicAdd per cell
//             and hands the chain to the next chunk.  Off the chain it writes
//             each kept point's index into its (voxel, slot) bucket entry.
//   k_gather  per voxel: rows -> voxels (zero padded), coordinates, num_points,
//             optional fused mean VFE; also returns the workspace to its clean
//             state (table entries of the frame's voxels, counts, chain state),
//             so no call ever needs a memset.
//
// Everything is fp32 / int32; nothing here is a contraction, hence no tensor
// cores: the bounds are HBM bandwidth and L2 atomic round trips on the chain.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvx {

constexpr unsigned long long kEmpty64 = 0xFFFFFFFFFFFFFFFFull;
constexpr unsigned int kEmpty32 = 0xFFFFFFFFu;
constexpr int kVidUnset = -1;                    // entry claimed, voxel number not written yet
constexpr int kVidDead = -2;                     // cell first seen when the frame was already full
constexpr int kBlock = 256;                      // gather / VFE / scatter block
constexpr int kWarps = kBlock / 32;

// chain kernel geometry
#ifndef BVX_CHAIN_THREADS
#define BVX_CHAIN_THREADS 1024
#endif
#ifndef BVX_CHAIN_PTS
#define BVX_CHAIN_PTS 4
#endif
constexpr int kCT = BVX_CHAIN_THREADS;           // threads per chain block
constexpr int kCP = BVX_CHAIN_PTS;               // points per thread
constexpr int kCChunk = kCT * kCP;               // points per chunk
constexpr int kCWarps = kCT / 32;
constexpr int kLSlots = 2 * kCChunk;             // shared-memory hash slots per chunk
constexpr int kLShift = 32 - (kCChunk == 4096 ? 13 : kCChunk == 2048 ? 12 : kCChunk == 8192 ? 14 : 11);
static_assert((1 << (32 - kLShift)) == kLSlots, "kCChunk must be 1024, 2048, 4096 or 8192");
static_assert(kCChunk <= 65535, "local indices are stored as uint16");
constexpr unsigned int kLEnd = 0xFFFFu;

// Table entry.  key = {cell:32 (high) | points offered to the cell so far:32 (low)};
// all ones = empty.  One 64-bit atomic on key both claims a cell and reserves slots.
struct __align__(16) Entry {
    unsigned long long key;
    int vid;        // voxel number of the cell in first-come order (kVidUnset / kVidDead)
    int pad_;
};

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
    int seg_begin;
    int seg_count;
    int pad_;
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

// Per-frame chain state.  `done` is the release/acquire word of the chain.
struct FrameState {
    int done;      // chunks of the frame that have been published; bit 30 = closed (cut-off found)
    int nvox;      // voxels created so far (<= cap)
    int istar;     // break cut-off: points with index >= istar are dropped
    int ndead;     // entries of the dead list (cells claimed beyond the cap)
    int m;         // result: voxels of the frame (read by k_gather; not reset)
    int pad_[3];
};
constexpr int kClosedBit = 1 << 30;

struct Pipe {
    GridParams g;
    const FrameDesc* frames;
    const SegDesc* segs;
    FrameState* state;
    Entry* table;
    unsigned int* vslot;         // [B, cap] table slot of the voxel (for the reset)
    int* bucket;                 // [B, cap, T] point index per (voxel, slot)
    unsigned int* dead;          // [B, kCChunk] table slots of dead cells
    int* counters;               // [0] ticket, [1] frames done
    int* frame_voxel_offsets;
    long long* trace;            // optional [tickets, 8] globaltimer stamps of thread 0 (diagnostics)
    int nframes;
    int rows;                    // chunk rows = max(1, ceil(max n / kCChunk))
    int f_in;
    int f_out;
    int T;
    int cap;                     // per-frame max_voxels of this call
    int overflow_continue;
    int coor_cols;               // 3 (single frame) or 4 (batched, batch index first)
    int vec4;                    // f_in == 4 and every segment pointer 16-byte aligned
    // outputs
    float* voxels;
    int* coors;
    int* num_points;
    float* mean_out;
    int mean_c;
};

__device__ __forceinline__ int ld_acquire(const int* p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ long long gtime()
{
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define BVX_TRACE(slot) do { if (p.trace && tid == 0) p.trace[(size_t)g * 8 + (slot)] = gtime(); } while (0)
__device__ __forceinline__ void st_release(int* p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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

// xyz of frame-local point i (sweep transform applied)
__device__ __forceinline__ void load_xyz(const Pipe& p, const FrameDesc& fr, int i, float& x, float& y,
                                         float& z)
{
    if (fr.pts0) {
        if (p.vec4) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(fr.pts0) + i);
            x = v.x; y = v.y; z = v.z;
        } else {
            const float* q = fr.pts0 + (size_t)i * p.f_in;
            x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
        }
        return;
    }
    const SegDesc& sg = find_seg(p, fr, i);
    const float* q = sg.pts + (size_t)(i - sg.start) * p.f_in;
    if (p.vec4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(q));
        x = v.x; y = v.y; z = v.z;
    } else {
        x = __ldg(q); y = __ldg(q + 1); z = __ldg(q + 2);
    }
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
        q = sg->pts + (size_t)(i - sg->start) * p.f_in;
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

// ---------------------------------------------------------------- k_chain
struct ChainSmem {
    unsigned int lkey[kLSlots];    // cell of the local slot; after the chain: slot base of the cell's voxel
    unsigned int lhead[kLSlots];   // newest point (local index) of the slot's list; after the chain: voxel id
    unsigned short lnext[kCChunk]; // next older list member
    int wcnt[kCP * kCWarps];       // new voxels per (k, warp), then their exclusive prefix
    int s_g, s_done, s_nvox, s_total, s_istar, s_last;
};
static_assert((kCP * kCWarps) % 32 == 0 && (kCP * kCWarps) / 32 <= 8, "scan shape");

template <int NT>
__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* smem /*[NT/32]*/)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int wt;
    const int ex = warp_exclusive_scan(v, lane, &wt);
    __syncthreads();
    if (lane == 31) smem[wid] = wt;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < NT / 32; ++k) {
        const int s = smem[k];
        if (k < wid) base += s;
        tot += s;
    }
    *total = tot;
    return base + ex;
}

// Wait until the frame's chain counter reaches `want` chunks (or the frame is closed).
__device__ __forceinline__ int chain_wait(const int* done, int want)
{
    int d;
    for (;;) {
        d = ld_acquire(done);
        if ((d & kClosedBit) || d >= want) break;
        __nanosleep(20);
    }
    return d;
}

// grid: persistent, block kCT; dynamic smem sizeof(ChainSmem).
// Ticket g -> (row r = g / B, frame f = g % B): chunk r of frame f.  A frame's chunks
// are ticketed in increasing order, so the block that holds chunk r-1 is always
// already running when chunk r waits for it (no deadlock, any grid size).
__global__ void __launch_bounds__(kCT, kCT >= 1024 ? 1 : 2) k_chain(const Pipe p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ChainSmem& sm = *reinterpret_cast<ChainSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int T = p.T, cap = p.cap;
    const int ntickets = p.rows * p.nframes;
    for (;;) {
        __syncthreads();                       // previous ticket is done with shared memory
        if (tid == 0) {
            int g = atomicAdd(&p.counters[0], 1);
            if (g < ntickets && ld_acquire(&p.counters[2])) g = ntickets;   // every frame is complete
            sm.s_g = g;
            // a frame whose cut-off is already known needs none of its later chunks
            sm.s_done = g < ntickets ? ld_acquire(&p.state[g % p.nframes].done) : 0;
        }
        {   // clear the local table while the ticket is in flight
            uint4* c4 = reinterpret_cast<uint4*>(sm.lkey);
            const uint4 e = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
#pragma unroll
            for (int k = 0; k < (2 * kLSlots / 4) / kCT; ++k) c4[k * kCT + tid] = e;
        }
        __syncthreads();
        const int g = sm.s_g;
        if (g >= ntickets) break;
        const int f = g % p.nframes, r = g / p.nframes;
        const FrameDesc fr = p.frames[f];
        const int base = r * kCChunk;
        if ((base >= fr.n && r > 0) || (sm.s_done & kClosedBit)) continue;
        FrameState* st = p.state + f;
        BVX_TRACE(0);

        // ---- phase 1: cell keys (coalesced: thread t owns points base + k*kCT + t)
        int cell[kCP];
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            const int i = base + k * kCT + tid;
            cell[k] = -1;
            if (i < fr.n) {
                float x, y, z;
                load_xyz(p, fr, i, x, y, z);
                cell[k] = cell_of(p.g, x, y, z);
            }
        }
        // ---- phase 2: the chunk's distinct cells (shared-memory hash) + per-cell point lists
        unsigned int pk[kCP];                  // local slot | (rank or multiplicity) << 16
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            pk[k] = 0;
            if (cell[k] >= 0) {
                unsigned int h = hash_cell((unsigned)cell[k], kLShift);
                for (;;) {
                    const unsigned int old = atomicCAS(&sm.lkey[h], kEmpty32, (unsigned)cell[k]);
                    if (old == kEmpty32 || old == (unsigned)cell[k]) break;
                    h = (h + 1) & (kLSlots - 1);
                }
                const unsigned int li = k * kCT + tid;
                const unsigned int prev = atomicExch(&sm.lhead[h], li);
                sm.lnext[li] = (unsigned short)prev;       // kEmpty32 -> kLEnd
                pk[k] = h;
            }
        }
        __syncthreads();
        // ---- phase 3: input-order rank of every point among the chunk's points of its cell
        unsigned int firstmask = 0;
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            if (cell[k] >= 0) {
                const unsigned int li = k * kCT + tid;
                unsigned int j = sm.lhead[pk[k]];
                unsigned int smaller = 0, total = 0;
                while (j != kLEnd) {
                    ++total;
                    smaller += (j < li);
                    j = sm.lnext[j];
                }
                if (smaller == 0) { firstmask |= 1u << k; pk[k] |= total << 16; }
                else pk[k] |= smaller << 16;
            }
        }
        BVX_TRACE(1);
        // ---- phase 4: when the chain is one chunk away, probe the frame's table (plain loads):
        // where does each distinct cell live, or which empty slot would it take?
        Entry* tbl = p.table + fr.tbl_off;
        unsigned int gh[kCP];
        int nvid[kCP];                         // voxel id when the probe already saw it (-1 unknown)
        unsigned int foundmask = 0;
        {
            if (r > 1 && lane == 0) chain_wait(&st->done, r - 1);
            __syncwarp();
            unsigned int unres = firstmask;
#pragma unroll
            for (int k = 0; k < kCP; ++k) {
                nvid[k] = kVidUnset;
                gh[k] = fr.tbl_shift < 0 ? (unsigned)cell[k] : hash_cell((unsigned)cell[k], fr.tbl_shift);
            }
            while (unres) {
#pragma unroll
                for (int k0 = 0; k0 < kCP; k0 += 4) {      // four probes in flight per step (registers)
                    uint4 e[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (unres & (1u << (k0 + q))) e[q] = __ldcg(reinterpret_cast<const uint4*>(tbl + gh[k0 + q]));
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int k = k0 + q;
                        const unsigned int bit = 1u << k;
                        if (!(unres & bit)) continue;
                        if (e[q].y == (unsigned)cell[k]) { foundmask |= bit; nvid[k] = (int)e[q].z; unres &= ~bit; }
                        else if (e[q].y == kEmpty32) unres &= ~bit;
                        else gh[k] = (gh[k] + 1) & fr.tbl_mask;
                    }
                }
            }
        }
        // ---- phase 5: my turn on the chain
        int d = 0, nvox0 = 0;
        if (r > 0) {
            if (lane == 0) { d = chain_wait(&st->done, r); nvox0 = __ldcg(&st->nvox); }
            d = __shfl_sync(0xffffffffu, d, 0);
            nvox0 = __shfl_sync(0xffffffffu, nvox0, 0);
        }
        if (d & kClosedBit) continue;          // the break cut-off lies before this chunk (block-uniform)
        BVX_TRACE(2);
        const bool full = nvox0 >= cap;        // table is frozen: never insert
        // one atomic per distinct cell claims the cell (if new) AND reserves the slots of
        // all its points in this chunk: key = {cell:32 | points offered so far:32}
        unsigned int unres = firstmask, newmask = 0, missmask = 0;
        while (unres) {
            unsigned long long rr[kCP];
#pragma unroll
            for (int k = 0; k < kCP; ++k) {
                const unsigned int bit = 1u << k;
                if (!(unres & bit)) continue;
                const unsigned long long m = (pk[k] >> 16);
                if (foundmask & bit) rr[k] = atomicAdd(&tbl[gh[k]].key, m);
                else if (full) rr[k] = __ldcg(&tbl[gh[k]].key);
                else rr[k] = atomicCAS(&tbl[gh[k]].key, kEmpty64, ((unsigned long long)(unsigned)cell[k] << 32) | m);
            }
#pragma unroll
            for (int k = 0; k < kCP; ++k) {
                const unsigned int bit = 1u << k;
                if (!(unres & bit)) continue;
                if (foundmask & bit) { cell[k] = (int)(unsigned)rr[k]; unres &= ~bit; }        // slot base
                else if (rr[k] == kEmpty64) {
                    if (full) missmask |= bit; else newmask |= bit;
                    cell[k] = 0; unres &= ~bit;
                }
                else if ((unsigned)(rr[k] >> 32) == (unsigned)cell[k]) foundmask |= bit;       // came in since the probe
                else gh[k] = (gh[k] + 1) & fr.tbl_mask;
            }
        }
        // from here cell[k] of a first point holds the slot base of its cell
        BVX_TRACE(3);
        // ---- phase 6: count the new voxels and hand the chain on
        const unsigned int cntmask = newmask | missmask;   // break + full: a miss is the cut-off point
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            const unsigned int b = __ballot_sync(0xffffffffu, (cntmask >> k) & 1u);
            if (lane == 0) sm.wcnt[k * kCWarps + warp] = __popc(b);
        }
        __syncthreads();
        if (warp == 0) {
            constexpr int per = kCP * kCWarps / 32;
            int v[per], s = 0;
#pragma unroll
            for (int q = 0; q < per; ++q) { v[q] = sm.wcnt[lane * per + q]; s += v[q]; }
            int tot;
            int ex = warp_exclusive_scan(s, lane, &tot);
            if (lane == 0) {
                const bool over0 = nvox0 + tot > cap;
                const bool cut0 = over0 && !p.overflow_continue;
                const bool lastchunk = (long long)base + kCChunk >= fr.n;
                const int nv1 = over0 ? cap : nvox0 + tot;
                st->nvox = nv1;
                if (over0 && !full) st->ndead = nvox0 + tot - cap;
                if (cut0 || lastchunk) st->m = nv1;
                st_release(&st->done, (r + 1) | (cut0 ? kClosedBit : 0));   // hand the chain on
                if (p.trace) p.trace[(size_t)g * 8 + 5] = gtime();
                sm.s_total = tot;
                int last = 0;
                if (cut0 || lastchunk) {
                    __threadfence();
                    last = (atomicAdd(&p.counters[1], 1) == p.nframes - 1);
                }
                sm.s_last = last;
                sm.s_istar = 0x7FFFFFFF;
            }
#pragma unroll
            for (int q = 0; q < per; ++q) { sm.wcnt[lane * per + q] = ex; ex += v[q]; }
        }
        __syncthreads();
        // ---- off the chain: number the new voxels in input order
        const int total_new = sm.s_total;
        const bool over = nvox0 + total_new > cap;
        const bool cut = over && !p.overflow_continue;
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            const unsigned int bit = 1u << k;
            const unsigned int b = __ballot_sync(0xffffffffu, (cntmask >> k) & 1u);
            if (!(cntmask & bit)) continue;
            const int vid = nvox0 + sm.wcnt[k * kCWarps + warp] + __popc(b & ((1u << lane) - 1u));
            if (cut && vid == cap) sm.s_istar = base + k * kCT + tid;   // the point that would open voxel #cap
            if (missmask & bit) { nvid[k] = kVidDead; continue; }
            if (vid < cap) {
                tbl[gh[k]].vid = vid;
                p.vslot[(size_t)f * cap + vid] = gh[k];
                nvid[k] = vid;
            } else {       // claimed beyond the cap: dead; remember the slot for the reset
                tbl[gh[k]].vid = kVidDead;
                p.dead[(size_t)f * kCChunk + (vid - cap)] = gh[k];
                nvid[k] = kVidDead;
            }
        }
        BVX_TRACE(4);
        if (cut) __syncthreads();
        const int istar = cut ? sm.s_istar : 0x7FFFFFFF;
        // per-cell voxel id / slot base for the other points of the cell
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            const unsigned int bit = 1u << k;
            if (!(firstmask & bit)) continue;
            const unsigned int lh = pk[k] & 0xFFFFu;
            int vid = nvid[k];
            if (vid == kVidUnset) {            // created by a recent chunk: its number is on its way
                const int* vp = &tbl[gh[k]].vid;
                while ((vid = __ldcg(vp)) == kVidUnset) __nanosleep(20);
            }
            if (cut && vid >= 0) {             // only points before the cut-off count: give the rest back
                int keep = 0;
                unsigned int j = sm.lhead[lh];
                while (j != kLEnd) { keep += (base + (int)j < istar); j = sm.lnext[j]; }
                const int m = (int)(pk[k] >> 16);
                if (keep < m) atomicAdd(&tbl[gh[k]].key, (unsigned long long)(long long)(keep - m));
            }
            nvid[k] = vid;
        }
        __syncthreads();                       // all list walks done before the lists are overwritten
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            if (firstmask & (1u << k)) {
                const unsigned int lh = pk[k] & 0xFFFFu;
                sm.lhead[lh] = (unsigned)nvid[k];
                sm.lkey[lh] = (unsigned)cell[k];
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kCP; ++k) {
            const int i = base + k * kCT + tid;
            const unsigned int lh = pk[k] & 0xFFFFu;
            const bool first = (firstmask >> k) & 1u;
            const unsigned int rk = first ? 0u : (pk[k] >> 16);
            const bool took = first || rk > 0;      // non-first points always have rank >= 1
            if (!took || i >= istar) continue;
            const int vid = (int)sm.lhead[lh];
            if (vid < 0) continue;
            const unsigned int slot = sm.lkey[lh] + rk;
            if (slot < (unsigned)T) p.bucket[((size_t)f * cap + vid) * T + slot] = i;
        }
        BVX_TRACE(6);
        if (sm.s_last) {
            // every frame is complete: exclusive prefix of M over frames
            __threadfence();
            int carry = 0;
            for (int b0 = 0; b0 < p.nframes; b0 += kCT) {
                const int ff = b0 + tid;
                const int m = ff < p.nframes ? __ldcg(&p.state[ff].m) : 0;
                int tot;
                const int ex = block_exclusive_scan<kCT>(m, &tot, sm.wcnt);
                if (ff < p.nframes) p.frame_voxel_offsets[ff] = carry + ex;
                carry += tot;
            }
            if (tid == 0) {
                p.frame_voxel_offsets[p.nframes] = carry;
                p.counters[1] = 0;
                __threadfence();
                st_release(&p.counters[2], 1);
            }
        }
    }
}

// ---------------------------------------------------------------- k_gather
__device__ __forceinline__ void reset_entry(Entry* e)
{
    *reinterpret_cast<uint4*>(e) = make_uint4(kEmpty32, kEmpty32, (unsigned)kVidUnset, 0u);
}

// Block 0 of every frame returns the chain state to its clean state for the next
// call: dead-cell table entries, the frame's chain words, the call-wide counters.
__device__ __forceinline__ void housekeeping(const Pipe& p, const FrameDesc& fr, int f)
{
    FrameState* st = p.state + f;
    const int nd = st->ndead;
    Entry* tbl = p.table + fr.tbl_off;
    for (int j = threadIdx.x; j < nd; j += blockDim.x) reset_entry(tbl + p.dead[(size_t)f * kCChunk + j]);
    __syncthreads();
    if (threadIdx.x == 0) {
        st->done = 0; st->nvox = 0; st->istar = 0x7FFFFFFF; st->ndead = 0;
        if (f == 0) { p.counters[0] = 0; p.counters[2] = 0; }
    }
}

__device__ __forceinline__ void write_coor(const Pipe& p, int f, size_t gv, int cell)
{
    const int gx = p.g.gs[0], gy = p.g.gs[1];
    const int x = cell % gx, r = cell / gx;
    const int y = r % gy, z = r / gy;
    if (p.coor_cols == 4) {
        reinterpret_cast<int4*>(p.coors)[gv] = make_int4(f, z, y, x);
    } else {
        int* c = p.coors + gv * 3;
        c[0] = z; c[1] = y; c[2] = x;
    }
}

// Fast path: T <= 8, 4 floats per point, 16-byte aligned rows, plain frames.  One
// thread per voxel: its table entry gives the cell (-> coordinates) and the number
// of points offered (-> num_points); T independent row loads in flight; the voxel
// rows go through a per-warp shared-memory transpose so the [M, T, 4] store is
// fully coalesced.  grid (tiles, B); dynamic smem = kWarps * 32 * T * 16 bytes.
template <int T>
__global__ void __launch_bounds__(kBlock) k_gather_small(const Pipe p)
{
    extern __shared__ float4 smem_v[];
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * kBlock;
    if (blockIdx.x == 0) housekeeping(p, fr, f);
    if (v0 >= mf) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int v = v0 + threadIdx.x;
    const bool live = v < mf;
    const size_t gbase = (size_t)p.frame_voxel_offsets[f];
    const size_t fv = (size_t)f * p.cap + v;
    int nvalid = 0, cell = 0;
    if (live) {
        Entry* e = p.table + fr.tbl_off + __ldcs(p.vslot + fv);
        const uint4 ev = __ldcg(reinterpret_cast<const uint4*>(e));
        nvalid = (int)min(ev.x, (unsigned)T);
        cell = (int)ev.y;
        reset_entry(e);                             // workspace back to clean
    }
    int idx[T];
    const int* gb = p.bucket + fv * T;
#pragma unroll
    for (int s = 0; s < T; ++s) idx[s] = s < nvalid ? __ldcs(gb + s) : -1;
    float4 row[T];
#pragma unroll
    for (int s = 0; s < T; ++s) {
        row[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (idx[s] >= 0) row[s] = __ldg(reinterpret_cast<const float4*>(fr.pts0) + idx[s]);
    }
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
    if (!live) return;
    const size_t gv = gbase + v;
    p.num_points[gv] = nvalid;
    write_coor(p, f, gv, cell);
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
// entries) staged through shared memory.
// grid (tiles, B); dynamic smem: round4(vt) ints + vt*T*f_out floats.
__global__ void __launch_bounds__(kBlock) k_gather(const Pipe p, int vt)
{
    extern __shared__ int smem_i[];
    const int f = blockIdx.y;
    const FrameDesc fr = p.frames[f];
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * vt;
    if (blockIdx.x == 0) housekeeping(p, fr, f);
    if (v0 >= mf) return;
    const int T = p.T, F = p.f_out;
    const int nv = min(vt, mf - v0);
    const int E = nv * T;
    int* scnt = smem_i;
    float* srow = reinterpret_cast<float*>(smem_i + ((vt + 3) & ~3));
    const size_t fv0 = (size_t)f * p.cap + v0;
    const size_t gout = (size_t)p.frame_voxel_offsets[f] + v0;
    for (int v = threadIdx.x; v < nv; v += kBlock) {
        Entry* e = p.table + fr.tbl_off + __ldcs(p.vslot + fv0 + v);
        const uint4 ev = __ldcg(reinterpret_cast<const uint4*>(e));
        const int c = (int)min(ev.x, (unsigned)T);
        scnt[v] = c;
        reset_entry(e);
        p.num_points[gout + v] = c;
        write_coor(p, f, gout + v, (int)ev.y);
    }
    __syncthreads();
    const int* gb = p.bucket + fv0 * T;
    for (int e = threadIdx.x; e < E; e += kBlock) {
        const int v = e / T, s = e - v * T;
        float* dst = srow + (size_t)e * F;
        if (s < scnt[v]) {
            fetch_row(p, fr, __ldcs(gb + e), dst);
        } else {
            for (int k = 0; k < F; ++k) dst[k] = 0.0f;
        }
    }
    __syncthreads();
    float* out = p.voxels + gout * T * F;
    const int tot = E * F;
    if ((tot & 3) == 0 && ((((size_t)gout * T * F) & 3) == 0)) {
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
            float s = 0.0f;
            for (int j = 0; j < T; ++j) s = __fadd_rn(s, srow[(size_t)(v * T + j) * F + c]);
            p.mean_out[(gout + v) * C + c] = __fdiv_rn(s, (float)scnt[v]);
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
