// bvx_kernels.cuh -- sm_100a kernels of the voxelization hot path.
//
// Pipeline (one pass of VoxelGenerator.generate over B frames; DESIGN.md has the
// data layout and the per-kernel roofline).  Work is cut into chunks of 1024
// consecutive points; a warp owns 128 consecutive points = one "unit" of four 32-point
// bit words; a k_insert block takes 512 points (4 warps x 4 points per thread) per ticket:
//   k_insert  per ticket: cell key -> table slot; first occurrence by 64-bit
//                         atomicCAS/atomicMin on {cell:32 | point index:32};
//                         candidate bit per point ("held the minimum when I left"),
//                         `lost` bit for a candidate displaced later; every in-range
//                         point that is NOT the first of its cell goes, with its
//                         table slot, onto the chunk's compact "extras" list
//   k_mid     per chunk : scan of the per-chunk first-occurrence counts (every block,
//                         in shared memory) -> voxel number of every first point
//                         (written as vfirst[voxel]), M = min(total, cap), the break
//                         cut-off i*; then only the EXTRAS (the 2nd, 3rd ... point of
//                         a cell) look their voxel up and take their slot by a sorted
//                         atomicMin chain (T <= 8) or count + bucket (larger T)
//   k_overflow          : (T > 8) T-smallest selection for crowded voxels
//   k_gather  per voxel : rows -> voxels (zero padded), coordinates from the first
//                         point, num_points, optional fused mean VFE; puts every
//                         per-call structure back into its between-calls state
// Everything is fp32 / int32; nothing here is a contraction, hence no tensor
// cores: the bounds are HBM bandwidth and L2 random-access / atomic throughput.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef BVX_BULK_LOAD
#define BVX_BULK_LOAD 0   // k_insert: ticket's points through cp.async.bulk into shared memory (experiment)
#endif
#ifndef BVX_TICKET
#define BVX_TICKET 512     // points per k_insert ticket (block iteration): 1024 or 512
#endif
#ifndef BVX_GATHER_VPT
#define BVX_GATHER_VPT 1   // measured cfg4: 2 voxels per thread 46.7 us (4 blocks/SM) .. 58.9 (8, spills) against 43.9 for 1
#endif

namespace bvx {

constexpr unsigned long long kEmpty64 = 0xFFFFFFFFFFFFFFFFull;
constexpr int kEmptyIdx = 0x7F7F7F7F;  // empty bucket entry (a cudaMemset pattern); > every point index
constexpr int kMaxFramePoints = 0x7F000000;
constexpr int kChainMaxT = 8;        // T <= 8: sorted atomicMin chain; else count+bucket
constexpr int kBlock = 256;
constexpr int kPts = 4;              // points per thread
constexpr int kChunk = kBlock * kPts;  // 1024 points per chunk
constexpr int kChunkWords = kChunk / 32;
constexpr int kWarps = kBlock / 32;
constexpr int kUnit = 32 * kPts;     // 128 points: what one warp owns of a chunk
constexpr int kTicket = BVX_TICKET;     // points a k_insert block takes per ticket: a part of a chunk
constexpr int kParts = kChunk / kTicket;  // parts per chunk; each has its own extras list
constexpr int kIThreads = kTicket / kPts, kIWarps = kIThreads / 32;
constexpr int kScanSmemChunks = 2048;  // frames up to this many chunks are scanned in shared memory by k_mid
constexpr int kGatherVpt = BVX_GATHER_VPT;   // voxels per thread of the specialised gather
constexpr int kInlineFrames = 64;      // calls of up to this many frames carry their frame descriptors as kernel arguments
#ifndef BVX_GATHER_PF
#define BVX_GATHER_PF 0
#endif
#ifndef BVX_GATHER_BLOCKS
#define BVX_GATHER_BLOCKS 6   // 40 registers; the kernel alone is 1.5 us slower than with 8 blocks of 32, the back-to-back step 6 us faster
#endif
#ifndef BVX_INSERT_BLOCKS_CONTINUE
#define BVX_INSERT_BLOCKS_CONTINUE 8
#endif
#ifndef BVX_INSERT_BLOCKS
#define BVX_INSERT_BLOCKS 5   // break mode: persistent insert blocks in flight = 148 x 5 x 256 threads (measured 4 / 5 / 6 / 8: see DESIGN)
#endif

struct GridParams {
    float lo[3];
    float vs[3];
    float rvs[3];  // 1 / vs rounded to nearest (NaN: always take the exact divide)
    float mg[3];   // 2^-21 (grid + 1), rounded up: margin of the divide-free cell index
    float gsf[3];  // grid size as float, x y z
    int gs[3];     // x y z
};

struct FrameDesc {
    long long pts_off;           // plain frame: float offset of its first row from Pipe::base
    unsigned long long tbl_off;  // entry offset of this frame's table region
    unsigned int tbl_mask;       // slots - 1 (hashed)
    int tbl_shift;               // 32 - log2(slots) (hashed); -1 => dense (slot = cell)
    int n;                       // points in the frame (all segments)
    int plain;                   // ONE segment, no transform / dt / permutation / crop: rows are read from base + pts_off
    int word_off;                // base into cand / lost (multiple of 4)
    int nwords;
    int chunk_off;               // base into chunkstat / extn / chunkbase; x8 into unit8; x1024 into extras
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
    int clear_claim;  // k_insert: table slices handed out so far (hashed grid)
    unsigned long long ticket;  // k_insert: bits 0-30 = next chunk of the frame to hand out (in index order),
                                // bit 31 = table known to be empty, high 32 bits = cells claimed so far --
                                // one atomic returns all three
    int n_eff;        // k_insert: start of the first chunk skipped by the break early exit
    int n_done;       // points that went through k_insert (prefix), published by k_mid
    int cleared;      // k_insert: table slices cleared so far (hashed grid)
    int rows;         // k_gather_small, compact-row mode: kept points of the frame (reset by k_mid)
};

struct Pipe {
    GridParams g;
    const FrameDesc* frames;
    const SegDesc* segs;
    FrameState* state;
    const float* base;   // plain frames: rows start at base + FrameDesc::pts_off
    unsigned long long* table;
    unsigned int* cand;
    unsigned int* lost;  // per point bit: was the cell's first point for a while, then displaced by an earlier one
    unsigned char* unit8;     // per chunk 16 bytes: bytes 0-7 = candidate bits of its eight 128-point units (0 .. 128 each),
                              // bytes 8-15 = displaced candidates per unit (two words, atomics; between calls: 0)
    unsigned int* chunkstat;  // per chunk: first occurrences (low 16 bits) | displaced candidates (high 16)
    int2* extras;        // per chunk, 1024 entries: {point index, table slot of its cell} of every in-range point
                         // that is not the first of its cell (as seen when it was inserted, or displaced later)
    int* extn;           // per chunk: entries on its extras list
    int* chunkbase;      // frames of more than kScanSmemChunks chunks only: exclusive prefix of the chunk counts
    int* vfirst;         // [B, cap] point index of the first point of every voxel
    int* bucket;         // [B, cap, T-1] the other kept points of a voxel (between calls: all kEmptyIdx for T <= 8)
    int* cnt;            // [B, cap] T > 8: points beyond the first seen by the voxel (between calls: all 0)
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
    int tail_clear;      // hashed tables, small batches: the gather's trailing blocks empty the tables for the NEXT call in
                         // this many slices per frame and leave tail_sig in counters[4..5]; a call that finds its own
                         // signature there (same table regions, nothing ran in between) skips the clear and its handshake
    unsigned int tail_sig[2];   // 64-bit signature of the call's table regions (never 0)
    int clear_rows;      // hashed tables: slices each table is cleared in = leading block rows of k_insert (0: dense grid)
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
    int inline_frames;   // frame descriptors are in inl[] (constant bank: no load chain at block start, no upload)
    FrameDesc inl[kInlineFrames];
};

__device__ __forceinline__ FrameDesc frame_desc(const Pipe& p, int f)
{
    if (p.inline_frames) return p.inl[f];
    return p.frames[f];
}

// Programmatic dependent launch: every pipeline kernel lets its successor start being scheduled as
// soon as all of its own blocks are resident (the successor's blocks then fill the SM slots the last
// wave frees), and itself waits here until its predecessor has completed and flushed.  Both
// instructions are no-ops for a launch without the attribute.
#ifdef BVX_TIMELINE
// experiment builds only (tools/variants.sh): first block start / last block end of every kernel role
__device__ unsigned long long g_tl[32];
struct TlScope {
    int k;
    __device__ __forceinline__ explicit TlScope(int k_) : k(k_)
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (threadIdx.x == 0) atomicMin(&g_tl[2 * k], t);
    }
    __device__ __forceinline__ ~TlScope()
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (threadIdx.x == 0) atomicMax(&g_tl[2 * k + 1], t);
    }
};
#define BVX_TL(k) TlScope tl_scope_(k)
#define BVX_TL_MARK(k) do { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
        if ((threadIdx.x & 31) == 0) { atomicMin(&g_tl[2 * (k)], t_); atomicMax(&g_tl[2 * (k) + 1], t_); } } while (0)
#else
#define BVX_TL(k)
#define BVX_TL_MARK(k)
#endif

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

// floor(RN((v - lo) / size)), the reference's fp32 expression, without the divide where its result
// is certain.  a = RN(v - lo); t = RN(a * RN(1/size)) differs from the correctly rounded quotient q
// by less than 2^-22 |t| (three roundings of 2^-24 each), so whenever t lies further than 2^-21 |t| from
// both neighbouring integers, floor(q) == floor(t).  The margin is the constant mg = 2^-21 (grid + 1):
// it covers every t in [-1, grid + 1], and outside that interval any floor on the right side of the grid
// boundary classifies the point as out of range, which is all that is needed there.  Otherwise (about one
// point in a thousand; also NaN, Inf, |t| >= 2^23, a size whose reciprocal is not a normal number: rvs = NaN)
// the IEEE divide decides: ONE branch per point.
// -1 if outside / NaN, else the linear cell index in the reference's zyx grid order.
__device__ __forceinline__ bool cell_xyz(const GridParams& g, float x, float y, float z, int& ix, int& iy, int& iz)
{
    const float ax = __fsub_rn(x, g.lo[0]), ay = __fsub_rn(y, g.lo[1]), az = __fsub_rn(z, g.lo[2]);
    const float tx = __fmul_rn(ax, g.rvs[0]), ty = __fmul_rn(ay, g.rvs[1]), tz = __fmul_rn(az, g.rvs[2]);
    float cx = floorf(tx), cy = floorf(ty), cz = floorf(tz);
    const float fx = __fsub_rn(tx, cx), fy = __fsub_rn(ty, cy), fz = __fsub_rn(tz, cz);     // exact
    // (a NaN makes its comparison false)
    const bool safe = (fx > g.mg[0]) & (__fsub_rn(1.0f, fx) > g.mg[0]) & (fy > g.mg[1]) & (__fsub_rn(1.0f, fy) > g.mg[1]) &
                      (fz > g.mg[2]) & (__fsub_rn(1.0f, fz) > g.mg[2]);
    if (!safe) {
        cx = floorf(__fdiv_rn(ax, g.vs[0]));
        cy = floorf(__fdiv_rn(ay, g.vs[1]));
        cz = floorf(__fdiv_rn(az, g.vs[2]));
    }
    const bool ok = (cx >= 0.0f) && (cx < g.gsf[0]) && (cy >= 0.0f) && (cy < g.gsf[1]) &&
                    (cz >= 0.0f) && (cz < g.gsf[2]);
    ix = (int)cx; iy = (int)cy; iz = (int)cz;
    return ok;
}

__device__ __forceinline__ int cell_of(const GridParams& g, float x, float y, float z)
{
    int ix, iy, iz;
    if (!cell_xyz(g, x, y, z, ix, iy, iz)) return -1;
    return (iz * g.gs[1] + iy) * g.gs[0] + ix;
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

__device__ __forceinline__ const float* plain_rows(const Pipe& p, const FrameDesc& fr)
{
    return p.base + fr.pts_off;
}

// xyz of frame-local point i (sweep transform applied).  CROP: also apply the frame's crop planes
// (only k_insert needs that: every later kernel only looks at points that got a cell).
template <bool CROP = false>
__device__ __forceinline__ void load_xyz(const Pipe& p, const FrameDesc& fr, int i, float& x, float& y,
                                         float& z)
{
    if (fr.plain) {
        const float* q0 = plain_rows(p, fr);
        if (p.vec4) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(q0) + i);  // streamed once
            x = v.x; y = v.y; z = v.z;
        } else {
            const float* q = q0 + (size_t)i * p.f_in;
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
    if (fr.plain) {
        q = plain_rows(p, fr) + (size_t)i * p.f_in;
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
// frames: first the rows that clear the tables, then the (persistent) chunk rows.  A block takes
// the frame's next chunk in index order from a ticket counter.  Warp w of chunk c owns points
// c*1024 + w*128 + k*32 + lane, k = 0..3: every warp-wide access is coalesced, a warp covers four
// consecutive 32-point mask words.
//
// Break early exit (the reference's serial loop stops at the point that would open
// voxel #cap): the ticket counter carries, in its high half, the number of distinct
// cells claimed so far (every warp adds its new cells when it is done).  A block
// gets its ticket and that count from ONE atomic.  If the count already exceeds cap,
// every cell it counts was claimed by a chunk with a smaller ticket, so the cut-off
// index i* lies before this chunk and the chunk is skipped (nothing it holds can be
// kept).  The count only grows, so skipping is monotone in the ticket and the
// inserted points always form a prefix [0, n_eff).  The ticket of the NEXT chunk is taken
// by the first warp as soon as its own insert rounds are over, so its round trip overlaps
// the epilogue and the wait for the other warps (one block barrier per chunk).  (Taking it
// at the START of the current chunk hid it completely but made the count one chunk old:
// 83 instead of 63 chunks per frame were processed before the exit, cfg4.)
//
// Table protocol, CAS first, in ROUNDS (one atomic per point in flight, 4 per
// thread):
//   atomicCAS(empty -> {cell, i}):  was empty   -> new cell, I am a candidate
//                                   same cell   -> atomicMin if my index is smaller
//                                   other cell  -> probe the next slot
// A point whose atomicMin displaces the entry of a later point marks that point in the
// `lost` bitmap, so after the kernel   first = cand & ~lost   holds exactly the points that
// are still table entries -- no pass over the table is needed to find them.  Every in-range
// point that is not a candidate when it leaves, and every displaced point (recorded by its
// displacer), goes on the chunk's extras list with the table slot of its cell: those are all
// the points that are not the first of their voxel, and only they need a pass of their own.
// (Recording the first point as seen instead of the slot saves k_mid the table access only for
// warps in which no lane's record went stale -- one warp in ten: measured slower.)
__device__ __forceinline__ void clear_slice(const Pipe& p, const FrameDesc& fr, int s, int nslices)
{
    const unsigned int n2 = (fr.tbl_mask + 1u) >> 1;        // uint4 = two slots
    const unsigned int per = (n2 + nslices - 1) / nslices;
    const unsigned int lo = min(n2, (unsigned)s * per), hi = min(n2, lo + per);
    uint4* t4 = reinterpret_cast<uint4*>(p.table + fr.tbl_off);
    const uint4 e = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
    for (unsigned int j = lo + threadIdx.x; j < hi; j += blockDim.x) t4[j] = e;
}

#if BVX_BULK_LOAD
// 1-D bulk copy global -> shared through the TMA unit (cp.async.bulk), completion on an mbarrier
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, unsigned bytes, unsigned long long* bar)
{
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar), d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(d), "l"(gsrc), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
    const unsigned b = (unsigned)__cvta_generic_to_shared(bar);
    asm volatile("{\n\t.reg .pred p;\n\tWAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE;\n\tbra WAIT;\n\tDONE:\n\t}"
                 ::"r"(b), "r"(parity) : "memory");
}
#endif

// OCC = resident 256-thread-equivalents per SM the kernel is compiled for (register cap): break mode runs best at 5-6
// (latency chains of fewer, fatter blocks; less overshoot past the cut-off), continue mode at 8 (more atomics in flight).
template <bool CROP, int OCC>
__global__ void __launch_bounds__(kIThreads, OCC * kParts) k_insert(const __grid_constant__ Pipe p)
{
    __shared__ int s_tk[2], s_rdy[2], s_extn[2], s_slice;
#if BVX_BULK_LOAD
    __shared__ __align__(128) float4 s_pts[2][kTicket];
    __shared__ __align__(8) unsigned long long s_bar[2];
#endif
    pdl_prologue();
    const int f = blockIdx.x;
    const FrameDesc fr = frame_desc(p, f);
    FrameState* st = p.state + f;
    BVX_TL((int)blockIdx.y < p.clear_rows ? 0 : 1);
    // tables already emptied by the previous call's gather (same table regions: its signature is still there)?
    const bool pre = __ldcg(p.counters + 4) == (int)p.tail_sig[0] && __ldcg(p.counters + 5) == (int)p.tail_sig[1];
    if ((int)blockIdx.y < p.clear_rows) {
        if (pre) return;
        // Clear rows: the frame's table is emptied here, right before it is used, with normal
        // stores -- whole sectors, so the lines are allocated in the L2 without a DRAM fill and
        // the atomics below find them there.  Slices are CLAIMED (not owned by block index): an
        // insert block that finds the table not ready claims slices too, so it only ever waits
        // for blocks that are already running -- no assumption about the dispatch order.
        if (fr.n > 0) {
            if (threadIdx.x == 0) s_slice = atomicAdd(&st->clear_claim, 1);
            __syncthreads();
            if (s_slice < p.clear_rows) {
                clear_slice(p, fr, s_slice, p.clear_rows);
                __threadfence();
                __syncthreads();
                if (threadIdx.x == 0) atomicAdd(&st->cleared, 1);
            }
        }
        return;
    }
    if ((long long)(blockIdx.y - p.clear_rows) * kTicket >= fr.n) return;
    const bool brk = !p.overflow_continue;
    unsigned long long* tbl = p.table + fr.tbl_off;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int toff = wid * kUnit + lane;                 // this thread's first point inside a chunk
    // ticket -> (chunk, table-ready flag); a chunk past the break cut-off is skipped
    auto take = [&](unsigned long long t, int q) {
        int c = (int)((unsigned int)t & 0x7FFFFFFFu);
        s_rdy[q] = (int)(((unsigned int)t >> 31) | (p.clear_rows == 0) | pre);   // bit 31: the table is known to be empty
        if (brk && (int)(t >> 32) > p.cap && (long long)c * kTicket < fr.n) {
            atomicMin(&st->n_eff, c * kTicket);
            c = 0x7FFFFFF;                             // skipped
        }
        s_tk[q] = c;
#if BVX_BULK_LOAD
        if (!CROP && fr.plain && p.vec4 && (long long)c * kTicket < fr.n) {
            // the ticket's points start their way into shared memory now: a barrier earlier than the loads they replace
            const int np = min(kTicket, fr.n - c * kTicket);
            bulk_load(s_pts[q], reinterpret_cast<const float4*>(plain_rows(p, fr)) + (size_t)c * kTicket, (unsigned)np * 16u, &s_bar[q]);
        }
#endif
    };
    if (threadIdx.x == 0) {
        s_extn[0] = 0; s_extn[1] = 0;
#if BVX_BULK_LOAD
        mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#endif
        take(atomicAdd(&st->ticket, 1ull), 0);
    }
    __syncthreads();
    // Chunk rows are persistent: a block keeps taking tickets of its frame until the frame is
    // exhausted or cut, so the launch carries (rows x frames) blocks, not one block per chunk.
  for (int it = 0;; ++it) {
    const int q = it & 1;
    const long long basel = (long long)s_tk[q] * kTicket;
    const int ready = s_rdy[q];
    if (basel >= fr.n) return;
    const int base = (int)basel;
    int cell[kPts];
    unsigned int h[kPts];
    if (!CROP && fr.plain && p.vec4) {
        // plain frame of float4 points: all four loads are issued before the first one is used
        // (one DRAM round trip per chunk, not four)
        float4 v[kPts];
#if BVX_BULK_LOAD
        mbar_wait(&s_bar[q], (unsigned)(it >> 1) & 1u);
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (base + toff + k * 32 < fr.n) v[k] = s_pts[q][toff + k * 32];
        }
#else
        const float4* qp = reinterpret_cast<const float4*>(plain_rows(p, fr)) + base + toff;
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (base + toff + k * 32 < fr.n) v[k] = __ldcs(qp + k * 32);   // streamed once
        }
#endif
#pragma unroll
        for (int k = 0; k < kPts; ++k)
            cell[k] = (base + toff + k * 32 < fr.n) ? cell_of(p.g, v[k].x, v[k].y, v[k].z) : -1;
    } else {
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const int i = base + toff + k * 32;
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
        // Help with whatever slices are still unclaimed, then wait for the claimed ones (their
        // blocks are running).  The first block of a frame to see all slices done says so in bit 31
        // of the ticket word; blocks that get the bit with their ticket skip all of this.  (The
        // clear stores were fenced to the L2 before `cleared` was bumped, and the table is only
        // ever touched by atomics, which act at the L2.)
        if (threadIdx.x == 0) s_slice = *(volatile int*)&st->cleared >= p.clear_rows ? -1 : 0;
        __syncthreads();
        if (s_slice != -1) {
            for (;;) {
                __syncthreads();
                if (threadIdx.x == 0) s_slice = atomicAdd(&st->clear_claim, 1);
                __syncthreads();
                if (s_slice >= p.clear_rows) break;
                clear_slice(p, fr, s_slice, p.clear_rows);
                __threadfence();
                __syncthreads();
                if (threadIdx.x == 0) atomicAdd(&st->cleared, 1);
            }
            if (threadIdx.x == 0) while (*(volatile int*)&st->cleared < p.clear_rows) { }
        }
        if (threadIdx.x == 0) {
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
    const int part = (base & (kChunk - 1)) / kTicket;           // which part of its chunk this ticket is
    const size_t lidx = (size_t)(fr.chunk_off + (base >> 10)) * kParts + part;     // its extras list
    int2* ext = p.extras + lidx * kTicket;       // (every point of the ticket adds at most one entry: itself, or the point it displaced)
    // A displaced point o (it was its cell's first point until point i arrived) becomes an extra of the
    // cell; its chunk counts one first point less and one displaced candidate more.
    auto displaced = [&](unsigned int o, unsigned int slot) {
        atomicOr(p.lost + fr.word_off + (o >> 5), 1u << (o & 31));
        atomicAdd(p.chunkstat + fr.chunk_off + (o >> 10), 0xFFFFu);
        atomicAdd(reinterpret_cast<unsigned int*>(p.unit8) + (size_t)(fr.chunk_off + (o >> 10)) * 4 + 2 + ((o >> 9) & 1u),
                  1u << (((o >> 7) & 3u) * 8));
        ext[atomicAdd(&s_extn[q], 1)] = make_int2((int)o, (int)slot);
    };
    {
        // first round: one CAS per in-range point, all four in flight
        unsigned long long r[kPts];
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            r[k] = 0;
            if (active & (1u << k))
                r[k] = atomicCAS(tbl + h[k], kEmpty64,
                                 ((unsigned long long)(unsigned)cell[k] << 32) | (unsigned)(base + toff + k * 32));
        }
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            const unsigned int bit = 1u << k;
            if (!(active & bit)) continue;
            if (r[k] == kEmpty64) {
                wonmask |= bit; ++nnew; active &= ~bit;
            } else if ((unsigned)(r[k] >> 32) == (unsigned)cell[k]) {
                if ((unsigned)r[k] > (unsigned)(base + toff + k * 32)) needmin |= bit; else active &= ~bit;
            } else {
                h[k] = (h[k] + 1) & fr.tbl_mask;
            }
        }
    }
    // Stragglers (a slot held by another cell: probe on; the cell's entry is a later point: atomicMin): a
    // couple of lanes per warp, so each lane works on ONE of its points per round -- the rounds cost a quarter
    // of a four-point round, and an atomic instruction with two active lanes costs as much as a full one.
    while (__any_sync(0xffffffffu, active)) {
        const int k = __ffs(active) - 1;                   // this lane's point of the round (-1: none left)
        const unsigned int bit = active & (0u - active);
        const unsigned int hk = k == 0 ? h[0] : (k == 1 ? h[1] : (k == 2 ? h[2] : h[3]));
        const unsigned int ck = (unsigned)(k == 0 ? cell[0] : (k == 1 ? cell[1] : (k == 2 ? cell[2] : cell[3])));
        const unsigned int i = (unsigned)(base + toff + k * 32);
        const unsigned long long mine = ((unsigned long long)ck << 32) | i;
        unsigned long long rr = 0;
        if (bit) rr = (needmin & bit) ? atomicMin(tbl + hk, mine) : atomicCAS(tbl + hk, kEmpty64, mine);
        if (bit) {
            if (needmin & bit) {
                const unsigned int o = (unsigned)rr;
                if (o > i) { wonmask |= bit; displaced(o, hk); }
                active &= ~bit;
            } else if (rr == kEmpty64) {
                wonmask |= bit; ++nnew; active &= ~bit;
            } else if ((unsigned)(rr >> 32) == ck) {
                if ((unsigned)rr > i) needmin |= bit; else active &= ~bit;
            } else {
                const unsigned int hn = (hk + 1) & fr.tbl_mask;
                if (k == 0) h[0] = hn; else if (k == 1) h[1] = hn; else if (k == 2) h[2] = hn; else h[3] = hn;
            }
        }
    }
    if (brk) {
        // publish the cells this warp claimed (per warp, as soon as it is done: measured better
        // than one add per block); later tickets see them in their ticket word
        nnew = __reduce_add_sync(0xffffffffu, nnew);
        if (nnew && lane == 0) atomicAdd(&st->ticket, (unsigned long long)nnew << 32);
    }
    // The next chunk's ticket is taken here, by the first warp when ITS rounds are over: the claimed-cell
    // count it carries is fresh, and its round trip overlaps the epilogue and the wait for the other warps.
    unsigned long long tnext = 0;
    if (threadIdx.x == 0) tnext = atomicAdd(&st->ticket, 1ull);
    {
        // candidate words of the warp's unit (one 16-byte store), their count, and the warp's own
        // extras: in range, not a candidate
        unsigned int cw[kPts], ew[kPts];
        int ncand = 0, next = 0;
#pragma unroll
        for (int k = 0; k < kPts; ++k) {
            cw[k] = __ballot_sync(0xffffffffu, (wonmask >> k) & 1u);
            ew[k] = __ballot_sync(0xffffffffu, cell[k] >= 0 && !((wonmask >> k) & 1u));
            ncand += __popc(cw[k]);
            next += __popc(ew[k]);
        }
        const int cidx = fr.chunk_off + (base >> 10);
        const int w0 = (base >> 5) + wid * kPts;                    // first word of the unit (frame-local)
        if (lane == 0) {
            if (w0 + kPts <= fr.nwords) {
                *reinterpret_cast<uint4*>(p.cand + fr.word_off + w0) = make_uint4(cw[0], cw[1], cw[2], cw[3]);
            } else {
#pragma unroll
                for (int k = 0; k < kPts; ++k)
                    if (w0 + k < fr.nwords) p.cand[fr.word_off + w0 + k] = cw[k];
            }
            p.unit8[(size_t)cidx * 16 + part * kIWarps + wid] = (unsigned char)ncand;
            if (ncand) atomicAdd(p.chunkstat + cidx, (unsigned)ncand);
        }
        if (next) {
            int eb = 0;
            if (lane == 0) eb = atomicAdd(&s_extn[q], next);
            eb = __shfl_sync(0xffffffffu, eb, 0);
            const unsigned int below = (1u << lane) - 1u;
#pragma unroll
            for (int k = 0; k < kPts; ++k) {
                if ((ew[k] >> lane) & 1u)
                    ext[eb + __popc(ew[k] & below)] = make_int2(base + toff + k * 32, (int)h[k]);
                eb += __popc(ew[k]);
            }
        }
    }
    if (threadIdx.x == 0) take(tnext, q ^ 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        p.extn[lidx] = s_extn[q];
        s_extn[q] = 0;                  // next used two iterations from now, behind the next barrier
    }
  }
}

// ---------------------------------------------------------------- scans
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

// Exclusive prefix over the frames of M (frame_voxel_offsets) by the block that finished the
// call's last frame; puts the call-level counters back.
__device__ void frames_done(const Pipe& p, int* sm)
{
    __shared__ int s_lastframe;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_lastframe = (atomicAdd(&p.counters[0], 1) == p.nframes - 1);
    __syncthreads();
    if (!s_lastframe) return;
    __threadfence();
    int carry = 0;
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
        p.counters[2] = 0;
    }
}

// The break cut-off: the point that would open voxel #cap is the (cap - cut_base)-th first
// occurrence of chunk cc.  One warp: 32 words, one per lane.
__device__ __forceinline__ int find_istar(const Pipe& p, const FrameDesc& fr, int cc, int need, int nwords)
{
    const int lane = threadIdx.x & 31;
    const int w = cc * kChunkWords + lane;
    unsigned int fw = 0;
    if (w < nwords) fw = __ldcg(p.cand + fr.word_off + w) & ~__ldcg(p.lost + fr.word_off + w);
    int tot;
    const int ex = warp_exclusive_scan(__popc(fw), lane, &tot);
    const unsigned int le = __ballot_sync(0xffffffffu, ex <= need);
    const int wl = 31 - __clz(le);                  // last word whose prefix is <= need (lane 0 always is)
    int r = 0;
    if (lane == wl) {
        unsigned int m = fw;
        for (int k = need - ex; k > 0; --k) m &= m - 1;
        r = w * 32 + (__ffs(m) - 1);
    }
    return __shfl_sync(0xffffffffu, r, wl);
}

// Frames of more than kScanSmemChunks chunks (> 4 M points): one block per frame scans the chunk
// counts into chunkbase[] and publishes the frame's results; k_mid<true> reads them from there.
__global__ void __launch_bounds__(kBlock) k_scan_big(const __grid_constant__ Pipe p)
{
    __shared__ int sm[kWarps + 1];
    __shared__ int s_cut[2];
    const int f = blockIdx.x;
    const FrameDesc fr = frame_desc(p, f);
    const int eff = min(fr.n, __ldcg(&p.state[f].n_eff));
    const int nch = (eff + kChunk - 1) / kChunk, nwords = (eff + 31) / 32;
    int* cb = p.chunkbase + fr.chunk_off;
    if (threadIdx.x == 0) s_cut[0] = -1;
    __syncthreads();
    int carry = 0;
    for (int b0 = 0; b0 < nch; b0 += kBlock) {
        const int c = b0 + threadIdx.x;
        const int v = c < nch ? (int)(__ldcg(p.chunkstat + fr.chunk_off + c) & 0xFFFFu) : 0;
        int tot;
        const int ex = carry + block_exclusive_scan(v, &tot, sm);
        if (c < nch) {
            cb[c] = ex;
            if (ex <= p.cap && p.cap < ex + v) { s_cut[0] = c; s_cut[1] = ex; }   // at most one thread
        }
        carry += tot;
    }
    __syncthreads();
    const int total = carry;
    const bool cut = !p.overflow_continue && total > p.cap;
    int istar = eff;
    if (cut && threadIdx.x < 32) istar = find_istar(p, fr, s_cut[0], p.cap - s_cut[1], nwords);
    if (threadIdx.x == 0) {
        FrameState* st = p.state + f;
        st->total = total;
        st->m = min(total, p.cap);
        st->istar = istar;
        st->n_done = eff;
        st->rows = 0;
    }
    frames_done(p, sm);
}

// ---------------------------------------------------------------- k_mid
// Voxel number of point fi if it is a first point = number of first points before it in the frame:
// chunk prefix + candidate counts of the units before its unit + bits before it in its unit,
// minus the displaced candidates before it (only chunks that have any: per-unit counts + the
// unit's own `lost` words).  -1 if fi is not a first point (it was displaced after it was recorded).
template <bool BIG>
__device__ __forceinline__ int rank_of(const Pipe& p, const FrameDesc& fr, unsigned int fi, const int* s_pre,
                                       const unsigned short* s_lost)
{
    const int c = (int)(fi >> 10), u = (int)(fi >> 7) & 7, wq = (int)(fi >> 5) & 3, b = (int)fi & 31;
    const uint4 u8 = __ldcg(reinterpret_cast<const uint4*>(p.unit8 + (size_t)(fr.chunk_off + c) * 16));
    const uint4 cw = __ldg(reinterpret_cast<const uint4*>(p.cand + fr.word_off + c * kChunkWords + u * kPts));
    int r, nlost;
    if (BIG) {
        r = __ldg(p.chunkbase + fr.chunk_off + c);
        nlost = (int)(__ldcg(p.chunkstat + fr.chunk_off + c) >> 16);
    } else {
        r = s_pre[c];
        nlost = s_lost[c];
    }
    // requested together with the two loads above (not after their results were used)
    uint4 lw = make_uint4(0u, 0u, 0u, 0u);
    if (nlost) lw = __ldcg(reinterpret_cast<const uint4*>(p.lost + fr.word_off + c * kChunkWords + u * kPts));
    // bytes 0 .. u-1 of the eight unit counts
    const unsigned int mlo = u >= 4 ? 0xFFFFFFFFu : ((1u << (8 * u)) - 1u);
    const unsigned int mhi = u <= 4 ? 0u : ((1u << (8 * (u - 4))) - 1u);
    r += (int)__vsadu4(u8.x & mlo, 0u) + (int)__vsadu4(u8.y & mhi, 0u);
    const unsigned int bit = 1u << b, below = bit - 1u;
    const unsigned int m0 = wq > 0 ? 0xFFFFFFFFu : below, m1 = wq > 1 ? 0xFFFFFFFFu : (wq == 1 ? below : 0u),
                       m2 = wq > 2 ? 0xFFFFFFFFu : (wq == 2 ? below : 0u), m3 = wq == 3 ? below : 0u;
    r += __popc(cw.x & m0) + __popc(cw.y & m1) + __popc(cw.z & m2) + __popc(cw.w & m3);
    unsigned int mine = wq == 0 ? cw.x : (wq == 1 ? cw.y : (wq == 2 ? cw.z : cw.w));
    if (nlost) {
        r -= (int)__vsadu4(u8.z & mlo, 0u) + (int)__vsadu4(u8.w & mhi, 0u);
        r -= __popc(lw.x & m0) + __popc(lw.y & m1) + __popc(lw.z & m2) + __popc(lw.w & m3);
        mine &= ~(wq == 0 ? lw.x : (wq == 1 ? lw.y : (wq == 2 ? lw.z : lw.w)));
    }
    return (mine & bit) ? r : -1;
}

// Extras {point index, table slot of its cell}: the cell's entry holds its first point; that
// point's voxel number is the extra's voxel; the extra takes its slot among the voxel's other points.
// (In break mode extras behind the cut-off may get a slot too: they are the largest indices of their
// voxel, never push a smaller one out, and the gather drops them.)  Two extras per lane at a time, so
// that their lookups and their first atomics are in flight together: the pass is a chain of dependent
// L2 / DRAM round trips, not a throughput problem.
template <bool BIG>
__device__ __forceinline__ int extra_rank(const Pipe& p, const FrameDesc& fr, int2 e, const int* s_pre,
                                          const unsigned short* s_lost)
{
    const unsigned int fi = (unsigned int)__ldcg(p.table + fr.tbl_off + (unsigned int)e.y);
    return rank_of<BIG>(p, fr, fi, s_pre, s_lost);
}

__device__ __forceinline__ void place_extra(const Pipe& p, int f, int rank, int i, int first_old, bool have_first)
{
    const int T1 = p.T - 1;
    const size_t g = (size_t)f * p.cap + rank;
    if (p.T <= kChainMaxT) {
        // sorted insertion over the voxel's T-1 slots: slot j ends up holding the (j+1)-th
        // smallest index among the voxel's extras (the atomic on slot 0 may have been issued already)
        int* b = p.bucket + g * T1;
        int cur = i;
        for (int s = 0; s < T1; ++s) {
            const int old = (s == 0 && have_first) ? first_old : atomicMin(b + s, cur);
            if (old == kEmptyIdx) break;
            cur = max(old, cur);
        }
    } else {
        const int k = atomicAdd(p.cnt + g, 1);
        if (k < T1) {
            p.bucket[g * T1 + k] = i;
        } else {
            const int o = atomicAdd(&p.counters[1], 1);
            if (o < p.ovf_cap) p.ovf[o] = make_int2((int)g, i);
        }
    }
}

template <bool BIG>
__device__ __forceinline__ void place_two(const Pipe& p, const FrameDesc& fr, int f, bool v0, int2 e0, bool v1, int2 e1,
                                          const int* s_pre, const unsigned short* s_lost)
{
    int r0 = v0 ? extra_rank<BIG>(p, fr, e0, s_pre, s_lost) : 0x7FFFFFFF;
    int r1 = v1 ? extra_rank<BIG>(p, fr, e1, s_pre, s_lost) : 0x7FFFFFFF;
    const bool chain = p.T <= kChainMaxT;
    int o0 = 0, o1 = 0;
    if (chain) {
        const int T1 = p.T - 1;
        if ((unsigned)r0 < (unsigned)p.cap) o0 = atomicMin(p.bucket + ((size_t)f * p.cap + r0) * T1, e0.x);
        if ((unsigned)r1 < (unsigned)p.cap) o1 = atomicMin(p.bucket + ((size_t)f * p.cap + r1) * T1, e1.x);
    }
    if ((unsigned)r0 < (unsigned)p.cap) place_extra(p, f, r0, e0.x, o0, chain);
    if ((unsigned)r1 < (unsigned)p.cap) place_extra(p, f, r1, e1.x, o1, chain);
}

// grid (gx + 1, B); dynamic smem: (max chunks per frame + 1) x (6 + 4 kParts) bytes (not BIG) + 16 KB.  Every block of
// frame f scans the frame's chunk counts in shared memory (<= 4096 ints: cheaper than a kernel boundary
// and a last-block chain).  Then the blocks of a frame split by role, so that the three latency chains
// run side by side instead of one after the other:
//   * the LAST block only publishes the frame's results (and, for the last frame to get there, the
//     voxel offsets of the call)
//   * blocks [0, n2): one warp per chunk: the chunk's first points get their voxel number from the
//     bitmaps and are written to vfirst; the first point of would-be voxel #cap is the break cut-off i*
//   * blocks [n2, gx): the frame's extras, spread evenly over their warps (the chunks' lists are very
//     uneven: late chunks hold most of them): voxel number of the recorded first point -> slot
// (one block per frame besides the publisher does both jobs)
#ifndef BVX_MID_BLOCKS
#define BVX_MID_BLOCKS 6   // resident blocks per SM the launch is sized for (one wave)
#endif
template <bool BIG>
__global__ void __launch_bounds__(kBlock, BVX_MID_BLOCKS) k_mid(const __grid_constant__ Pipe p, int smem_chunks, int n2)
{
    extern __shared__ int s_dyn[];
    __shared__ int sm[kWarps + 1];
    __shared__ int s_cut;
    pdl_prologue();
    BVX_TL(2);
    BVX_TL_MARK(7);
    const int f = blockIdx.y;
    const FrameDesc fr = frame_desc(p, f);
    FrameState* st = p.state + f;
    int* s_pre = s_dyn;                                  // [smem_chunks] exclusive prefix of the first-point counts
    int* s_epre = s_dyn + smem_chunks;                   // [kParts * smem_chunks + 1] exclusive prefix of the extras lists' lengths
    unsigned short* s_lost = reinterpret_cast<unsigned short*>(s_dyn + (1 + kParts) * smem_chunks + 1);
    unsigned short* s_stage = s_lost + ((smem_chunks + 1) & ~1);   // [kWarps][kChunk] first points of a chunk, per warp
    const int cap = p.cap;
    const bool brk = !p.overflow_continue;
    const int gx = gridDim.x - 1;
    const bool publisher = (int)blockIdx.x == gx;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // n2 == 0: every block does both jobs (numbering, then extras); else blocks [0, n2) number, [n2, gx) place
    const bool role2 = n2 == 0 || (int)blockIdx.x < n2, role3 = (n2 == 0 || (int)blockIdx.x >= n2) && !publisher;
    const int nw2 = (n2 == 0 ? gx : n2) * kWarps, nw3 = (n2 == 0 ? gx : gx - n2) * kWarps;
    const int w2 = blockIdx.x * kWarps + wid, w3 = (n2 == 0 ? (int)blockIdx.x : (int)blockIdx.x - n2) * kWarps + wid;
    // the bitmap words of the first chunk this warp numbers are fetched before anything else
    unsigned int pf_c = 0, pf_l = 0;
    if (!BIG && role2 && !publisher && w2 < fr.nchunks) {
        const int w = w2 * kChunkWords + lane;
        if (w < fr.nwords) { pf_c = __ldcg(p.cand + fr.word_off + w); pf_l = __ldcg(p.lost + fr.word_off + w); }
    }
    int total, eff, nch, cut_chunk;
    if (BIG) {
        eff = __ldcg(&st->n_done);
        nch = (eff + kChunk - 1) / kChunk;
        total = __ldcg(&st->total);
        cut_chunk = total > cap ? (brk ? __ldcg(&st->istar) / kChunk : nch - 1) : nch - 1;
        if (publisher) return;
    } else {
        // thread t scans chunks [t*per, (t+1)*per); the loads do not wait for n_eff
        const int per = (fr.nchunks + kBlock - 1) / kBlock;
        const int c0 = min(fr.nchunks, (int)threadIdx.x * per), c1 = min(fr.nchunks, c0 + per);
        eff = min(fr.n, __ldcg(&st->n_eff));     // prefix that went through k_insert
        int sum = 0, esum = 0;
        // the list lengths are requested together with n_eff and the counts, and the ones of parts behind the
        // inserted prefix (stale values of an earlier call) are dropped afterwards: one round trip, not two
        for (int c = c0; c < c1; ++c) {
            const unsigned int v = __ldcg(p.chunkstat + fr.chunk_off + c);
            int ne[kParts];
#pragma unroll
            for (int h = 0; h < kParts; ++h) ne[h] = __ldcg(p.extn + (size_t)(fr.chunk_off + c) * kParts + h);
            s_pre[c] = (int)(v & 0xFFFFu);
            s_lost[c] = (unsigned short)(v >> 16);
            sum += (int)(v & 0xFFFFu);
#pragma unroll
            for (int h = 0; h < kParts; ++h) {
                if (c * kChunk + h * kTicket >= eff) ne[h] = 0;
                s_epre[c * kParts + h] = ne[h];
                esum += ne[h];
            }
        }
        nch = (eff + kChunk - 1) / kChunk;
        BVX_TL_MARK(8);
        if (threadIdx.x == 0) s_cut = nch - 1;
        int etotal;
        int eex = block_exclusive_scan(esum, &etotal, sm);
        int ex = block_exclusive_scan(sum, &total, sm);
        for (int c = c0; c < c1; ++c) {
            const int v = s_pre[c];
            s_pre[c] = ex;
            if (ex <= cap && cap < ex + v) s_cut = c;   // at most one thread: the chunk that holds first point #cap
            ex += v;
#pragma unroll
            for (int h = 0; h < kParts; ++h) {
                const int ne = s_epre[c * kParts + h];
                s_epre[c * kParts + h] = eex;
                eex += ne;
            }
        }
        if (threadIdx.x == 0) s_epre[fr.nchunks * kParts] = etotal;
        __syncthreads();
        BVX_TL_MARK(4);
        cut_chunk = s_cut;
        if (publisher) {
            if (threadIdx.x == 0) {
                st->total = total;
                st->m = min(total, cap);
                if (!(brk && total > cap)) st->istar = eff;      // else: written by the warp that meets first point #cap
                st->n_done = eff;
                st->rows = 0;
            }
            frames_done(p, sm);
            BVX_TL_MARK(6);
            return;
        }
    }
    const int nwords = (eff + 31) / 32;
    const int nlim = (brk && total > cap) ? cut_chunk + 1 : nch;   // chunks that can hold anything that is kept
    if (role2) {
        for (int c = w2; c < nlim; c += nw2) {
            const int cidx = fr.chunk_off + c;
            int pre, nlost;
            if (BIG) {
                pre = __ldg(p.chunkbase + cidx);
                nlost = (int)(__ldcg(p.chunkstat + cidx) >> 16);
            } else {
                pre = s_pre[c];
                nlost = s_lost[c];
            }
            if (pre <= cap) {
                // first points of the chunk -> vfirst[voxel number]
                const int w = c * kChunkWords + lane;
                unsigned int fw = 0;
                if (w < nwords) {
                    if (!BIG && c == w2) {
                        fw = pf_c & ~pf_l;
                    } else {
                        fw = __ldcg(p.cand + fr.word_off + w);
                        if (nlost) fw &= ~__ldcg(p.lost + fr.word_off + w);
                    }
                }
                // the chunk's first points, compacted through shared memory (local index, 16 bits), so that
                // vfirst is written with coalesced stores: one scattered 4-byte store per voxel cost a
                // sector transaction each (1.3 M per step: 4.5 us of L1/L2 transaction time)
                int tot;
                int k = warp_exclusive_scan(__popc(fw), lane, &tot);
                unsigned short* stage = s_stage + wid * kChunk;
                while (fw) {
                    stage[k++] = (unsigned short)(lane * 32 + (__ffs(fw) - 1));
                    fw &= fw - 1;
                }
                __syncwarp();
                int* vf = p.vfirst + (size_t)f * cap;
                for (int j = lane; j < tot; j += 32) {
                    const int v = pre + j, i = c * kChunk + stage[j];
                    if (v < cap) vf[v] = i;
                    else if (v == cap && brk && !BIG) st->istar = i;     // the point that would open voxel #cap
                }
                __syncwarp();
            }
            if (BIG && p.T > 1) {
                // (frames of > 4 M points) the chunk's own extras, by the same warp
                for (int h = 0; h < kParts && c * kChunk + h * kTicket < eff; ++h) {
                    const int ne = __ldcg(p.extn + (size_t)cidx * kParts + h);
                    const int2* ext = p.extras + ((size_t)cidx * kParts + h) * kTicket;
                    for (int j = lane; j < ne; j += 32)
                        place_two<BIG>(p, fr, f, true, __ldcs(ext + j), false, make_int2(0, 0), s_pre, s_lost);
                }
            }
        }
    }
    BVX_TL_MARK(5);
    if (BIG || p.T <= 1 || !role3) return;
    // extras of chunks [0, nlim), flattened: item g of the frame belongs to the chunk c with
    // s_epre[c] <= g < s_epre[c + 1]; consecutive lanes take consecutive items, two per lane at a time
    const int E = s_epre[nlim * kParts];
    auto fetch = [&](int g) {
        int lo = 0, hi = nlim * kParts - 1;    // largest list l with s_epre[l] <= g
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_epre[mid] <= g) lo = mid; else hi = mid - 1;
        }
        return __ldcs(p.extras + ((size_t)fr.chunk_off * kParts + lo) * kTicket + (g - s_epre[lo]));
    };
    for (int g0 = w3 * 64; g0 < E; g0 += nw3 * 64) {
        const int ga = g0 + lane, gb = g0 + 32 + lane;
        const bool va = ga < E, vb = gb < E;
        const int2 ea = va ? fetch(ga) : make_int2(0, 0);
        const int2 eb = vb ? fetch(gb) : make_int2(0, 0);
        place_two<BIG>(p, fr, f, va, ea, vb, eb, s_pre, s_lost);
    }
}

// ---------------------------------------------------------------- k_overflow
// Large-T path: a crowded voxel's bucket holds T-1 arbitrary extras; every other
// extra is offered here and the bucket keeps the T-1 smallest indices.
__global__ void __launch_bounds__(kBlock) k_overflow(const __grid_constant__ Pipe p)
{
    pdl_prologue();
    const int n = min(((volatile int*)p.counters)[1], p.ovf_cap);
    const int T1 = p.T - 1;
    for (int j = blockIdx.x * kBlock + threadIdx.x; j < n; j += gridDim.x * kBlock) {
        const int2 it = p.ovf[j];
        int* b = p.bucket + (size_t)it.x * T1;
        int mx = 0;
        for (int k = 0; k < T1; ++k) mx = max(mx, __ldcg(b + k));
        if (it.y > mx) continue;  // entries only decrease: cannot be among the T-1 smallest
        int cur = it.y;
        for (int k = 0; k < T1; ++k) {
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

// Trailing blocks of both gathers: dense grid -> reset the entries the frame created; hashed grid with tail_clear ->
// empty one slice of the frame's table for the next call (nothing reads the table any more: k_mid is complete).
__device__ __forceinline__ void tail_block(const Pipe& p, const FrameDesc& fr, int idx)
{
    if (fr.tbl_shift < 0) reset_chunk(p, fr, idx);
    else if (idx < p.tail_clear && fr.n > 0) clear_slice(p, fr, idx, p.tail_clear);
}

// Every gather block first does its share of putting the frame's per-call structures back into
// their between-calls state (the gather is the call's last kernel): displaced-candidate bits and
// chunk counters to zero, k_insert's per-frame counters to their start values.
__device__ __forceinline__ void tidy_frame(const Pipe& p, const FrameDesc& fr, int f)
{
    FrameState* st = p.state + f;
    const int eff = st->n_done;
    const int nwords = (eff + 31) / 32, nch = (eff + kChunk - 1) / kChunk;
    const int stride = gridDim.x * kBlock;
    for (int w = blockIdx.x * kBlock + threadIdx.x; w < nwords; w += stride) p.lost[fr.word_off + w] = 0u;
    for (int c = blockIdx.x * kBlock + threadIdx.x; c < nch; c += stride) {
        p.chunkstat[fr.chunk_off + c] = 0u;
        reinterpret_cast<uint2*>(p.unit8)[(size_t)(fr.chunk_off + c) * 2 + 1] = make_uint2(0u, 0u);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->ticket = 0ull;
        st->cleared = 0;
        st->clear_claim = 0;
        st->n_eff = 0x7FFFFFFF;
        if (f == 0) {
            p.counters[1] = 0;
            p.counters[4] = p.tail_clear ? (int)p.tail_sig[0] : 0;     // what the tables look like after this call
            p.counters[5] = p.tail_clear ? (int)p.tail_sig[1] : 0;
        }
    }
}

// Coordinates of a voxel from its first point (a kept point is in range by construction); no detour over the
// linear cell index: de-linearising it costs two integer divisions per voxel.
__device__ __forceinline__ void write_coor(const Pipe& p, int f, size_t gv, float px, float py, float pz)
{
    int x, y, z;
    cell_xyz(p.g, px, py, pz, x, y, z);
    if (p.coor_cols == 4) {
        reinterpret_cast<int4*>(p.coors)[gv] = make_int4(f + p.frame_base, z, y, x);
    } else {
        int* c = p.coors + gv * 3;
        c[0] = z; c[1] = y; c[2] = x;
    }
}

// 16-byte asynchronous copy global -> shared (LDGSTS); src_bytes == 0 writes 16 zero bytes
// without reading global memory: the zero padding of the voxel rows.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes)
{
    const unsigned int d = (unsigned int)__cvta_generic_to_shared(smem_dst);
#if BVX_GATHER_PF == 128
    asm volatile("cp.async.ca.shared.global.L2::128B [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
#elif BVX_GATHER_PF == 256
    asm volatile("cp.async.ca.shared.global.L2::256B [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
#elif BVX_GATHER_PF == 64
    asm volatile("cp.async.ca.shared.global.L2::64B [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
#else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
#endif
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// Fast path: T <= 8 (extras already sorted by the atomicMin chain), 4 floats per
// point, 16-byte aligned rows.  One thread per voxel, kGatherVpt voxels per thread (tiles of 256 voxels
// blockIdx.x * kGatherVpt + u): a block is a chain of dependent round trips (indices -> rows -> stores), and
// what bounds the kernel is how many of the random 16-byte row reads are in flight per SM, so a thread
// issues the index loads of all its voxels, then the row reads of all of them, before it waits for any.
// The T rows of a voxel are copied global -> shared ASYNCHRONOUSLY (cp.async, zero fill for the empty slots),
// straight to their place in the warp's [32 voxels, T, 4] tile, so no registers are held while they are in
// flight; the tile then goes out with fully coalesced 16-byte streaming stores, and the thread reads its
// own rows back for the coordinates and the fused mean.
// grid (ceil(tiles / kGatherVpt) + reset chunks, B); dynamic smem = kGatherVpt * kWarps * 32 * T * 16 bytes.
template <int T>
__global__ void __launch_bounds__(kBlock, BVX_GATHER_BLOCKS) k_gather_small(const __grid_constant__ Pipe p, int tiles)
{
    extern __shared__ float4 smem_v[];
    constexpr int V = kGatherVpt;
    constexpr int TX = T > 1 ? T - 1 : 1;
    pdl_prologue();
    BVX_TL(3);
    const int f = blockIdx.y;
    const FrameDesc fr = frame_desc(p, f);
    const int nblk = (tiles + V - 1) / V;
    if ((int)blockIdx.x >= nblk) { tidy_frame(p, fr, f); tail_block(p, fr, blockIdx.x - nblk); return; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int tile0 = blockIdx.x * V;
    // The point indices of the voxels are fetched BEFORE the frame's voxel count is known (their addresses
    // do not depend on it; rows behind the count hold nothing that is used).
    int raw0[V];
    int rawx[V][TX];
#pragma unroll
    for (int u = 0; u < V; ++u) {
        const int v = (tile0 + u) * kBlock + threadIdx.x;
        raw0[u] = kEmptyIdx;
#pragma unroll
        for (int s = 0; s < TX; ++s) rawx[u][s] = kEmptyIdx;
        if (v < p.cap) {
            raw0[u] = __ldcs(p.vfirst + (size_t)f * p.cap + v);
            const int* gb = p.bucket + ((size_t)f * p.cap + v) * (T - 1);
            if (T == 5) {
                const int4 e = *reinterpret_cast<const int4*>(gb);
                rawx[u][0] = e.x; rawx[u][1 % TX] = e.y; rawx[u][2 % TX] = e.z; rawx[u][3 % TX] = e.w;
            } else {
#pragma unroll
                for (int s = 1; s < T; ++s) rawx[u][s - 1] = gb[s - 1];
            }
        }
    }
    const int mf = p.state[f].m;
    const int istar = p.state[f].istar;
    const size_t gbase = (size_t)p.frame_voxel_offsets[f];
    if (tile0 * kBlock >= mf) { tidy_frame(p, fr, f); return; }
    const float4* rows = reinterpret_cast<const float4*>(plain_rows(p, fr));
    int nvalid[V];
    int keep[V][TX];                  // the voxel's other points in front of the break cut-off (for idx_out)
#pragma unroll
    for (int u = 0; u < V; ++u) {
        const int v = (tile0 + u) * kBlock + threadIdx.x;
        const bool live = v < mf;
        // voxel-major float4 index lane*T + s inside the warp's tile (even T: xor swizzle against bank conflicts)
        float4* sw = smem_v + (u * kWarps + wid) * (32 * T);
        int* gb = p.bucket + ((size_t)f * p.cap + v) * (T - 1);
        // a bucket row that held any entry is put back to empty for the next call; entries behind the
        // break cut-off i* (a suffix of the row) are dropped
        if (T == 5 && live && rawx[u][0] != kEmptyIdx)
            *reinterpret_cast<int4*>(gb) = make_int4(kEmptyIdx, kEmptyIdx, kEmptyIdx, kEmptyIdx);
        int nv = 0;
#pragma unroll
        for (int s = 0; s < T; ++s) {
            int id;
            if (s == 0) {
                id = live ? raw0[u] : kEmptyIdx;
            } else {
                const int raw = live ? rawx[u][s - 1] : kEmptyIdx;
                if (T != 5 && raw != kEmptyIdx) gb[s - 1] = kEmptyIdx;
                id = raw < istar ? raw : kEmptyIdx;
                keep[u][s - 1] = id;
            }
            int q = lane * T + s;
            if ((T & 1) == 0) q ^= (q >> 3) & 7;
            const bool on = id != kEmptyIdx;
            cp_async16(sw + q, rows + (on ? id : 0), on ? 16 : 0);
            nv += on;
        }
        nvalid[u] = nv;
    }
    cp_async_wait_all();
    __syncwarp();
#pragma unroll
    for (int u = 0; u < V; ++u) {
        const int v0 = (tile0 + u) * kBlock;
        if (v0 >= mf) break;                               // (block-uniform)
        const int v = v0 + threadIdx.x;
        const bool live = v < mf;
        float4* sw = smem_v + (u * kWarps + wid) * (32 * T);
        if (p.voxels) {
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
        auto own = [&](int slot) {                            // this voxel's rows, back from the tile
            int q = lane * T + slot;
            if ((T & 1) == 0) q ^= (q >> 3) & 7;
            return sw[q];
        };
        if (p.crow) {
            // compact rows for the host form: the tile reserves a range for its kept points, each voxel
            // writes its rows (slots 0 .. nvalid-1) at its prefix inside the range
            __shared__ int s_scan[kWarps + 1];
            __shared__ int s_base;
            int tot;
            const int ex = block_exclusive_scan(nvalid[u], &tot, s_scan);
            if (threadIdx.x == 0) {
                s_base = atomicAdd(&p.counters[2], tot);
                p.tile_base[f * tiles + tile0 + u] = s_base;
                atomicAdd(&p.state[f].rows, tot);
            }
            __syncthreads();
            float4* dst = p.crow + s_base + ex;
#pragma unroll
            for (int s = 0; s < T; ++s)
                if (s < nvalid[u]) __stcs(dst + s, own(s));
            __syncthreads();                                  // s_base is rewritten by the next tile
        }
        if (!live) continue;
        const size_t gv = gbase + v;
        if (p.idx_out) {
            p.idx_out[gv * T] = raw0[u];
#pragma unroll
            for (int s = 1; s < T; ++s) p.idx_out[gv * T + s] = keep[u][s - 1] == kEmptyIdx ? -1 : keep[u][s - 1];
        }
        p.num_points[gv] = nvalid[u];
        const float4 r0 = own(0);
        write_coor(p, f, gv, r0.x, r0.y, r0.z);
        if (p.mean_out) {
            // sum over the T slots in slot order (zeros included), like the reference's features.sum(dim=1)
            float sx = __fadd_rn(0.f, r0.x), sy = __fadd_rn(0.f, r0.y), sz = __fadd_rn(0.f, r0.z), swt = __fadd_rn(0.f, r0.w);
#pragma unroll
            for (int s = 1; s < T; ++s) {
                const float4 r = own(s);
                sx = __fadd_rn(sx, r.x); sy = __fadd_rn(sy, r.y);
                sz = __fadd_rn(sz, r.z); swt = __fadd_rn(swt, r.w);
            }
            const float nf = (float)nvalid[u];
            const float m4[4] = {__fdiv_rn(sx, nf), __fdiv_rn(sy, nf), __fdiv_rn(sz, nf), __fdiv_rn(swt, nf)};
            if (p.mean_c == 4) {
                reinterpret_cast<float4*>(p.mean_out)[gv] = make_float4(m4[0], m4[1], m4[2], m4[3]);
            } else {
                for (int c = 0; c < p.mean_c; ++c) p.mean_out[gv * p.mean_c + c] = m4[c];
            }
        }
    }
    tidy_frame(p, fr, f);        // last: nothing of this block waits for it
}

// The same one-thread-per-voxel gather for frames made of SWEEPS (or transformed / permuted / cropped single
// sweeps): T <= 8, 4 input floats per point in 16-byte aligned rows, F_out = 4 or 5 (dt appended).  The rows
// cannot go global -> shared asynchronously here -- they are transformed on the way (sweep -> key frame, fixed
// operation order, no FMA) --, so a thread resolves the addresses of all its T rows (segment, permutation),
// issues the T loads back to back, then transforms and stages them in the warp's [32 voxels, T, F_out] tile,
// which goes out coalesced.  (The general k_gather below keeps one to four row reads per thread in flight and
// spends its time on index arithmetic: cfg5 x 64 frames 983 us there.)
// grid (tiles of 256 voxels + trailing blocks, B); dynamic smem = kWarps * 32 * T * F_out * 4 bytes.
template <int T>
__global__ void __launch_bounds__(kBlock) k_gather_sweeps(const __grid_constant__ Pipe p, int tiles)
{
    extern __shared__ float smem_sw[];
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = frame_desc(p, f);
    if ((int)blockIdx.x >= tiles) { tidy_frame(p, fr, f); tail_block(p, fr, blockIdx.x - tiles); return; }
    const int F = p.f_out;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int v0 = blockIdx.x * kBlock, v = v0 + threadIdx.x;
    // the voxel's point indices, fetched before the frame's voxel count is known (see k_gather_small)
    int idx[T];
    int* gb = p.bucket + ((size_t)f * p.cap + v) * (T - 1);
#pragma unroll
    for (int s = 0; s < T; ++s) idx[s] = kEmptyIdx;
    if (v < p.cap) {
        idx[0] = __ldcs(p.vfirst + (size_t)f * p.cap + v);
#pragma unroll
        for (int s = 1; s < T; ++s) idx[s] = gb[s - 1];
    }
    const int mf = p.state[f].m;
    const int istar = p.state[f].istar;
    const size_t gbase = (size_t)p.frame_voxel_offsets[f];
    if (v0 >= mf) { tidy_frame(p, fr, f); return; }
    const bool live = v < mf;
    float4 r4[T];
    const SegDesc* sg[T];
    int nvalid = 0;
#pragma unroll
    for (int s = 0; s < T; ++s) {
        int id = live ? idx[s] : kEmptyIdx;
        if (s > 0) {
            // a bucket slot that held an entry goes back to empty for the next call; entries behind the
            // break cut-off i* (a suffix of the row) are dropped
            if (id != kEmptyIdx) gb[s - 1] = kEmptyIdx;
            if (id >= istar) id = kEmptyIdx;
        }
        idx[s] = id;
        sg[s] = nullptr;
        if (id != kEmptyIdx) {
            ++nvalid;
            const float* q;
            if (fr.plain) {
                q = plain_rows(p, fr) + (size_t)id * 4;
            } else {
                sg[s] = &find_seg(p, fr, id);
                const int jj = sg[s]->perm ? __ldg(sg[s]->perm + (id - sg[s]->start)) : id - sg[s]->start;
                q = sg[s]->pts + (size_t)jj * 4;
            }
            r4[s] = __ldg(reinterpret_cast<const float4*>(q));
        }
    }
    float* sw = smem_sw + (size_t)wid * (32 * T) * F;
#pragma unroll
    for (int s = 0; s < T; ++s) {
        float* dst = sw + (size_t)(lane * T + s) * F;
        if (idx[s] != kEmptyIdx) {
            float x = r4[s].x, y = r4[s].y, z = r4[s].z;
            if (sg[s]) {
                xform(*sg[s], x, y, z);
                if (sg[s]->has_dt) dst[4] = sg[s]->dt;
            }
            dst[0] = x; dst[1] = y; dst[2] = z; dst[3] = r4[s].w;
        } else {
            for (int c = 0; c < F; ++c) dst[c] = 0.0f;
        }
    }
    __syncwarp();
    if (p.voxels) {
        const int wv0 = v0 + wid * 32;                    // first voxel of this warp
        const int wn = min(32, mf - wv0);                 // live voxels of this warp (may be <= 0)
        float* out = p.voxels + (gbase + wv0) * (size_t)T * F;
        const int n = wn * T * F;
        for (int j = lane; j < n; j += 32) __stcs(out + j, sw[j]);
    }
    tidy_frame(p, fr, f);
    if (!live) return;
    const size_t gv = gbase + v;
    if (p.idx_out) {
#pragma unroll
        for (int s = 0; s < T; ++s) p.idx_out[gv * T + s] = idx[s] == kEmptyIdx ? -1 : idx[s];
    }
    p.num_points[gv] = nvalid;
    const float* own = sw + (size_t)(lane * T) * F;
    write_coor(p, f, gv, own[0], own[1], own[2]);
    if (p.mean_out) {
        // sum over the T slots in slot order (zeros included), like the reference's features.sum(dim=1)
        const float nf = (float)nvalid;
        for (int c = 0; c < p.mean_c; ++c) {
            float sum = 0.0f;
#pragma unroll
            for (int s = 0; s < T; ++s) sum = __fadd_rn(sum, own[s * F + c]);
            p.mean_out[gv * p.mean_c + c] = __fdiv_rn(sum, nf);
        }
    }
}

// General path (any T, any feature count, sweeps): a tile of vt voxels (vt*T entries: the first
// point + T-1 bucket slots) staged through shared memory; entries are rank-sorted in shared memory
// (the large-T bucket is unordered).
// grid (tiles + reset chunks, B); dynamic smem: round4(E) + 2*round4(vt) ints + E*f_out floats, E = vt*T.
template <int kGB>
__global__ void __launch_bounds__(kBlock) k_gather(const __grid_constant__ Pipe p, int tiles, int vt)
{
    extern __shared__ int smem_i[];
    pdl_prologue();
    const int f = blockIdx.y;
    const FrameDesc fr = frame_desc(p, f);
    tidy_frame(p, fr, f);
    if ((int)blockIdx.x >= tiles) { tail_block(p, fr, blockIdx.x - tiles); return; }
    const int mf = p.state[f].m;
    const int v0 = blockIdx.x * vt;
    if (v0 >= mf) return;
    const int T = p.T, T1 = p.T - 1, F = p.f_out;
    const int nv = min(vt, mf - v0);
    const int E = nv * T;
    int* sidx = smem_i;
    int* scnt = smem_i + ((vt * T + 3) & ~3);                  // live entries per voxel
    float* srow = reinterpret_cast<float*>(scnt + 2 * ((vt + 3) & ~3));
    const size_t g0 = (size_t)f * p.cap + v0;
    int* gb = p.bucket + g0 * T1;
    const unsigned int tmagic = (unsigned int)((0x100000000ull + (unsigned)T - 1) / (unsigned)T);   // ceil(2^32 / T) (T = 1: does not fit, handled apart)
    for (int e = threadIdx.x; e < E; e += kBlock) {
        const int v = T == 1 ? e : (int)__umulhi((unsigned)e, tmagic), s = e - v * T;
        sidx[e] = s == 0 ? __ldcs(p.vfirst + g0 + v) : gb[(size_t)v * T1 + s - 1];
    }
    __syncthreads();
    // scnt[v]: live entries (the first point + the extras the bucket holds); snv[v]: those kept, i.e.
    // in front of the break cut-off i* -- they are the smallest indices of the voxel
    int* snv = scnt + ((vt + 3) & ~3);
    const int istar = p.state[f].istar;
    if (T > kChainMaxT) {
        // count + bucket path: the counter says how many extras the voxel saw; the live entries
        // are the first min(count, T-1) positions of its bucket, in arrival order (what lies
        // behind them is left over from earlier calls); the counter goes back to zero
        for (int v = threadIdx.x; v < nv; v += kBlock) {
            const int nl = 1 + min(p.cnt[g0 + v], T1);
            p.cnt[g0 + v] = 0;
            int c = 1;
            for (int k = 1; k < nl; ++k) c += (sidx[v * T + k] < istar);
            scnt[v] = nl;
            snv[v] = c;
        }
    } else {
        // chain path: slots 0 .. k-1 are filled in ascending order; they go back to empty
        for (int v = threadIdx.x; v < nv; v += kBlock) {
            int nl = 1, c = 1;
            for (int k = 1; k < T; ++k) {
                const int e = sidx[v * T + k];
                if (e != kEmptyIdx) { ++nl; gb[(size_t)v * T1 + k - 1] = kEmptyIdx; }
                c += (e < istar);
            }
            scnt[v] = nl;
            snv[v] = c;
        }
    }
    __syncthreads();
    const size_t gout = (size_t)p.frame_voxel_offsets[f] + v0;
    // Entries in groups of kGB per thread: first every entry of the group finds its slot (rank of its index in the
    // voxel) and the address of its row (segment, permutation), then all row loads of the group are issued, then the
    // rows are transformed and stored -- kGB random row reads in flight per thread instead of one (kGB = 4 for T <= 32;
    // long buckets hold few live entries per tile: kGB = 1 there, measured cfg3 x 64: 921 us against 1251).
    for (int e0 = 0; e0 < E; e0 += kGB * kBlock) {
        const float* q[kGB];
        const SegDesc* sg[kGB];
        int slot[kGB], vox[kGB];       // slot: (voxel, rank) as voxel * T + rank, -1 = nothing to fetch; vox: voxel if rank 0, else -1
#pragma unroll
        for (int k = 0; k < kGB; ++k) {
            const int e = e0 + k * kBlock + threadIdx.x;
            slot[k] = -1;
            if (e >= E) continue;
            const int v = T == 1 ? e : (int)__umulhi((unsigned)e, tmagic), s = e - v * T;     // e / T, exact for e * T < 2^32
            const int my = sidx[e];
            const int nlive = scnt[v], nvalid = snv[v];
            if (s < nlive && (s == 0 || my < istar)) {
                const int* vb = sidx + v * T;
                int r = 0;
                for (int j = 0; j < nlive; ++j) r += (vb[j] < my);    // slot = rank of the index in the voxel
                slot[k] = v * T + r;
                vox[k] = r == 0 ? v : -1;
                sg[k] = nullptr;
                if (fr.plain) {
                    q[k] = plain_rows(p, fr) + (size_t)my * p.f_in;
                } else {
                    sg[k] = &find_seg(p, fr, my);
                    const int jj = sg[k]->perm ? __ldg(sg[k]->perm + (my - sg[k]->start)) : my - sg[k]->start;
                    q[k] = sg[k]->pts + (size_t)jj * p.f_in;
                }
                if (p.idx_out) p.idx_out[(gout + v) * T + r] = my;
            }
            if (s >= nvalid) {
                float* z = srow + (size_t)(v * T + s) * F;
                for (int j = 0; j < F; ++j) z[j] = 0.0f;
                if (p.idx_out) p.idx_out[(gout + v) * T + s] = -1;
            }
            if (s == 0) p.num_points[gout + v] = nvalid;
        }
        float4 r4[kGB];
        if (p.vec4) {
#pragma unroll
            for (int k = 0; k < kGB; ++k)
                if (slot[k] >= 0) r4[k] = __ldg(reinterpret_cast<const float4*>(q[k]));
        }
#pragma unroll
        for (int k = 0; k < kGB; ++k) {
            if (slot[k] < 0) continue;
            float* dst = srow + (size_t)slot[k] * F;
            float x, y, z;
            if (p.vec4) {
                x = r4[k].x; y = r4[k].y; z = r4[k].z;
                dst[3] = r4[k].w;
            } else {
                x = __ldg(q[k]); y = __ldg(q[k] + 1); z = __ldg(q[k] + 2);
                for (int j = 3; j < p.f_in; ++j) dst[j] = __ldg(q[k] + j);
            }
            if (sg[k]) {
                xform(*sg[k], x, y, z);
                if (sg[k]->has_dt) dst[p.f_in] = sg[k]->dt;
            }
            dst[0] = x; dst[1] = y; dst[2] = z;
            if (vox[k] >= 0) write_coor(p, f, gout + vox[k], x, y, z);
        }
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
            const int nvalid = snv[v];
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

// canvas[b, c*D + z, y, x] = feats[i, c] for coors4[i] = (b, z, y, x), zeros elsewhere
// (torchplus.scatter_nd + the view of middle.py:522-525), as a GATHER: the canvas is channel-major, so a
// voxel's C values lie H*W floats apart and scattering them costs one 4-byte transaction each, after a
// separate zero fill.  Instead
//   k_bev_map    : map[b, z, y, x] = i  (one 4-byte store per voxel; the map is NOT cleared first)
//   k_bev_canvas : every thread owns 4 consecutive x of one (b, z, y) and writes them for a slice of the
//                  channels as float4: each canvas byte is written exactly once, coalesced, zeros included.
//                  A map entry v counts iff v < M and coors4[v] is this very cell -- that check is what makes
//                  stale or uninitialised map contents harmless.
// m_dev (optional): the number of voxels lives on the device (frame_voxel_offsets[B] of the call that
// produced them); m is then the capacity the launch was sized for.
__global__ void __launch_bounds__(kBlock) k_bev_map(const int4* __restrict__ coors, int m, const int* __restrict__ m_dev,
                                                    int b, int d, int ny, int nx, int* __restrict__ map)
{
    const int i = blockIdx.x * kBlock + threadIdx.x;
    if (m_dev) m = min(m, __ldg(m_dev));
    if (i >= m) return;
    const int4 q = __ldg(coors + i);
    if ((unsigned)q.x >= (unsigned)b || (unsigned)q.y >= (unsigned)d ||
        (unsigned)q.z >= (unsigned)ny || (unsigned)q.w >= (unsigned)nx)
        return;
    map[(((size_t)q.x * d + q.y) * ny + q.z) * nx + q.w] = i;
}

// VEC = 4: nx % 4 == 0, float4 stores; VEC = 1: any nx.  grid (groups / 256, channel slices).
template <int VEC>
__global__ void __launch_bounds__(kBlock) k_bev_canvas(const float* __restrict__ feats, const int4* __restrict__ coors,
                                                       int m, const int* __restrict__ m_dev, int c, int b, int d,
                                                       int ny, int nx, const int* __restrict__ map,
                                                       float* __restrict__ canvas, int ch_per)
{
    const int nxg = nx / VEC;
    const long long g = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (g >= (long long)b * d * ny * nxg) return;
    if (m_dev) m = min(m, __ldg(m_dev));
    const int xg = (int)(g % nxg);
    long long r = g / nxg;
    const int y = (int)(r % ny); r /= ny;
    const int z = (int)(r % d);
    const int bb = (int)(r / d);
    const size_t cell0 = (((size_t)bb * d + z) * ny + y) * nx + (size_t)xg * VEC;
    int v[VEC];
    if (VEC == 4) {
        const int4 mv = __ldg(reinterpret_cast<const int4*>(map + cell0));
        v[0] = mv.x; v[1 % VEC] = mv.y; v[2 % VEC] = mv.z; v[3 % VEC] = mv.w;
    } else {
        v[0] = __ldg(map + cell0);
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        bool ok = (unsigned)v[j] < (unsigned)m;
        if (ok) {
            const int4 q = __ldg(coors + v[j]);
            ok = q.x == bb && q.y == z && q.z == y && q.w == xg * VEC + j;
        }
        if (!ok) v[j] = -1;
    }
    const int ch0 = blockIdx.y * ch_per, ch1 = min(c, ch0 + ch_per);
    const size_t plane = (size_t)d * ny * nx;
    float* out = canvas + ((size_t)bb * c + ch0) * plane + ((size_t)z * ny + y) * nx + (size_t)xg * VEC;
    for (int ch = ch0; ch < ch1; ++ch, out += plane) {
        float o[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) o[j] = v[j] >= 0 ? __ldg(feats + (size_t)v[j] * c + ch) : 0.0f;
        if (VEC == 4) *reinterpret_cast<float4*>(out) = make_float4(o[0], o[1 % VEC], o[2 % VEC], o[3 % VEC]);
        else out[0] = o[0];
    }
}

// ---------------------------------------------------------------- learned VFE: the bandwidth part
// VoxelFeatureExtractor / V2 (second/pytorch/models/voxel_encoder.py:100-112, 165-176) before their first
// linear layer: points_mean = sum_t xyz / num; features = cat([features, xyz - points_mean (, |xyz|_2)]),
// for ALL T slots (the reference does not mask its input: a padded slot gets -points_mean), and the
// padding indicator mask[v, t] = t < num[v] (get_paddings_indicator, voxel_encoder.py:12-33) as floats.
// A tile of vt voxels is staged in shared memory so that both the [M, T, F] read and the [M, T, Fo] write
// are fully coalesced.  dynamic smem: vt * (T * F + 3) floats.
__global__ void __launch_bounds__(kBlock) k_vfe_augment(const float* __restrict__ voxels, const int* __restrict__ num,
                                                        int m, int T, int F, int with_distance, int vt,
                                                        float* __restrict__ out, float* __restrict__ mask)
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
    for (int k = threadIdx.x; k < nv * 3; k += kBlock) {
        const int v = k / 3, c = k - v * 3;
        float s = 0.0f;
        for (int j = 0; j < T; ++j) s = __fadd_rn(s, smem_f[(size_t)(v * T + j) * F + c]);
        smean[k] = __fdiv_rn(s, (float)num[v0 + v]);
    }
    __syncthreads();
    const int Fo = F + 3 + (with_distance ? 1 : 0);
    float* o = out + (size_t)v0 * T * Fo;
    for (int k = threadIdx.x; k < nv * T * Fo; k += kBlock) {
        const int row = k / Fo, c = k - row * Fo;            // row = v * T + t
        const float* in = smem_f + (size_t)row * F;
        float r;
        if (c < F) r = in[c];
        else if (c < F + 3) r = __fsub_rn(in[c - F], smean[(row / T) * 3 + (c - F)]);
        else r = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(in[0], in[0]), __fmul_rn(in[1], in[1])), __fmul_rn(in[2], in[2])));
        o[k] = r;
    }
    if (mask) {
        for (int k = threadIdx.x; k < nv * T; k += kBlock) {
            const int v = k / T, t = k - v * T;
            mask[(size_t)v0 * T + k] = t < num[v0 + v] ? 1.0f : 0.0f;
        }
    }
}

// The tail of the learned VFEs (voxel_encoder.py:124-129, 184-189): x *= mask; voxelwise = max over T.
// out[v, c] = max_t (x[v, t, c] * (t < num[v])); arg[v, c] = the t that won (first one on ties), for the backward.
__global__ void __launch_bounds__(kBlock) k_masked_max(const float* __restrict__ x, const int* __restrict__ num, int m, int T,
                                                       int C, float* __restrict__ out, int* __restrict__ arg)
{
    const long long k = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= (long long)m * C) return;
    const int v = (int)(k / C), c = (int)(k - (long long)v * C);
    const int n = num[v];
    const float* xv = x + (size_t)v * T * C + c;
    float best = __fmul_rn(__ldg(xv), 0 < n ? 1.0f : 0.0f);
    int bi = 0;
    for (int t = 1; t < T; ++t) {
        const float val = __fmul_rn(__ldg(xv + (size_t)t * C), t < n ? 1.0f : 0.0f);
        if (val > best) { best = val; bi = t; }
    }
    out[k] = best;
    if (arg) arg[k] = bi;
}

// ---------------------------------------------------------------- index scatter to the BEV plane
// The camera -> BEV projection of the fusion RPN (second/pytorch/models/rpn.py:957-962):
//   projected[b, :, dst_y[b, j], dst_x[b, j]] = src[b, :, src0[j], src1[j]]   for j = 0 .. P-1, zeros elsewhere,
// a Python loop over the batch of advanced-index assignments in the reference (last j wins where several pixels
// fall into one cell, as on the CPU).  Same gather structure as the voxel scatter: a [B, ny, nx] map of the
// winning j (atomicMax, so the map is filled with -1 first), then every output element is written exactly once.
__global__ void __launch_bounds__(kBlock) k_fill_i32(int* __restrict__ p, size_t n, int value)
{
    const size_t stride = (size_t)gridDim.x * kBlock;
    for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += stride) p[i] = value;
}

__global__ void __launch_bounds__(kBlock) k_idx_map(const int2* __restrict__ dst_xy, int b, int np, int ny, int nx,
                                                    int* __restrict__ map)
{
    const long long k = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (k >= (long long)b * np) return;
    const int bb = (int)(k / np), j = (int)(k - (long long)bb * np);
    const int2 q = __ldg(dst_xy + k);
    if ((unsigned)q.x >= (unsigned)nx || (unsigned)q.y >= (unsigned)ny) return;
    atomicMax(map + ((size_t)bb * ny + q.y) * nx + q.x, j);
}

template <int VEC>
__global__ void __launch_bounds__(kBlock) k_idx_canvas(const float* __restrict__ src, const int2* __restrict__ src_idx,
                                                       int c, int b, int s0, int s1, int ny, int nx,
                                                       const int* __restrict__ map, float* __restrict__ out, int ch_per)
{
    const int nxg = nx / VEC;
    const long long g = (long long)blockIdx.x * kBlock + threadIdx.x;
    if (g >= (long long)b * ny * nxg) return;
    const int xg = (int)(g % nxg);
    const long long r = g / nxg;
    const int y = (int)(r % ny), bb = (int)(r / ny);
    const size_t cell0 = ((size_t)bb * ny + y) * nx + (size_t)xg * VEC;
    int off[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int v = __ldg(map + cell0 + j);
        off[j] = -1;
        if (v >= 0) {
            int2 sp = make_int2(v / s1, v % s1);             // implicit mesh: pixel j = (j / s1, j % s1)
            if (src_idx) sp = __ldg(src_idx + v);
            if ((unsigned)sp.x < (unsigned)s0 && (unsigned)sp.y < (unsigned)s1) off[j] = sp.x * s1 + sp.y;
        }
    }
    const int ch0 = blockIdx.y * ch_per, ch1 = min(c, ch0 + ch_per);
    const size_t splane = (size_t)s0 * s1, oplane = (size_t)ny * nx;
    const float* sb = src + ((size_t)bb * c + ch0) * splane;
    float* o = out + ((size_t)bb * c + ch0) * oplane + (size_t)y * nx + (size_t)xg * VEC;
    for (int ch = ch0; ch < ch1; ++ch, sb += splane, o += oplane) {
        float val[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) val[j] = off[j] >= 0 ? __ldg(sb + off[j]) : 0.0f;
        if (VEC == 4) *reinterpret_cast<float4*>(o) = make_float4(val[0], val[1 % VEC], val[2 % VEC], val[3 % VEC]);
        else o[0] = val[0];
    }
}

}  // namespace bvx
