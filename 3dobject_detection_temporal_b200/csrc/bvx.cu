// bvx.cu -- host side of the C ABI declared in include/bvx.h.
//
// The library owns only per-handle workspace; inputs and outputs are caller
// owned device memory (the *_host forms stage through pinned memory).  There is
// no CPU fallback: without a CUDA device bvx_create fails with BVX_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "bvx.h"
#include "bvx_kernels.cuh"
#if defined(__SSE2__)
#include <emmintrin.h>
#include <xmmintrin.h>
#endif
#include <sched.h>
#include <unistd.h>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include "bvx_anchors.cuh"

using namespace bvx;

static std::atomic<long long> g_launches{0};
static thread_local std::string g_last_error;

// Fork-join pool of host threads for the host form's row assembly (see assemble_rows): the
// workers sleep on a condition variable between jobs; the calling thread takes share 0.
struct HostPool {
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_go, cv_done;
    std::function<void(int, int)> job;   // (share, shares)
    long long generation = 0;
    int pending = 0;
    bool stop = false;
    int shares = 1;

    void start(int n)
    {
        shares = std::max(1, n);
        for (int w = 1; w < shares; ++w)
            workers.emplace_back([this, w] {
                long long seen = 0;
                for (;;) {
                    std::function<void(int, int)> fn;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        cv_go.wait(lk, [&] { return stop || generation != seen; });
                        if (stop) return;
                        seen = generation;
                        fn = job;
                    }
                    fn(w, shares);
                    {
                        std::lock_guard<std::mutex> lk(mu);
                        if (--pending == 0) cv_done.notify_one();
                    }
                }
            });
    }
    void run(const std::function<void(int, int)>& fn)
    {
        if (shares > 1) {
            std::lock_guard<std::mutex> lk(mu);
            job = fn;
            pending = shares - 1;
            ++generation;
            cv_go.notify_all();
        }
        fn(0, shares);
        if (shares > 1) {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return pending == 0; });
        }
    }
    ~HostPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
            cv_go.notify_all();
        }
        for (auto& t : workers) t.join();
    }
};

struct bvx_handle {
    bvx_config cfg;
    GridParams grid;
    long long ncells;
    int device;
    // capacities
    long long table_slots;
    long long words_cap;
    long long voxel_cap;   // max_frames * max_voxels
    // device workspace
    unsigned long long* table = nullptr;
    unsigned int* cand = nullptr;
    unsigned int* lost = nullptr;
    unsigned char* unit8 = nullptr;
    unsigned int* chunkstat = nullptr;
    int2* extras = nullptr;
    int* extn = nullptr;
    int* chunkbase = nullptr;
    long long chunks_cap = 0;
    int* vfirst = nullptr;
    int* bucket = nullptr;
    int* cnt = nullptr;
    int2* ovf = nullptr;
    int* counters = nullptr;
    FrameDesc* d_frames = nullptr;
    SegDesc* d_segs = nullptr;
    FrameState* d_state = nullptr;
    int* d_offsets_scratch = nullptr;  // frame_voxel_offsets when the caller passes none
    size_t ws_bytes = 0;
    int seg_cap = 0;
    // descriptor upload
    std::vector<FrameDesc> h_frames;
    std::vector<SegDesc> h_segs;
    std::vector<char> uploaded;  // byte image of the last uploaded layout (plain frames: relative to layout_base)
    const float* layout_base = nullptr;   // plain frames of the current layout start at layout_base + FrameDesc::pts_off
    bool all_plain = false;      // every frame of the current layout is one plain segment
    int scan_smem_chunks = kScanSmemChunks;   // frames with more chunks take the k_scan_big path (BVX_SCAN_SMEM_CHUNKS: tests)
    static const int kRing = 32;   // >= sub-batches of a pipelined host call, so enqueueing them never waits for the device
    char* pin[kRing] = {};
    cudaEvent_t pin_ev[kRing];
    size_t pin_bytes = 0;
    int pin_next = 0;
    // pipelined host form: copy streams, per-sub-batch events, pinned scratch for the counts
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    std::vector<cudaEvent_t> ev_h2d, ev_done;
    int32_t* pin_offs = nullptr;
    size_t pin_offs_count = 0;
    FrameState* pin_state = nullptr;   // per-frame results of the last pipelined host call
    size_t pin_state_count = 0;
    bool pdl = getenv("BVX_NO_PDL") == nullptr;   // programmatic dependent launch between the pipeline kernels
    int max_tbl_lg = 0;                // log2(slots) of the largest table of the current layout
    bool prefix_upload = true;         // break mode: upload only the prefix of a frame the cut-off can reach
    bool host_assemble = true;         // pipelined host form: bring home point indices, assemble the padded rows on the host
    HostPool* pool = nullptr;          // host threads of the row assembly (created on first use)
    int host_threads = 0;              // 0: one per core this process may run on (at most 32)
    int* dev_idx = nullptr;            // [voxel capacity, T] point index per (voxel, slot) of the host form
    int32_t* pin_idx = nullptr;        // pinned mirror of dev_idx
    size_t pin_idx_count = 0;
    std::vector<cudaEvent_t> ev_out;   // per sub-batch: its D2H copies have landed
    float* pin_rows = nullptr;         // pinned: compact kept rows of the host form (specialised gather)
    size_t pin_rows_floats = 0;
    int* dev_tile_base = nullptr;      // [frames, tiles] start of each gather tile's rows
    int32_t* pin_tile_base = nullptr;
    size_t tile_base_count = 0;
    float cut_ratio = 0.f;             // largest (cut-off index / points) over the frames of the last call; 0 = unknown
    unsigned long long last_h2d = 0, last_d2h = 0;   // bytes moved by the last host-form call
    int last_redo = 0;                 // sub-batches of the last call that had to be redone with full frames
    // host-form staging
    void* stage_in = nullptr;
    size_t stage_in_bytes = 0;
    void* stage_out = nullptr;
    size_t stage_out_bytes = 0;
    float* dev_points = nullptr;
    size_t dev_points_bytes = 0;
    float* dev_voxels = nullptr;
    int* dev_coors = nullptr;
    int* dev_num = nullptr;
    int* dev_offs = nullptr;
    int* dev_offs_sub = nullptr;   // per-sub-batch offsets of the pipelined host form
    size_t dev_out_voxels = 0;  // voxel capacity of the dev_* outputs
    int dev_out_tf = 0;
    // optional per-kernel timing (bvx_profile_*): CUDA events around every launch
    bool prof = false;
    std::vector<cudaEvent_t> prof_pool;      // all events ever created
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof_pending;
    size_t prof_used = 0;
    double prof_ms[BVX_PROFILE_SLOTS] = {0};
    long long prof_n[BVX_PROFILE_SLOTS] = {0};
    std::string err;
};

static cudaEvent_t prof_event(bvx_handle* h)
{
    if (h->prof_used == h->prof_pool.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        h->prof_pool.push_back(e);
    }
    return h->prof_pool[h->prof_used++];
}

struct ProfScope {
    bvx_handle* h;
    int id;
    cudaStream_t st;
    cudaEvent_t a, b;
    ProfScope(bvx_handle* h_, int id_, cudaStream_t st_) : h(h_), id(id_), st(st_), a(nullptr), b(nullptr)
    {
        if (h->prof) { a = prof_event(h); b = prof_event(h); cudaEventRecord(a, st); }
    }
    ~ProfScope()
    {
        if (h->prof) { cudaEventRecord(b, st); h->prof_pending.push_back({id, {a, b}}); }
    }
};

static int fail(bvx_handle* h, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    g_last_error = buf;
    return code;
}

#define CK(h, call)                                                                          \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess)                                                               \
            return fail(h, BVX_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                                 \
    } while (0)

template <typename T>
static cudaError_t dalloc(bvx_handle* h, T** p, size_t count)
{
    const size_t bytes = count * sizeof(T);
    cudaError_t e = cudaMalloc((void**)p, bytes ? bytes : 16);
    if (e == cudaSuccess) h->ws_bytes += bytes;
    return e;
}

static int ilog2_ceil(unsigned long long v)
{
    int l = 0;
    while ((1ull << l) < v) ++l;
    return l;
}

extern "C" int bvx_abi_version(void) { return BVX_ABI_VERSION; }
extern "C" int64_t bvx_launch_count(void) { return g_launches.load(); }

extern "C" const char* bvx_last_error(const bvx_handle* h)
{
    return h ? h->err.c_str() : g_last_error.c_str();
}

extern "C" int bvx_grid_size(const float vs[3], const float rg[6], int64_t out[3])
{
    if (!vs || !rg || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    for (int j = 0; j < 3; ++j) {
        volatile float ext = rg[3 + j] - rg[j];
        volatile float q = ext / vs[j];
        if (!(vs[j] > 0.0f) || !std::isfinite(q))
            return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "voxel_size[%d] must be positive and finite", j);
        out[j] = (int64_t)rintf(q);  // half-to-even, like numpy.round
    }
    return BVX_OK;
}

extern "C" int bvx_host_alloc(void** out, size_t bytes)
{
    if (!out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 16);
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "cudaMallocHost: %s", cudaGetErrorString(e));
    return BVX_OK;
}

extern "C" int bvx_host_free(void* p)
{
    if (p) cudaFreeHost(p);
    return BVX_OK;
}

extern "C" int bvx_destroy(bvx_handle* h)
{
    if (!h) return BVX_OK;
    cudaSetDevice(h->device);
    cudaFree(h->table); cudaFree(h->cand); cudaFree(h->lost); cudaFree(h->unit8); cudaFree(h->chunkstat);
    cudaFree(h->extras); cudaFree(h->extn); cudaFree(h->chunkbase); cudaFree(h->vfirst); cudaFree(h->bucket);
    cudaFree(h->cnt); cudaFree(h->ovf);
    cudaFree(h->counters); cudaFree(h->d_frames); cudaFree(h->d_segs); cudaFree(h->d_state);
    cudaFree(h->d_offsets_scratch);
    cudaFree(h->dev_points); cudaFree(h->dev_voxels); cudaFree(h->dev_coors);
    cudaFree(h->dev_num); cudaFree(h->dev_offs); cudaFree(h->dev_offs_sub);
    for (int i = 0; i < bvx_handle::kRing; ++i) {
        if (h->pin[i]) { cudaFreeHost(h->pin[i]); cudaEventDestroy(h->pin_ev[i]); }
    }
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    for (cudaEvent_t e : h->ev_h2d) cudaEventDestroy(e);
    for (cudaEvent_t e : h->ev_done) cudaEventDestroy(e);
    for (cudaEvent_t e : h->ev_out) cudaEventDestroy(e);
    delete h->pool;
    cudaFree(h->dev_idx);
    if (h->pin_idx) cudaFreeHost(h->pin_idx);
    cudaFree(h->dev_tile_base);
    if (h->pin_tile_base) cudaFreeHost(h->pin_tile_base);
    if (h->pin_rows) cudaFreeHost(h->pin_rows);
    if (h->pin_offs) cudaFreeHost(h->pin_offs);
    if (h->pin_state) cudaFreeHost(h->pin_state);
    if (h->stage_in) cudaFreeHost(h->stage_in);
    if (h->stage_out) cudaFreeHost(h->stage_out);
    for (cudaEvent_t e : h->prof_pool) cudaEventDestroy(e);
    delete h;
    return BVX_OK;
}

extern "C" int bvx_create(const bvx_config* cfg, bvx_handle** out)
{
    if (!cfg || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fail(nullptr, BVX_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= ndev)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "device %d out of range (have %d)", cfg->device, ndev);
    if (cfg->max_num_points < 1 || cfg->max_voxels < 1 || cfg->max_points < 1 || cfg->max_frames < 1 ||
        cfg->max_point_features < 3)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT,
                    "max_num_points/max_voxels/max_points/max_frames must be >= 1 and max_point_features >= 3");
    if (cfg->overflow_mode != BVX_OVERFLOW_BREAK && cfg->overflow_mode != BVX_OVERFLOW_CONTINUE)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad overflow_mode %d", cfg->overflow_mode);
    if (cfg->grid_kind != BVX_GRID_HASHED && cfg->grid_kind != BVX_GRID_DENSE)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad grid_kind %d", cfg->grid_kind);
    int64_t gs[3];
    int rc = bvx_grid_size(cfg->voxel_size, cfg->point_cloud_range, gs);
    if (rc) return rc;
    for (int j = 0; j < 3; ++j)
        if (gs[j] < 1 || gs[j] > (1 << 24))
            return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "grid_size[%d] = %lld out of range", j, (long long)gs[j]);
    const long long ncells = (long long)gs[0] * gs[1] * gs[2];
    if (ncells >= (1ll << 31))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "grid has %lld cells; at most 2^31-1 supported", ncells);

    bvx_handle* h = new (std::nothrow) bvx_handle();
    if (!h) return fail(nullptr, BVX_ERR_CUDA, "out of host memory");
    h->cfg = *cfg;
    h->device = cfg->device;
    if (const char* e = getenv("BVX_SCAN_SMEM_CHUNKS")) h->scan_smem_chunks = std::max(1, std::min(kScanSmemChunks, atoi(e)));
    h->ncells = ncells;
    for (int j = 0; j < 3; ++j) {
        h->grid.lo[j] = cfg->point_cloud_range[j];
        h->grid.vs[j] = cfg->voxel_size[j];
        {
            // reciprocal for the divide-free fast path of cell_axis(): only if it is a normal number
            volatile float r = 1.0f / cfg->voxel_size[j];
            h->grid.rvs[j] = std::isnormal((float)r) ? (float)r : std::nanf("");
        }
        {
            // margin of the divide-free cell index: 2^-21 (grid + 1), rounded up
            const double m = ((double)gs[j] + 1.0) / 2097152.0;
            h->grid.mg[j] = std::nextafterf((float)m, INFINITY);
        }
        h->grid.gs[j] = (int)gs[j];
        h->grid.gsf[j] = (float)gs[j];
    }
    auto bail = [&](cudaError_t e, const char* what) {
        int code = fail(nullptr, BVX_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
        bvx_destroy(h);
        return code;
    };
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return bail(e, "cudaSetDevice");
    const long long np = cfg->max_points, nf = cfg->max_frames;
    if (cfg->grid_kind == BVX_GRID_DENSE) h->table_slots = ncells * nf;
    else h->table_slots = 2 * (np + np / 11 + 2) + 1024 * nf;     // per frame: next power of two >= n * 12/11 + 1
    h->words_cap = np / 32 + 6 * nf + 8;                              // per frame: rounded up to 4 words
    h->voxel_cap = (long long)cfg->max_voxels * nf;
    h->seg_cap = (int)std::max<long long>(nf, 64);
    const long long t1 = std::max(1, cfg->max_num_points - 1);        // bucket slots per voxel (the first point lives in vfirst)
    if ((e = dalloc(h, &h->table, (size_t)h->table_slots)) != cudaSuccess) return bail(e, "alloc table");
    if ((e = dalloc(h, &h->cand, (size_t)h->words_cap)) != cudaSuccess) return bail(e, "alloc cand");
    if ((e = dalloc(h, &h->lost, (size_t)h->words_cap)) != cudaSuccess) return bail(e, "alloc lost");
    if ((e = cudaMemset(h->lost, 0, (size_t)h->words_cap * 4)) != cudaSuccess) return bail(e, "memset lost");
    h->chunks_cap = np / kChunk + 2 * nf + 8;
    if ((e = dalloc(h, &h->unit8, (size_t)h->chunks_cap * 16)) != cudaSuccess) return bail(e, "alloc unit8");
    if ((e = cudaMemset(h->unit8, 0, (size_t)h->chunks_cap * 16)) != cudaSuccess) return bail(e, "memset unit8");
    if ((e = dalloc(h, &h->chunkstat, (size_t)h->chunks_cap)) != cudaSuccess) return bail(e, "alloc chunkstat");
    if ((e = cudaMemset(h->chunkstat, 0, (size_t)h->chunks_cap * 4)) != cudaSuccess) return bail(e, "memset chunkstat");
    if ((e = dalloc(h, &h->extras, (size_t)h->chunks_cap * kChunk)) != cudaSuccess) return bail(e, "alloc extras");
    if ((e = dalloc(h, &h->extn, (size_t)h->chunks_cap * kParts)) != cudaSuccess) return bail(e, "alloc extn");
    if ((e = dalloc(h, &h->chunkbase, (size_t)h->chunks_cap)) != cudaSuccess) return bail(e, "alloc chunkbase");
    if ((e = dalloc(h, &h->vfirst, (size_t)h->voxel_cap)) != cudaSuccess) return bail(e, "alloc vfirst");
    if ((e = dalloc(h, &h->bucket, (size_t)h->voxel_cap * t1)) != cudaSuccess) return bail(e, "alloc bucket");
    // between calls every bucket slot is empty (kEmptyIdx = 0x7F7F7F7F) and every voxel counter zero:
    // the gather puts back what its call used
    if ((e = cudaMemset(h->bucket, 0x7F, (size_t)h->voxel_cap * t1 * 4)) != cudaSuccess) return bail(e, "memset bucket");
    if ((e = dalloc(h, &h->cnt, (size_t)h->voxel_cap)) != cudaSuccess) return bail(e, "alloc cnt");
    if ((e = cudaMemset(h->cnt, 0, (size_t)h->voxel_cap * 4)) != cudaSuccess) return bail(e, "memset cnt");
    if (cfg->max_num_points > kChainMaxT)
        if ((e = dalloc(h, &h->ovf, (size_t)np)) != cudaSuccess) return bail(e, "alloc ovf");
    if ((e = dalloc(h, &h->counters, 8)) != cudaSuccess) return bail(e, "alloc counters");
    if ((e = dalloc(h, &h->d_frames, (size_t)nf)) != cudaSuccess) return bail(e, "alloc frames");
    if ((e = dalloc(h, &h->d_segs, (size_t)h->seg_cap)) != cudaSuccess) return bail(e, "alloc segs");
    if ((e = dalloc(h, &h->d_state, (size_t)nf)) != cudaSuccess) return bail(e, "alloc state");
    if ((e = dalloc(h, &h->d_offsets_scratch, (size_t)nf + 1)) != cudaSuccess) return bail(e, "alloc offsets");
    // one-time initialisation of what a call expects clean (hashed tables, bucket rows and voxel
    // counters are emptied by the call itself right before use)
    if (cfg->grid_kind == BVX_GRID_DENSE)
        if ((e = cudaMemset(h->table, 0xFF, (size_t)h->table_slots * 8)) != cudaSuccess) return bail(e, "memset table");
    if ((e = cudaMemset(h->counters, 0, 32)) != cudaSuccess) return bail(e, "memset counters");
    {
        FrameState z;
        memset(&z, 0, sizeof z);
        z.n_eff = 0x7FFFFFFF;
        std::vector<FrameState> init((size_t)nf, z);
        if ((e = cudaMemcpy(h->d_state, init.data(), sizeof(FrameState) * (size_t)nf, cudaMemcpyHostToDevice)) != cudaSuccess)
            return bail(e, "init state");
    }
    h->pin_bytes = sizeof(FrameDesc) * (size_t)nf + sizeof(SegDesc) * (size_t)h->seg_cap;
    for (int i = 0; i < bvx_handle::kRing; ++i) {
        if ((e = cudaMallocHost((void**)&h->pin[i], h->pin_bytes)) != cudaSuccess) return bail(e, "alloc pinned ring");
        if ((e = cudaEventCreateWithFlags(&h->pin_ev[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
    }
    if ((e = cudaDeviceSynchronize()) != cudaSuccess) return bail(e, "sync");
    *out = h;
    return BVX_OK;
}

extern "C" int bvx_workspace_bytes(const bvx_handle* h, size_t* out)
{
    if (!h || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    *out = h->ws_bytes;
    return BVX_OK;
}

// ---------------------------------------------------------------------------
// Build the frame / segment descriptors for a call and upload them if they
// differ from what the device already holds.
struct FrameSpec {
    const bvx_sweep* sweeps;  // k sweeps (host structs, device point pointers)
    int k;
    int with_dt;
    const bvx_crop* crop = nullptr;   // optional convex crop on the raw coordinates
};

// Chunk rows of k_insert = blocks of one frame in flight at a time (persistent; see run_pipeline).

static int upload_layout(bvx_handle* h, const std::vector<FrameSpec>& frames, cudaStream_t st,
                         int* max_n_out, bool* all_aligned)
{
    const int B = (int)frames.size();
    h->max_tbl_lg = 0;
    h->h_frames.assign(B, FrameDesc());
    h->h_segs.clear();
    long long wd = 0, tb = 0, ck = 0, pt = 0;
    int max_n = 0;
    bool all_plain = true;
    *all_aligned = true;
    // Plain frames are described relative to one base pointer, which travels as a kernel argument:
    // a new input buffer with the same frame sizes needs no descriptor upload.
    const float* base = nullptr;
    auto is_plain = [&](const FrameSpec& fs) {
        return fs.k == 1 && !fs.with_dt && !fs.sweeps[0].has_transform && !fs.sweeps[0].perm &&
               !(fs.crop && fs.crop->n_planes > 0);
    };
    for (int f = 0; f < B && !base; ++f)
        if (is_plain(frames[f]) && frames[f].sweeps[0].points) base = frames[f].sweeps[0].points;
    for (int f = 0; f < B; ++f) {
        FrameDesc& fd = h->h_frames[f];
        memset(&fd, 0, sizeof fd);
        fd.seg_begin = (int)h->h_segs.size();
        fd.seg_count = frames[f].k;
        const bool plain = is_plain(frames[f]);
        long long n = 0;
        for (int s = 0; s < frames[f].k; ++s) {
            const bvx_sweep& sw = frames[f].sweeps[s];
            if (sw.n < 0 || (sw.n > 0 && !sw.points))
                return fail(h, BVX_ERR_BAD_ARGUMENT, "frame %d sweep %d: bad points/n", f, s);
            SegDesc sd;
            memset(&sd, 0, sizeof sd);
            sd.pts = plain ? nullptr : sw.points;          // a plain frame is read through base + pts_off
            sd.perm = sw.perm;
            sd.n = sw.n;
            sd.start = (int)n;
            sd.has_transform = sw.has_transform ? 1 : 0;
            sd.has_dt = frames[f].with_dt;
            memcpy(sd.R, sw.R, sizeof sd.R);
            memcpy(sd.t, sw.t, sizeof sd.t);
            sd.dt = sw.dt;
            if (frames[f].crop && frames[f].crop->n_planes > 0) {
                if (frames[f].crop->n_planes > 8)
                    return fail(h, BVX_ERR_BAD_ARGUMENT, "frame %d: at most 8 crop planes", f);
                sd.n_planes = frames[f].crop->n_planes;
                memcpy(sd.planes, frames[f].crop->planes, sizeof(double) * 4 * sd.n_planes);
            }
            if (((uintptr_t)sw.points & 15) != 0) *all_aligned = false;
            h->h_segs.push_back(sd);
            n += sw.n;
        }
        if (frames[f].k == 0) {  // empty frame still needs one (empty) segment
            SegDesc sd;
            memset(&sd, 0, sizeof sd);
            h->h_segs.push_back(sd);
            fd.seg_count = 1;
        }
        if (n > (long long)kMaxFramePoints) return fail(h, BVX_ERR_BAD_ARGUMENT, "frame %d too large", f);
        fd.n = (int)n;
        fd.plain = 0;
        fd.pts_off = 0;
        if (plain) {
            fd.plain = 1;
            if (frames[f].sweeps[0].points)
                fd.pts_off = ((long long)(intptr_t)frames[f].sweeps[0].points - (long long)(intptr_t)base) / 4;
        } else if (n > 0) {
            all_plain = false;
        }
        fd.chunk_off = (int)ck;
        fd.nchunks = (int)((n + kChunk - 1) / kChunk);
        ck += fd.nchunks;
        fd.word_off = (int)wd;
        fd.nwords = (int)((n + 31) / 32);
        fd.tbl_off = (unsigned long long)tb;
        if (h->cfg.grid_kind == BVX_GRID_DENSE) {
            fd.tbl_shift = -1;
            fd.tbl_mask = 0xFFFFFFFFu;
            tb += h->ncells;
        } else {
            // load factor <= 11/12 even when every point opens a cell of its own
            int lg = std::max(10, ilog2_ceil((unsigned long long)n + (unsigned long long)n / 11 + 1));
            if ((1ll << lg) > h->ncells && h->ncells >= 1024) {
                // table never needs more slots than there are cells
                lg = std::max(10, ilog2_ceil((unsigned long long)h->ncells + 1));
            }
            h->max_tbl_lg = std::max(h->max_tbl_lg, lg);
            fd.tbl_shift = 32 - lg;
            fd.tbl_mask = (1u << lg) - 1u;
            tb += (1ll << lg);
        }
        pt += n;
        wd += (fd.nwords + 3) & ~3;
        max_n = std::max(max_n, fd.n);
    }
    if (pt > h->cfg.max_points)
        return fail(h, BVX_ERR_WORKSPACE, "%lld points exceed the handle capacity max_points=%d", pt, h->cfg.max_points);
    if (B > h->cfg.max_frames)
        return fail(h, BVX_ERR_WORKSPACE, "%d frames exceed the handle capacity max_frames=%d", B, h->cfg.max_frames);
    h->all_plain = all_plain;
    h->layout_base = base;
    if (tb > h->table_slots || wd > h->words_cap || ck > h->chunks_cap || (int)h->h_segs.size() > h->seg_cap)
        return fail(h, BVX_ERR_WORKSPACE, "layout exceeds the workspace (table %lld/%lld, words %lld/%lld, segs %zu/%d)",
                    tb, h->table_slots, wd, h->words_cap, h->h_segs.size(), h->seg_cap);
    *max_n_out = max_n;
    // Up to kInlineFrames frames travel as kernel arguments (run_pipeline copies h_frames into the
    // Pipe); plain frames never look at the segment descriptors, so such a call uploads nothing.
    if (B <= kInlineFrames && all_plain) return BVX_OK;

    const size_t fb = sizeof(FrameDesc) * (size_t)B, sb = sizeof(SegDesc) * h->h_segs.size();
    std::vector<char> img(fb + sb);
    memcpy(img.data(), h->h_frames.data(), fb);
    memcpy(img.data() + fb, h->h_segs.data(), sb);
    if (img == h->uploaded) return BVX_OK;  // device already holds this layout
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cs);
    if (cs != cudaStreamCaptureStatusNone)
        return fail(h, BVX_ERR_BAD_ARGUMENT,
                    "layout changed during stream capture: run the call once outside capture first");
    const int r = h->pin_next;
    h->pin_next = (r + 1) % bvx_handle::kRing;
    CK(h, cudaEventSynchronize(h->pin_ev[r]));
    memcpy(h->pin[r], img.data(), img.size());
    // The descriptors go to the device through a kernel that reads the pinned host buffer directly,
    // not through the copy engine: a cudaMemcpyAsync of more than a few hundred bytes queues behind
    // every bulk H2D copy already submitted (the pipelined host form has the whole batch in flight),
    // which would hold this call's kernels back until all of them are done.
    k_upload<<<1, kBlock, 0, st>>>(reinterpret_cast<const int*>(h->pin[r]), reinterpret_cast<int*>(h->d_frames),
                                   (int)(fb / 4), reinterpret_cast<const int*>(h->pin[r] + fb),
                                   reinterpret_cast<int*>(h->d_segs), (int)(sb / 4));
    g_launches += 1;
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->pin_ev[r], st));
    h->uploaded.swap(img);
    return BVX_OK;
}

struct Outputs {
    float* voxels;
    int32_t* coors;
    int32_t* num_points;
    int32_t* frame_voxel_offsets;  // [B+1] device (may be null -> scratch)
    int32_t* voxel_num;            // [1] device (single-frame form; may be null)
    float* mean;
    int mean_c;
    int coor_cols;
    int frame_base;
    int32_t* idx = nullptr;        // optional [sum M, T] point indices (host form); voxels may then be null
    float* crow = nullptr;         // optional compact kept rows + per-tile bases (host form, specialised gather)
    int* tile_base = nullptr;
};

// Pipeline kernels are launched with programmatic stream serialization (see pdl_prologue()).
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(bvx_handle* h, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = h->pdl ? 1 : 0;
    g_launches += 1;
    return cudaLaunchKernelEx(&cfg, kern, args...);
}

static int run_pipeline(bvx_handle* h, const std::vector<FrameSpec>& frames, int f_in, int f_out,
                        int max_voxels, const Outputs& o, cudaStream_t st)
{
    const int B = (int)frames.size();
    if (B < 1) return fail(h, BVX_ERR_BAD_ARGUMENT, "need at least one frame");
    if (f_in < 3) return fail(h, BVX_ERR_BAD_ARGUMENT, "points need at least 3 features, got %d", f_in);
    if (f_out > h->cfg.max_point_features)
        return fail(h, BVX_ERR_WORKSPACE, "%d output features exceed max_point_features=%d", f_out,
                    h->cfg.max_point_features);
    const int cap = max_voxels > 0 ? max_voxels : h->cfg.max_voxels;
    if (cap > h->cfg.max_voxels)
        return fail(h, BVX_ERR_WORKSPACE, "max_voxels=%d exceeds the handle capacity %d", cap, h->cfg.max_voxels);
    if ((!o.voxels && !o.idx && !o.crow && !o.mean) || !o.coors || !o.num_points)
        return fail(h, BVX_ERR_BAD_ARGUMENT, "null output pointer");
    if (((uintptr_t)o.voxels & 15) || ((uintptr_t)o.coors & 15))
        return fail(h, BVX_ERR_BAD_ARGUMENT, "voxels/coors must be 16-byte aligned");
    if (o.mean && (o.mean_c < 1 || o.mean_c > f_out))
        return fail(h, BVX_ERR_BAD_ARGUMENT, "mean_channels=%d must be in [1, %d]", o.mean_c, f_out);
    if (B > 65535) return fail(h, BVX_ERR_BAD_ARGUMENT, "at most 65535 frames per call");
    CK(h, cudaSetDevice(h->device));
    // Everything that can fail is checked before the first launch: the kernels hand their structures
    // from one to the next and the last one puts them back, so a call either runs whole or not at all.
    const int T = h->cfg.max_num_points;
    int gather_vt = 0;
    size_t gather_smem = 0;
    {
        // general gather: voxels per tile ~256 bucket entries, ~1024 for long buckets (few of their entries are live)
        int vt = std::max(1, 1024 / T);            // ~1024 entries per tile: four row reads in flight per thread
        auto tile_smem = [&](int v) {
            return (size_t)(((v * T + 3) & ~3) + 2 * ((v + 3) & ~3)) * 4 + (size_t)v * T * f_out * 4;
        };
        while (vt > 1 && tile_smem(vt) > 48 * 1024) vt >>= 1;
        gather_vt = vt;
        gather_smem = tile_smem(vt);
    }
    int max_n = 0;
    bool aligned = true;
    int rc = upload_layout(h, frames, st, &max_n, &aligned);
    if (rc) return rc;
    const bool vec4 = f_in == 4 && aligned;
    const bool small = T <= kChainMaxT && f_out == 4 && vec4 && h->all_plain;
    if (o.crow && !small)
        return fail(h, BVX_ERR_BAD_ARGUMENT, "compact rows need the specialised gather (T <= 8, 4 aligned features)");
    if (!small && gather_smem > 200 * 1024)
        return fail(h, BVX_ERR_BAD_ARGUMENT, "max_num_points * features too large for the gather tile (%zu B)", gather_smem);
    const int chunks = (max_n + kChunk - 1) / kChunk;   // chunks of the largest frame
    if (chunks > (65535 - 256) / kParts)
        return fail(h, BVX_ERR_BAD_ARGUMENT, "at most %d M points per frame", 67 / kParts);

    Pipe p;
    memset(&p, 0, sizeof p);
    p.g = h->grid;
    p.frames = h->d_frames;
    p.segs = h->d_segs;
    p.state = h->d_state;
    p.base = h->layout_base;
    p.table = h->table;
    p.cand = h->cand;
    p.lost = h->lost;
    p.unit8 = h->unit8;
    p.chunkstat = h->chunkstat;
    p.extras = h->extras;
    p.extn = h->extn;
    p.chunkbase = h->chunkbase;
    p.vfirst = h->vfirst;
    p.bucket = h->bucket;
    p.cnt = h->cnt;
    p.ovf = h->ovf;
    p.counters = h->counters;
    p.frame_voxel_offsets = o.frame_voxel_offsets ? o.frame_voxel_offsets : h->d_offsets_scratch;
    p.voxel_num = o.voxel_num;
    p.nframes = B;
    p.f_in = f_in;
    p.f_out = f_out;
    p.T = T;
    p.cap = cap;
    p.overflow_continue = h->cfg.overflow_mode == BVX_OVERFLOW_CONTINUE;
    p.coor_cols = o.coor_cols;
    p.frame_base = o.frame_base;
    p.ovf_cap = h->cfg.max_points;
    p.vec4 = vec4 ? 1 : 0;
    p.clear_rows = 0;
    p.voxels = o.voxels;
    p.idx_out = o.idx;
    p.crow = reinterpret_cast<float4*>(o.crow);
    p.tile_base = o.tile_base;
    p.coors = o.coors;
    p.num_points = o.num_points;
    p.mean_out = o.mean;
    p.mean_c = o.mean_c;
    p.inline_frames = B <= kInlineFrames ? 1 : 0;
    if (p.inline_frames) memcpy(p.inl, h->h_frames.data(), sizeof(FrameDesc) * (size_t)B);

    const bool hashed = h->cfg.grid_kind == BVX_GRID_HASHED;
    const int rchunks = hashed ? 0 : chunks;            // per-entry reset blocks (dense grid only)
    if (hashed && chunks > 0) {
        // one clear slice per 64 KB of the largest frame's table (measured: 16 KB .. 256 KB slices are
        // within 2 % of each other at 64 frames; small slices keep the single-frame latency down)
        const unsigned long long tbytes = (1ull << h->max_tbl_lg) * 8;
        p.clear_rows = (int)std::max<unsigned long long>(1, std::min<unsigned long long>(256, tbytes / (64 * 1024)));
    }
    int tail = 0;
    if (hashed && chunks > 0 && B <= 8 && !getenv("BVX_NO_TAIL_CLEAR")) {
        // small batches: the gather empties the tables for the next call (64 KB slices); signature = FNV-1a over the regions
        tail = p.clear_rows;
        unsigned long long sg = 1469598103934665603ull;
        auto mix = [&](unsigned long long v) { for (int k = 0; k < 8; ++k) { sg ^= (v >> (8 * k)) & 0xFF; sg *= 1099511628211ull; } };
        mix((unsigned long long)B);
        for (int f = 0; f < B; ++f) { mix(h->h_frames[f].tbl_off); mix(h->h_frames[f].tbl_mask); mix((unsigned long long)(h->h_frames[f].n > 0)); }
        if ((unsigned int)sg == 0 && (unsigned int)(sg >> 32) == 0) sg = 1;
        p.tail_clear = tail;
        p.tail_sig[0] = (unsigned int)sg;
        p.tail_sig[1] = (unsigned int)(sg >> 32);
    } else {
        p.tail_sig[0] = 0xFFFFFFFFu; p.tail_sig[1] = 0xFFFFFFFFu;      // never stored: counters[4..5] hold 0 after such a call
    }
    const int extra = hashed ? tail : rchunks;           // trailing blocks of the gather (table reset / clear)
    cudaError_t le = cudaSuccess;
    auto note = [&](cudaError_t e) { if (le == cudaSuccess) le = e; };
    if (chunks > 0) {
        ProfScope ps(h, 0, st);
        bool crop = false;
        for (const FrameSpec& fs : frames) crop |= fs.crop && fs.crop->n_planes > 0;
        // persistent ticket rows: 148 SMs x occ x kParts blocks in flight over all frames (occ per mode: DESIGN section 6), never
        // more rows than the largest frame has chunks
        const int tickets = (max_n + kTicket - 1) / kTicket;
        const int occ = crop ? 4 : (p.overflow_continue ? BVX_INSERT_BLOCKS_CONTINUE : BVX_INSERT_BLOCKS);
        const int rows = std::min(tickets, std::max(1, (148 * occ * kParts + B - 1) / B));
        const dim3 grid(B, p.clear_rows + rows);
        if (crop) note(launch_k(h, k_insert<true, 4>, grid, kIThreads, 0, st, p));
        else if (p.overflow_continue) note(launch_k(h, k_insert<false, BVX_INSERT_BLOCKS_CONTINUE>, grid, kIThreads, 0, st, p));
        else note(launch_k(h, k_insert<false, BVX_INSERT_BLOCKS>, grid, kIThreads, 0, st, p));
    }
    {   // always runs: an empty frame still needs its state and the offsets written
        ProfScope ps(h, 1, st);
        // one warp per chunk for the numbering; small batches get up to 4x as many blocks, so that the extras (spread
        // over all warps of a frame) are one round of dependent atomics per warp instead of several; all blocks
        // resident at once (148 SMs x BVX_MID_BLOCKS blocks), + 1 publisher per frame
        const int n2 = std::max(1, (chunks + kWarps - 1) / kWarps);
        const int gx = std::max(1, std::min(4 * n2, std::max(2, 148 * BVX_MID_BLOCKS / B - 1)));
        const int n2e = 0;
        if (chunks > h->scan_smem_chunks) {
            k_scan_big<<<B, kBlock, 0, st>>>(p);
            g_launches += 1;
            note(cudaGetLastError());
            note(launch_k(h, k_mid<true>, dim3(gx + 1, B), kBlock, 16 + (size_t)kWarps * kChunk * 2, st, p, 0, n2e));
        } else {
            const int sc = (std::max(chunks, 1) + 3) & ~3;
            note(launch_k(h, k_mid<false>, dim3(gx + 1, B), kBlock, (((size_t)sc * (6 + 4 * kParts) + 8 + 15) & ~(size_t)15) + (size_t)kWarps * kChunk * 2, st, p, sc, n2e));
        }
    }
    if (chunks > 0 && T > kChainMaxT) {
        const int ob = std::min(148 * 4, (h->cfg.max_points + kBlock - 1) / kBlock);
        ProfScope ps(h, 4, st);
        note(launch_k(h, k_overflow, dim3(ob), kBlock, 0, st, p));
    }
    if (small) {
        const int tiles = (cap + kBlock - 1) / kBlock;
        const size_t smem = (size_t)kGatherVpt * kWarps * 32 * T * 16;
        const dim3 grid((tiles + kGatherVpt - 1) / kGatherVpt + extra, B);
        if (smem > 48 * 1024) {
            switch (T) {
#define BVX_GA(TT) case TT: note(cudaFuncSetAttribute(k_gather_small<TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); break;
                BVX_GA(1) BVX_GA(2) BVX_GA(3) BVX_GA(4) BVX_GA(5) BVX_GA(6) BVX_GA(7) BVX_GA(8)
#undef BVX_GA
            }
        }
        ProfScope ps(h, 5, st);
        switch (T) {
#define BVX_GS(TT) case TT: note(launch_k(h, k_gather_small<TT>, grid, kBlock, smem, st, p, tiles)); break;
            BVX_GS(1) BVX_GS(2) BVX_GS(3) BVX_GS(4) BVX_GS(5) BVX_GS(6) BVX_GS(7) BVX_GS(8)
#undef BVX_GS
        }
    } else if (T <= kChainMaxT && vec4 && (f_out == 4 || f_out == 5) && !getenv("BVX_NO_SWEEP_GATHER")) {
        // frames made of sweeps (transform, dt, permutation, crop): one thread per voxel, T row reads in flight
        const int tiles = (cap + kBlock - 1) / kBlock;
        const size_t smem = (size_t)kBlock * T * f_out * 4;
        const dim3 grid(tiles + extra, B);
        ProfScope ps(h, 5, st);
        switch (T) {
#define BVX_GW(TT) case TT: note(launch_k(h, k_gather_sweeps<TT>, grid, kBlock, smem, st, p, tiles)); break;
            BVX_GW(1) BVX_GW(2) BVX_GW(3) BVX_GW(4) BVX_GW(5) BVX_GW(6) BVX_GW(7) BVX_GW(8)
#undef BVX_GW
        }
    } else {
        const int tiles = (cap + gather_vt - 1) / gather_vt;
        ProfScope ps(h, 5, st);
        if (T <= 32) {
            if (gather_smem > 48 * 1024)
                note(cudaFuncSetAttribute(k_gather<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gather_smem));
            note(launch_k(h, k_gather<4>, dim3(tiles + extra, B), kBlock, gather_smem, st, p, tiles, gather_vt));
        } else {
            if (gather_smem > 48 * 1024)
                note(cudaFuncSetAttribute(k_gather<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gather_smem));
            note(launch_k(h, k_gather<1>, dim3(tiles + extra, B), kBlock, gather_smem, st, p, tiles, gather_vt));
        }
    }
    if (le != cudaSuccess) return fail(h, BVX_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(le));
    CK(h, cudaGetLastError());
    return BVX_OK;
}

static bvx_sweep plain_sweep(const float* pts, int n)
{
    bvx_sweep s;
    memset(&s, 0, sizeof s);
    s.points = pts;
    s.n = n;
    return s;
}

extern "C" int bvx_generate(bvx_handle* h, const float* points, int32_t n, int32_t f, int32_t max_voxels,
                            float* voxels, int32_t* coors, int32_t* num_points, int32_t* voxel_num,
                            void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (n < 0 || (n > 0 && !points)) return fail(h, BVX_ERR_BAD_ARGUMENT, "bad points/n");
    bvx_sweep s = plain_sweep(points, n);
    std::vector<FrameSpec> fr(1, FrameSpec{&s, 1, 0});
    Outputs o{voxels, coors, num_points, nullptr, voxel_num, nullptr, 0, 3, 0};
    return run_pipeline(h, fr, f, f, max_voxels, o, (cudaStream_t)stream);
}

extern "C" int bvx_generate_sweeps(bvx_handle* h, const bvx_sweep* sweeps, int32_t k, int32_t f_in,
                                   int32_t max_voxels, float* voxels, int32_t* coors, int32_t* num_points,
                                   int32_t* voxel_num, void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (k < 1 || !sweeps) return fail(h, BVX_ERR_BAD_ARGUMENT, "need at least one sweep");
    std::vector<FrameSpec> fr(1, FrameSpec{sweeps, k, 1});
    Outputs o{voxels, coors, num_points, nullptr, voxel_num, nullptr, 0, 3, 0};
    return run_pipeline(h, fr, f_in, f_in + 1, max_voxels, o, (cudaStream_t)stream);
}

extern "C" int bvx_generate_batched(bvx_handle* h, const float* points, const int32_t* offs, int32_t b,
                                    int32_t f, int32_t max_voxels, float* voxels, int32_t* coors4,
                                    int32_t* num_points, int32_t* frame_voxel_offsets, float* mean,
                                    int32_t mean_c, void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (b < 1 || !offs) return fail(h, BVX_ERR_BAD_ARGUMENT, "need frame_offsets and b >= 1");
    if (f < 3) return fail(h, BVX_ERR_BAD_ARGUMENT, "points need at least 3 features, got %d", f);
    std::vector<bvx_sweep> sw(b);
    std::vector<FrameSpec> fr(b);
    for (int i = 0; i < b; ++i) {
        if (offs[i + 1] < offs[i] || offs[i] < 0) return fail(h, BVX_ERR_BAD_ARGUMENT, "frame_offsets must be non-decreasing");
        sw[i] = plain_sweep(points + (size_t)offs[i] * f, offs[i + 1] - offs[i]);
        if (sw[i].n > 0 && !points) return fail(h, BVX_ERR_BAD_ARGUMENT, "null points");
        fr[i] = FrameSpec{&sw[i], 1, 0};
    }
    Outputs o{voxels, coors4, num_points, frame_voxel_offsets, nullptr, mean, mean_c, 4, 0};
    return run_pipeline(h, fr, f, f, max_voxels, o, (cudaStream_t)stream);
}

extern "C" int bvx_generate_batched_sweeps(bvx_handle* h, const bvx_sweep* sweeps, const int32_t* sweep_offsets,
                                           int32_t b, int32_t f_in, int32_t append_dt, const bvx_crop* crops,
                                           int32_t max_voxels,
                                           float* voxels, int32_t* coors4, int32_t* num_points,
                                           int32_t* frame_voxel_offsets, float* mean, int32_t mean_c, void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (b < 1 || !sweep_offsets || !sweeps) return fail(h, BVX_ERR_BAD_ARGUMENT, "need sweeps, sweep_offsets and b >= 1");
    if (f_in < 3) return fail(h, BVX_ERR_BAD_ARGUMENT, "points need at least 3 features, got %d", f_in);
    if (sweep_offsets[0] != 0) return fail(h, BVX_ERR_BAD_ARGUMENT, "sweep_offsets must start at 0");
    std::vector<FrameSpec> fr(b);
    for (int i = 0; i < b; ++i) {
        const int k = sweep_offsets[i + 1] - sweep_offsets[i];
        if (k < 1) return fail(h, BVX_ERR_BAD_ARGUMENT, "frame %d has no sweep", i);
        fr[i] = FrameSpec{sweeps + sweep_offsets[i], k, append_dt ? 1 : 0, crops ? crops + i : nullptr};
    }
    Outputs o{voxels, coors4, num_points, frame_voxel_offsets, nullptr, mean, mean_c, 4, 0};
    return run_pipeline(h, fr, f_in, f_in + (append_dt ? 1 : 0), max_voxels, o, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ host forms
static bool is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

static int ensure_host_buffers(bvx_handle* h, size_t in_bytes, size_t voxels, int tf)
{
    if (in_bytes > h->dev_points_bytes) {
        cudaFree(h->dev_points);
        h->dev_points = nullptr;
        CK(h, cudaMalloc((void**)&h->dev_points, in_bytes));
        h->dev_points_bytes = in_bytes;
    }
    if (voxels > h->dev_out_voxels || tf > h->dev_out_tf) {
        cudaFree(h->dev_voxels); cudaFree(h->dev_coors); cudaFree(h->dev_num); cudaFree(h->dev_offs); cudaFree(h->dev_offs_sub);
        h->dev_voxels = nullptr; h->dev_coors = nullptr; h->dev_num = nullptr; h->dev_offs = nullptr;
        const size_t v = std::max(voxels, h->dev_out_voxels);
        const int t2 = std::max(tf, h->dev_out_tf);
        CK(h, cudaMalloc((void**)&h->dev_voxels, v * t2 * 4));
        CK(h, cudaMalloc((void**)&h->dev_coors, v * 16));
        CK(h, cudaMalloc((void**)&h->dev_num, v * 4));
        CK(h, cudaMalloc((void**)&h->dev_offs, ((size_t)h->cfg.max_frames + 1) * 4));
        CK(h, cudaMalloc((void**)&h->dev_offs_sub, ((size_t)h->cfg.max_frames + 16) * 2 * 4));
        h->dev_out_voxels = v;
        h->dev_out_tf = t2;
    }
    return BVX_OK;
}

static int stage_in(bvx_handle* h, const void* src, size_t bytes, cudaStream_t st)
{
    if (bytes == 0) return BVX_OK;
    if (is_pinned(src)) {
        CK(h, cudaMemcpyAsync(h->dev_points, src, bytes, cudaMemcpyHostToDevice, st));
        return BVX_OK;
    }
    if (bytes > h->stage_in_bytes) {
        if (h->stage_in) cudaFreeHost(h->stage_in);
        h->stage_in = nullptr;
        CK(h, cudaMallocHost(&h->stage_in, bytes));
        h->stage_in_bytes = bytes;
    }
    CK(h, cudaStreamSynchronize(st));  // staging buffer may still feed an earlier copy
    memcpy(h->stage_in, src, bytes);
    CK(h, cudaMemcpyAsync(h->dev_points, h->stage_in, bytes, cudaMemcpyHostToDevice, st));
    return BVX_OK;
}

// D2H of three result arrays; pinned destinations are written in place.
static int stage_out(bvx_handle* h, void* dst[3], const void* src[3], const size_t bytes[3], cudaStream_t st)
{
    size_t need = 0;
    bool pinned[3];
    for (int i = 0; i < 3; ++i) {
        pinned[i] = bytes[i] == 0 || is_pinned(dst[i]);
        if (!pinned[i]) need += (bytes[i] + 255) & ~(size_t)255;
    }
    if (need > h->stage_out_bytes) {
        if (h->stage_out) cudaFreeHost(h->stage_out);
        h->stage_out = nullptr;
        CK(h, cudaMallocHost(&h->stage_out, need));
        h->stage_out_bytes = need;
    }
    size_t off = 0;
    char* tmp[3] = {nullptr, nullptr, nullptr};
    for (int i = 0; i < 3; ++i) {
        if (bytes[i] == 0) continue;
        if (pinned[i]) {
            CK(h, cudaMemcpyAsync(dst[i], src[i], bytes[i], cudaMemcpyDeviceToHost, st));
        } else {
            tmp[i] = (char*)h->stage_out + off;
            off += (bytes[i] + 255) & ~(size_t)255;
            CK(h, cudaMemcpyAsync(tmp[i], src[i], bytes[i], cudaMemcpyDeviceToHost, st));
        }
    }
    CK(h, cudaStreamSynchronize(st));
    for (int i = 0; i < 3; ++i)
        if (tmp[i]) memcpy(dst[i], tmp[i], bytes[i]);
    return BVX_OK;
}

// Pinned host buffers, B >= 2 frames: the batch is cut into sub-batches and the three legs
// run on three streams -- H2D of sub-batch k+1, the kernels of sub-batch k and the D2H of
// sub-batch k-1 overlap, so the call costs about max(H2D, D2H) on the full-duplex PCIe link
// instead of H2D + kernels + D2H.  Results land in the caller's buffers exactly as in the
// unpipelined form (concatenated in frame order, global batch index in the coordinates).
// Kernels of sub-batch k (frames [f0, f1), the first pre[f] points of each) on st, then the
// D2H of its offsets and per-frame state into the pinned mirrors.
static int pipelined_sub_batch(bvx_handle* h, const int32_t* offs, const std::vector<int>& pre, int f0, int f1,
                               int k, int fs, int f, int cap, int mode, cudaStream_t st)
{
    const int T = h->cfg.max_num_points;
    const int bk = f1 - f0;
    std::vector<bvx_sweep> sw(bk);
    std::vector<FrameSpec> fr(bk);
    for (int i = 0; i < bk; ++i) {
        sw[i] = plain_sweep(h->dev_points + (size_t)offs[f0 + i] * f, pre[f0 + i]);
        fr[i] = FrameSpec{&sw[i], 1, 0};
    }
    const size_t v0 = (size_t)f0 * cap;
    Outputs o{h->dev_voxels + v0 * T * f, h->dev_coors + v0 * 4, h->dev_num + v0,
              h->dev_offs_sub + (size_t)k * (fs + 1), nullptr, nullptr, 0, 4, f0};
    if (mode == 1) {           // indices instead of rows: the host assembles the rows (assemble_rows)
        o.voxels = nullptr;
        o.idx = h->dev_idx + v0 * T;
    } else if (mode == 2) {    // compact kept rows (in the sub-batch's share of the row buffer) + tile bases
        o.crow = o.voxels;
        o.voxels = nullptr;
        o.tile_base = h->dev_tile_base + (size_t)f0 * ((cap + kBlock - 1) / kBlock);
    }
    int rc = run_pipeline(h, fr, f, f, cap, o, st);
    if (rc) return rc;
    // offsets and per-frame state go home by SM stores into the pinned mirrors (same reason as the
    // descriptor upload: a small cudaMemcpyAsync would wait for the bulk D2H copy in progress)
    k_upload<<<1, kBlock, 0, st>>>(h->dev_offs_sub + (size_t)k * (fs + 1), h->pin_offs + (size_t)k * (fs + 1), bk + 1,
                                   reinterpret_cast<const int*>(h->d_state), reinterpret_cast<int*>(h->pin_state + f0),
                                   (int)((size_t)bk * sizeof(FrameState) / 4));
    g_launches += 1;
    CK(h, cudaGetLastError());
    h->last_d2h += ((size_t)bk + 1) * 4 + (size_t)bk * sizeof(FrameState);
    return BVX_OK;
}

// Host half of the pipelined host form.  The padded [M, T, f] rows are 4-5x their information
// (cfg4: 1.2 points per voxel in T = 5 slots) and the PCIe link is what bounds the call, so only
// the frame-local point index of every kept (voxel, slot) comes home (4*T bytes per voxel instead
// of 4*T*f); the rows are copied here from the caller's own points, zero padded, by the handle's
// host threads, while the copies and kernels of the next sub-batch run.  Every row of the
// reference's voxels is a verbatim copy of an input row, so this is bit-exact by construction.
// Voxels [run0, run0 + m) are those of frames [f0, f1); fvo = global voxel offsets per frame.
static void assemble_rows(HostPool* pool, const float* points, const int32_t* offs, const int32_t* fvo, int f0, int f1,
                          size_t run0, size_t m, int T, int f, float* voxels, const int32_t* num_points,
                          const int32_t* idx)
{
    if (m == 0) return;
    pool->run([=](int share, int shares) {
        const size_t per = (m + shares - 1) / shares;
        const size_t lo = run0 + std::min(m, per * share), hi = run0 + std::min(m, per * (share + 1));
        int fi = f0;
        for (size_t gv = lo; gv < hi; ++gv) {
            while (fi + 1 < f1 && (size_t)fvo[fi + 1] <= gv) ++fi;
            const float* src = points + (size_t)offs[fi] * f;
            const int n = num_points[gv];
            const int32_t* id = idx + gv * T;
            float* dst = voxels + gv * (size_t)T * f;
#if defined(__SSE2__)
            if (f == 4 && ((((uintptr_t)voxels) | ((uintptr_t)points)) & 15) == 0) {
                // 16-byte rows: streaming stores (the output is written once and not read here, so no
                // read-for-ownership of its lines), and the first row of a voxel a few voxels ahead is
                // prefetched -- the row reads are random accesses into the caller's points
                constexpr size_t kAhead = 24;
                if (gv + kAhead < hi) {
                    int fa = fi;
                    while (fa + 1 < f1 && (size_t)fvo[fa + 1] <= gv + kAhead) ++fa;
                    const int ia = idx[(gv + kAhead) * T];
                    if (ia >= 0) _mm_prefetch((const char*)(points + ((size_t)offs[fa] + ia) * 4), _MM_HINT_T0);
                }
                const __m128 zero = _mm_setzero_ps();
                for (int t = 0; t < n; ++t) _mm_stream_ps(dst + 4 * t, _mm_load_ps(src + 4 * (size_t)id[t]));
                for (int t = n; t < T; ++t) _mm_stream_ps(dst + 4 * t, zero);
                continue;
            }
#endif
            for (int t = 0; t < n; ++t) memcpy(dst + (size_t)t * f, src + (size_t)id[t] * f, (size_t)f * 4);
            if (n < T) memset(dst + (size_t)n * f, 0, (size_t)(T - n) * f * 4);
        }
#if defined(__SSE2__)
        _mm_sfence();
#endif
    });
}

// Same, for the specialised gather (T <= 8, 4 features): the kept rows themselves come home,
// compact -- 16 bytes per kept point, tile after tile of 256 voxels, voxel after voxel -- so the
// host reads them front to back (no random access into the points) and only pads.
static void assemble_compact(HostPool* pool, const int32_t* fvo, int f0, int f1, int tiles, const int32_t* tile_base,
                             const float* rows, int T, float* voxels, const int32_t* num_points)
{
    const int ntile = (f1 - f0) * tiles;
    if (ntile <= 0) return;
    pool->run([=](int share, int shares) {
        const int per = (ntile + shares - 1) / shares;
        const int lo = std::min(ntile, per * share), hi = std::min(ntile, per * (share + 1));
        for (int j = lo; j < hi; ++j) {
            const int i = f0 + j / tiles, t = j % tiles;
            const size_t g0 = (size_t)fvo[i] + (size_t)t * kBlock;
            const size_t g1 = std::min<size_t>((size_t)fvo[i + 1], g0 + kBlock);
            if (g0 >= g1) continue;
            const float* src = rows + (size_t)tile_base[(size_t)(i - f0) * tiles + t] * 4;
            for (size_t gv = g0; gv < g1; ++gv) {
                const int n = num_points[gv];
                float* dst = voxels + gv * (size_t)T * 4;
#if defined(__SSE2__)
                const __m128 zero = _mm_setzero_ps();
                for (int s = 0; s < n; ++s) _mm_stream_ps(dst + 4 * s, _mm_load_ps(src + 4 * s));
                for (int s = n; s < T; ++s) _mm_stream_ps(dst + 4 * s, zero);
#else
                memcpy(dst, src, (size_t)n * 16);
                memset(dst + (size_t)n * 4, 0, (size_t)(T - n) * 16);
#endif
                src += (size_t)n * 4;
            }
        }
#if defined(__SSE2__)
        _mm_sfence();
#endif
    });
}

// The host half on its own (host memory only, no device needed): what the pipelined host form runs
// per sub-batch, exported so that it can be used -- and tested -- without a GPU.
extern "C" int bvx_host_assemble_rows(const float* points, const int32_t* frame_offsets, const int32_t* frame_voxel_offsets,
                                      int32_t b, int32_t t, int32_t f, const int32_t* num_points, const int32_t* idx,
                                      float* voxels, int32_t threads)
{
    if (b < 1 || t < 1 || f < 1 || !frame_offsets || !frame_voxel_offsets || !num_points || !idx || !voxels)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad argument");
    HostPool pool;
    pool.start(std::min(32, std::max(1, (int)threads)));
    assemble_rows(&pool, points, frame_offsets, frame_voxel_offsets, 0, b, 0, (size_t)frame_voxel_offsets[b], t, f, voxels,
                  num_points, idx);
    return BVX_OK;
}

extern "C" int bvx_host_assemble_compact(const float* rows, const int32_t* tile_base, const int32_t* frame_voxel_offsets,
                                         int32_t b, int32_t max_voxels, int32_t t, const int32_t* num_points,
                                         float* voxels, int32_t threads)
{
    if (b < 1 || t < 1 || max_voxels < 1 || !tile_base || !frame_voxel_offsets || !num_points || !voxels ||
        ((uintptr_t)rows & 15) || ((uintptr_t)voxels & 15))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad argument (rows and voxels must be 16-byte aligned)");
    HostPool pool;
    pool.start(std::min(32, std::max(1, (int)threads)));
    assemble_compact(&pool, frame_voxel_offsets, 0, b, (max_voxels + kBlock - 1) / kBlock, tile_base, rows, t, voxels,
                     num_points);
    return BVX_OK;
}

static int generate_batched_pipelined(bvx_handle* h, const float* points, const int32_t* offs, int b, int f,
                                      int cap, float* voxels, int32_t* coors4, int32_t* num_points,
                                      int32_t* fvo_host, cudaStream_t st)
{
    const int T = h->cfg.max_num_points;
    const auto t_entry = std::chrono::steady_clock::now();
    // sub-batches: 8 when the rows come home by DMA; 16 with host assembly, where the first and the last
    // sub-batch are not overlapped with anything (measured cfg4: 4 / 8 / 12 / 16 -> 3.16 / 2.71 / 2.58 / 2.52 ms)
    const int nsub = std::min(b, h->host_assemble ? 16 : 8);
    const int fs = (b + nsub - 1) / nsub;                 // frames per sub-batch
    const int ns = (b + fs - 1) / fs;
    if (!h->s_h2d) CK(h, cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
    if (!h->s_d2h) CK(h, cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
    while ((int)h->ev_h2d.size() < ns) {
        cudaEvent_t e;
        CK(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_h2d.push_back(e);
    }
    while ((int)h->ev_done.size() < ns + 1) {
        cudaEvent_t e;
        CK(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_done.push_back(e);
    }
    const bool assemble = h->host_assemble;
    // specialised gather (T <= 8, 4 features, 16-byte aligned rows): the kept rows come home compact;
    // otherwise the point indices do and the rows are read from the caller's points
    const bool compact = assemble && T <= kChainMaxT && f == 4 && ((uintptr_t)voxels & 15) == 0;
    const int mode = compact ? 2 : (assemble ? 1 : 0);
    const int tiles = (cap + kBlock - 1) / kBlock;
    if (assemble) {
        if (!h->pool) {
            h->pool = new HostPool();
            int nt = h->host_threads;
            if (nt <= 0) {
                // the cores this process may run on (a rank bound to its GPU's NUMA node sees only those)
                cpu_set_t set;
                CPU_ZERO(&set);
                nt = sched_getaffinity(0, sizeof set, &set) == 0 ? CPU_COUNT(&set)
                                                                  : (int)std::thread::hardware_concurrency();
            }
            h->pool->start(std::min(32, std::max(1, nt)));
        }
        while ((int)h->ev_out.size() < ns) {
            cudaEvent_t e;
            CK(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            h->ev_out.push_back(e);
        }
        const size_t cnt = (size_t)b * cap * T;
        if (compact) {
            // kept rows of the call: at most one per point and at most T per voxel
            const size_t rf = std::min((size_t)offs[b], cnt) * 4;
            if (rf > h->pin_rows_floats) {
                if (h->pin_rows) cudaFreeHost(h->pin_rows);
                h->pin_rows = nullptr;
                h->pin_rows_floats = 0;
                CK(h, cudaMallocHost((void**)&h->pin_rows, rf * 4));
                h->pin_rows_floats = rf;
            }
            const size_t tb = (size_t)b * tiles;
            if (tb > h->tile_base_count) {
                if (h->pin_tile_base) cudaFreeHost(h->pin_tile_base);
                h->pin_tile_base = nullptr;
                cudaFree(h->dev_tile_base);
                h->dev_tile_base = nullptr;
                h->tile_base_count = 0;
                CK(h, cudaMallocHost((void**)&h->pin_tile_base, tb * 4));
                CK(h, cudaMalloc((void**)&h->dev_tile_base, tb * 4));
                h->tile_base_count = tb;
            }
        } else if (cnt > h->pin_idx_count) {
            if (h->pin_idx) cudaFreeHost(h->pin_idx);
            h->pin_idx = nullptr;
            cudaFree(h->dev_idx);
            h->dev_idx = nullptr;
            h->pin_idx_count = 0;
            CK(h, cudaMallocHost((void**)&h->pin_idx, cnt * 4));
            CK(h, cudaMalloc((void**)&h->dev_idx, cnt * 4));
            h->pin_idx_count = cnt;
        }
    }
    const size_t need = (size_t)ns * (fs + 1);
    if (need > h->pin_offs_count) {
        if (h->pin_offs) cudaFreeHost(h->pin_offs);
        h->pin_offs = nullptr;
        CK(h, cudaMallocHost((void**)&h->pin_offs, need * 4));
        h->pin_offs_count = need;
    }
    if ((size_t)b > h->pin_state_count) {
        if (h->pin_state) cudaFreeHost(h->pin_state);
        h->pin_state = nullptr;
        CK(h, cudaMallocHost((void**)&h->pin_state, (size_t)b * sizeof(FrameState)));
        h->pin_state_count = (size_t)b;
    }
    // Break mode: nothing after the cut-off point i* of a frame can reach the output (the
    // reference's loop stops there), so only a prefix of each frame is uploaded -- as long as the
    // last call's cut-offs say where i* is likely to be (x 1.25 + one chunk).  The frame's own
    // result says whether the prefix was enough (more distinct cells than max_voxels inside it);
    // if not, the rest of the frame follows and the sub-batch is redone.  Results are exact either way.
    const bool brk = h->cfg.overflow_mode == BVX_OVERFLOW_BREAK;
    const bool spec = brk && h->prefix_upload && h->cut_ratio > 0.f && h->cut_ratio * 1.25f < 0.9f;
    std::vector<int> pre(b);
    for (int i = 0; i < b; ++i) {
        const int n = offs[i + 1] - offs[i];
        pre[i] = n;
        if (spec) {
            const long long want = ((long long)std::ceil((double)n * h->cut_ratio * 1.25) / kChunk + 2) * kChunk;
            pre[i] = (int)std::min<long long>(n, want);
        }
    }
    // the copy streams start after whatever the caller queued on its stream
    CK(h, cudaEventRecord(h->ev_done[ns], st));
    CK(h, cudaStreamWaitEvent(h->s_h2d, h->ev_done[ns], 0));
    // leg 1: every H2D copy, back to back on the copy-in stream
    for (int k = 0; k < ns; ++k) {
        const int f0 = k * fs, f1 = std::min(b, f0 + fs);
        if (!spec) {
            const size_t p0 = (size_t)offs[f0], p1 = (size_t)offs[f1];
            if (p1 > p0) {
                CK(h, cudaMemcpyAsync(h->dev_points + p0 * f, points + p0 * f, (p1 - p0) * f * 4,
                                      cudaMemcpyHostToDevice, h->s_h2d));
                h->last_h2d += (p1 - p0) * f * 4;
            }
        } else {
            for (int i = f0; i < f1; ++i) {
                if (pre[i] == 0) continue;
                const size_t p0 = (size_t)offs[i];
                CK(h, cudaMemcpyAsync(h->dev_points + p0 * f, points + p0 * f, (size_t)pre[i] * f * 4,
                                      cudaMemcpyHostToDevice, h->s_h2d));
                h->last_h2d += (size_t)pre[i] * f * 4;
            }
        }
        CK(h, cudaEventRecord(h->ev_h2d[k], h->s_h2d));
    }
    // (Letting the DMA engine bring the rows of every second / third / fourth sub-batch while the threads
    // assemble the others measured slower than assembling all of them: 3.5 / 3.0 / 2.85 vs 2.87 ms.)
    // leg 2: the kernels of each sub-batch on the caller's stream, outputs at capacity stride
    for (int k = 0; k < ns; ++k) {
        const int f0 = k * fs, f1 = std::min(b, f0 + fs);
        CK(h, cudaStreamWaitEvent(st, h->ev_h2d[k], 0));
        int rc = pipelined_sub_batch(h, offs, pre, f0, f1, k, fs, f, cap, mode, st);
        if (rc) return rc;
        CK(h, cudaEventRecord(h->ev_done[k], st));
    }
    // leg 3: as each sub-batch finishes, its M rows go home on the copy-out stream
    const bool trace = getenv("BVX_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::micro>(b - a).count(); };
    double t_done = 0, t_out = 0, t_asm = 0;
    const auto t_leg3 = now();
    std::vector<size_t> sub_run(ns, 0), sub_m(ns, 0), sub_rows(ns, 0);
    size_t run = 0, rows_run = 0;
    auto assemble_sub = [&](int j) -> int {
        { const auto a = now(); CK(h, cudaEventSynchronize(h->ev_out[j])); t_out += us(a, now()); }
        const int g0 = j * fs, g1 = std::min(b, g0 + fs);
        const auto a = now();
        if (sub_m[j] == 0) return BVX_OK;
        if (mode == 2)
            assemble_compact(h->pool, fvo_host, g0, g1, tiles, h->pin_tile_base + (size_t)g0 * tiles,
                             h->pin_rows + sub_rows[j] * 4, T, voxels, num_points);
        else
            assemble_rows(h->pool, points, offs, fvo_host, g0, g1, sub_run[j], sub_m[j], T, f, voxels, num_points,
                          h->pin_idx);
        t_asm += us(a, now());
        return BVX_OK;
    };
    fvo_host[0] = 0;
    for (int k = 0; k < ns; ++k) {
        const int f0 = k * fs, f1 = std::min(b, f0 + fs), bk = f1 - f0;
        { const auto a = now(); CK(h, cudaEventSynchronize(h->ev_done[k])); t_done += us(a, now()); }
        bool redo = false;
        for (int i = f0; i < f1; ++i)
            if (pre[i] < offs[i + 1] - offs[i] && !(h->pin_state[i].total > cap)) redo = true;
        if (redo) {
            // a prefix held no cut-off: bring the rest of the short frames and run the sub-batch again
            for (int i = f0; i < f1; ++i) {
                const int n = offs[i + 1] - offs[i];
                if (pre[i] == n) continue;
                const size_t p0 = (size_t)offs[i] + pre[i];
                CK(h, cudaMemcpyAsync(h->dev_points + p0 * f, points + p0 * f, (size_t)(n - pre[i]) * f * 4,
                                      cudaMemcpyHostToDevice, st));
                h->last_h2d += (size_t)(n - pre[i]) * f * 4;
                pre[i] = n;
            }
            int rc = pipelined_sub_batch(h, offs, pre, f0, f1, k, fs, f, cap, mode, st);
            if (rc) return rc;
            CK(h, cudaStreamSynchronize(st));
            h->last_redo += 1;
        }
        const int32_t* lo = h->pin_offs + (size_t)k * (fs + 1);
        const size_t m = (size_t)lo[bk], v0 = (size_t)f0 * cap;
        for (int i = 1; i <= bk; ++i) fvo_host[f0 + i] = (int32_t)(run + (size_t)lo[i]);
        size_t rows_k = 0;
        if (compact)
            for (int i = f0; i < f1; ++i) rows_k += (size_t)h->pin_state[i].rows;
        if (m > 0) {
            if (mode == 2) {
                CK(h, cudaMemcpyAsync(h->pin_rows + rows_run * 4, h->dev_voxels + v0 * T * f, rows_k * 16,
                                      cudaMemcpyDeviceToHost, h->s_d2h));
                CK(h, cudaMemcpyAsync(h->pin_tile_base + (size_t)f0 * tiles, h->dev_tile_base + (size_t)f0 * tiles,
                                      (size_t)bk * tiles * 4, cudaMemcpyDeviceToHost, h->s_d2h));
                h->last_d2h += rows_k * 16 + (size_t)bk * tiles * 4;
            } else if (mode == 1) {
                CK(h, cudaMemcpyAsync(h->pin_idx + run * T, h->dev_idx + v0 * T, m * T * 4, cudaMemcpyDeviceToHost,
                                      h->s_d2h));
                h->last_d2h += m * (size_t)T * 4;
            } else {
                CK(h, cudaMemcpyAsync(voxels + run * T * f, h->dev_voxels + v0 * T * f, m * T * f * 4,
                                      cudaMemcpyDeviceToHost, h->s_d2h));
                h->last_d2h += m * (size_t)T * f * 4;
            }
            CK(h, cudaMemcpyAsync(coors4 + run * 4, h->dev_coors + v0 * 4, m * 16, cudaMemcpyDeviceToHost, h->s_d2h));
            CK(h, cudaMemcpyAsync(num_points + run, h->dev_num + v0, m * 4, cudaMemcpyDeviceToHost, h->s_d2h));
            h->last_d2h += m * (16 + 4);
        }
        sub_run[k] = run; sub_m[k] = m; sub_rows[k] = rows_run;
        if (assemble) {
            // the previous sub-batch is assembled while this one's copies are in flight
            CK(h, cudaEventRecord(h->ev_out[k], h->s_d2h));
            if (k > 0) {
                const int rc = assemble_sub(k - 1);
                if (rc) return rc;
            }
        }
        run += m;
        rows_run += rows_k;
    }
    if (assemble) {
        const int rc = assemble_sub(ns - 1);
        if (rc) return rc;
    }
    if (trace)
        fprintf(stderr, "[bvx] pid %d mode %d: enqueue %.0f us, leg3 %.0f us: wait kernels %.0f, wait copies %.0f, assemble %.0f (%d threads)\n",
                (int)getpid(), mode, us(t_entry, t_leg3), us(t_leg3, now()), t_done, t_out, t_asm, h->pool ? h->pool->shares : 0);
    // where the cut-offs were this time: the next call's prefix
    if (brk) {
        float r = 0.f;
        bool any = false;
        for (int i = 0; i < b; ++i) {
            const int n = offs[i + 1] - offs[i];
            if (n == 0) continue;
            any = true;
            const FrameState& s = h->pin_state[i];
            r = std::max(r, s.total > cap ? (float)(s.istar + 1) / (float)n : 1.0f);
        }
        if (any) h->cut_ratio = r;
    }
    CK(h, cudaStreamSynchronize(h->s_d2h));
    return BVX_OK;
}

extern "C" int bvx_generate_batched_host(bvx_handle* h, const float* points, const int32_t* offs, int32_t b,
                                         int32_t f, int32_t max_voxels, float* voxels, int32_t* coors4,
                                         int32_t* num_points, int32_t* fvo_host, void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (b < 1 || !offs || !fvo_host) return fail(h, BVX_ERR_BAD_ARGUMENT, "need frame_offsets, b >= 1 and an offsets output");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->device));
    const int cap = max_voxels > 0 ? max_voxels : h->cfg.max_voxels;
    const int T = h->cfg.max_num_points;
    const size_t n = (size_t)offs[b];
    int rc = ensure_host_buffers(h, n * f * 4, (size_t)cap * b, T * f);
    if (rc) return rc;
    h->last_h2d = 0; h->last_d2h = 0; h->last_redo = 0;
    if (b >= 2 && n > 0 && is_pinned(points) && is_pinned(voxels) && is_pinned(coors4) && is_pinned(num_points))
        return generate_batched_pipelined(h, points, offs, b, f, cap, voxels, coors4, num_points, fvo_host, st);
    if ((rc = stage_in(h, points, n * f * 4, st))) return rc;
    rc = bvx_generate_batched(h, h->dev_points, offs, b, f, cap, h->dev_voxels, h->dev_coors, h->dev_num,
                              h->dev_offs, nullptr, 0, st);
    if (rc) return rc;
    CK(h, cudaMemcpyAsync(fvo_host, h->dev_offs, ((size_t)b + 1) * 4, cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    const size_t m = (size_t)fvo_host[b];
    void* dst[3] = {voxels, coors4, num_points};
    const void* src[3] = {h->dev_voxels, h->dev_coors, h->dev_num};
    const size_t bytes[3] = {m * T * f * 4, m * 16, m * 4};
    h->last_h2d = n * f * 4;
    h->last_d2h = bytes[0] + bytes[1] + bytes[2] + ((size_t)b + 1) * 4;
    return stage_out(h, dst, src, bytes, st);
}

extern "C" int bvx_generate_host(bvx_handle* h, const float* points, int32_t n, int32_t f, int32_t max_voxels,
                                 float* voxels, int32_t* coors, int32_t* num_points, int32_t* voxel_num_host,
                                 void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (!voxel_num_host || n < 0) return fail(h, BVX_ERR_BAD_ARGUMENT, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->device));
    const int cap = max_voxels > 0 ? max_voxels : h->cfg.max_voxels;
    const int T = h->cfg.max_num_points;
    int rc = ensure_host_buffers(h, (size_t)n * f * 4, (size_t)cap, T * f);
    if (rc) return rc;
    if ((rc = stage_in(h, points, (size_t)n * f * 4, st))) return rc;
    rc = bvx_generate(h, h->dev_points, n, f, cap, h->dev_voxels, h->dev_coors, h->dev_num, h->dev_offs, st);
    if (rc) return rc;
    CK(h, cudaMemcpyAsync(voxel_num_host, h->dev_offs, 4, cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    const size_t m = (size_t)*voxel_num_host;
    void* dst[3] = {voxels, coors, num_points};
    const void* src[3] = {h->dev_voxels, h->dev_coors, h->dev_num};
    const size_t bytes[3] = {m * T * f * 4, m * 12, m * 4};
    h->last_h2d = (size_t)n * f * 4;
    h->last_d2h = bytes[0] + bytes[1] + bytes[2] + 4;
    h->last_redo = 0;
    return stage_out(h, dst, src, bytes, st);
}

extern "C" int bvx_set_prefix_upload(bvx_handle* h, int32_t on)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    h->prefix_upload = on != 0;
    h->cut_ratio = 0.f;
    return BVX_OK;
}

extern "C" int bvx_set_host_assemble(bvx_handle* h, int32_t on)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    h->host_assemble = on != 0;
    return BVX_OK;
}

extern "C" int bvx_set_host_threads(bvx_handle* h, int32_t n)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (n < 0) return fail(h, BVX_ERR_BAD_ARGUMENT, "host threads must be >= 0");
    if (h->pool && n != h->host_threads) { delete h->pool; h->pool = nullptr; }
    h->host_threads = n;
    return BVX_OK;
}

extern "C" int bvx_last_transfer(const bvx_handle* h, uint64_t* h2d_bytes, uint64_t* d2h_bytes, int32_t* redone)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (h2d_bytes) *h2d_bytes = h->last_h2d;
    if (d2h_bytes) *d2h_bytes = h->last_d2h;
    if (redone) *redone = h->last_redo;
    return BVX_OK;
}

extern "C" int bvx_frame_stats(bvx_handle* h, int32_t b, int32_t* cells, int32_t* voxels, int32_t* cutoff,
                               int32_t* processed, void* stream)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    if (b < 1 || b > h->cfg.max_frames) return fail(h, BVX_ERR_BAD_ARGUMENT, "b must be in [1, max_frames]");
    CK(h, cudaSetDevice(h->device));
    std::vector<FrameState> st((size_t)b);
    CK(h, cudaMemcpyAsync(st.data(), h->d_state, sizeof(FrameState) * (size_t)b, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CK(h, cudaStreamSynchronize((cudaStream_t)stream));
    for (int i = 0; i < b; ++i) {
        if (cells) cells[i] = st[i].total;
        if (voxels) voxels[i] = st[i].m;
        if (cutoff) cutoff[i] = st[i].istar;
        if (processed) processed[i] = st[i].n_done;
    }
    return BVX_OK;
}

#ifdef BVX_TIMELINE
extern "C" int bvx_debug_timeline(unsigned long long* out, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_tl, sizeof(unsigned long long) * 32);
    if (reset) {
        unsigned long long init[32];
        for (int i = 0; i < 32; ++i) init[i] = (i & 1) ? 0ull : ~0ull;
        cudaMemcpyToSymbol(g_tl, init, sizeof init);
    }
    return 0;
}
#endif

// ------------------------------------------------------------------ profiling
extern "C" int bvx_profile_enable(bvx_handle* h, int32_t on)
{
    if (!h) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null handle");
    h->prof = on != 0;
    return BVX_OK;
}

extern "C" int bvx_profile_read(bvx_handle* h, double* ms, int64_t* launches, int32_t reset)
{
    if (!h || !ms || !launches) return fail(h, BVX_ERR_BAD_ARGUMENT, "null argument");
    for (auto& pe : h->prof_pending) {
        float t = 0.f;
        CK(h, cudaEventSynchronize(pe.second.second));
        CK(h, cudaEventElapsedTime(&t, pe.second.first, pe.second.second));
        h->prof_ms[pe.first] += t;
        h->prof_n[pe.first] += 1;
    }
    h->prof_pending.clear();
    h->prof_used = 0;
    for (int i = 0; i < BVX_PROFILE_SLOTS; ++i) {
        ms[i] = h->prof_ms[i];
        launches[i] = h->prof_n[i];
        if (reset) { h->prof_ms[i] = 0; h->prof_n[i] = 0; }
    }
    return BVX_OK;
}

// ------------------------------------------------------------------ VFE / scatter
extern "C" int bvx_mean_vfe(const float* voxels, const int32_t* num_points, int32_t m, int32_t t, int32_t f,
                            int32_t c_out, int32_t simple, float* out, void* stream)
{
    if (m < 0 || t < 1 || f < 1 || c_out < 1 || c_out > f)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad mean_vfe shape m=%d t=%d f=%d c_out=%d", m, t, f, c_out);
    if (simple && c_out < 4) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "SimpleVoxel needs c_out >= 4");
    if (m == 0) return BVX_OK;
    if (!voxels || !num_points || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    int vt = std::max(1, std::min(64, 8192 / (t * f)));
    const size_t smem = ((size_t)vt * t * f + (size_t)vt * c_out) * 4;
    if (smem > 200 * 1024) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "t*f too large for the VFE tile");
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_mean_vfe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    const int blocks = (m + vt - 1) / vt;
    k_mean_vfe<<<blocks, kBlock, smem, (cudaStream_t)stream>>>(voxels, num_points, m, t, f, c_out, simple, vt, out);
    g_launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "k_mean_vfe launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

static int scatter_bev(const float* feats, const int32_t* coors4, int32_t m, const int32_t* m_dev, int32_t c,
                       int32_t b, int32_t d, int32_t ny, int32_t nx, float* canvas, void* stream);

extern "C" int bvx_scatter_bev(const float* feats, const int32_t* coors4, int32_t m, int32_t c, int32_t b,
                               int32_t d, int32_t ny, int32_t nx, float* canvas, void* stream)
{
    return scatter_bev(feats, coors4, m, nullptr, c, b, d, ny, nx, canvas, stream);
}

extern "C" int bvx_scatter_bev_dev(const float* feats, const int32_t* coors4, const int32_t* m_dev, int32_t m_cap,
                                   int32_t c, int32_t b, int32_t d, int32_t ny, int32_t nx, float* canvas,
                                   void* stream)
{
    if (!m_dev) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null voxel count pointer");
    return scatter_bev(feats, coors4, m_cap, m_dev, c, b, d, ny, nx, canvas, stream);
}

// Voxel-index map of the BEV scatter, one per (device, stream) -- calls on one stream are ordered, calls on
// different streams must not share it --, grown on demand and kept: its contents never need clearing
// (see k_bev_canvas), so a call costs no fill.
struct ScatterMap { int dev; cudaStream_t st; int* p; size_t count; };
static std::mutex g_map_mu;
static std::vector<ScatterMap> g_maps;
static int* scatter_map(cudaStream_t st, size_t cells);

static int scatter_bev(const float* feats, const int32_t* coors4, int32_t m, const int32_t* m_dev, int32_t c,
                       int32_t b, int32_t d, int32_t ny, int32_t nx, float* canvas, void* stream)
{
    if (m < 0 || c < 1 || b < 1 || d < 1 || ny < 1 || nx < 1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad scatter shape");
    if (!canvas || (m > 0 && (!feats || !coors4))) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if (((uintptr_t)canvas & 15) || ((uintptr_t)coors4 & 15))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "canvas/coors must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)b * d * ny * nx;
    if (m == 0) {
        // nothing to place: the canvas is all zeros
        const size_t total = cells * c, n4 = total >> 2;
        const int fb = (int)std::min<size_t>((n4 + kBlock - 1) / kBlock + 1, 148 * 16);
        k_fill_zero<<<fb, kBlock, 0, st>>>(reinterpret_cast<float4*>(canvas), n4, canvas + (n4 << 2), (int)(total & 3));
        g_launches += 1;
    } else {
        int* map = scatter_map(st, cells);
        if (!map) return fail(nullptr, BVX_ERR_CUDA, "scatter: cannot allocate the map");
        k_bev_map<<<(m + kBlock - 1) / kBlock, kBlock, 0, st>>>(reinterpret_cast<const int4*>(coors4), m, m_dev, b, d, ny, nx, map);
        // channel slices: enough blocks to fill the GPU, at most 8 channels per thread
        const bool vec = (nx & 3) == 0;
        const long long groups = (long long)(cells / (vec ? 4 : 1));
        const int gb = (int)((groups + kBlock - 1) / kBlock);
        int ch_per = c;
        while (ch_per > 8 || (ch_per > 1 && (long long)gb * ((c + ch_per - 1) / ch_per) < 148 * 8)) ch_per = (ch_per + 1) / 2;
        const dim3 grid(gb, (c + ch_per - 1) / ch_per);
        if (vec)
            k_bev_canvas<4><<<grid, kBlock, 0, st>>>(feats, reinterpret_cast<const int4*>(coors4), m, m_dev, c, b, d, ny, nx,
                                                     map, canvas, ch_per);
        else
            k_bev_canvas<1><<<grid, kBlock, 0, st>>>(feats, reinterpret_cast<const int4*>(coors4), m, m_dev, c, b, d, ny, nx,
                                                     map, canvas, ch_per);
        g_launches += 2;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "scatter launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

static int* scatter_map(cudaStream_t st, size_t cells)
{
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_map_mu);
    ScatterMap* sm = nullptr;
    for (ScatterMap& e : g_maps)
        if (e.dev == dev && e.st == st) sm = &e;
    if (!sm) {
        g_maps.push_back(ScatterMap{dev, st, nullptr, 0});
        sm = &g_maps.back();
    }
    if (sm->count < cells + 4) {
        if (sm->p) { cudaStreamSynchronize(st); cudaFree(sm->p); sm->p = nullptr; sm->count = 0; }
        if (cudaMalloc((void**)&sm->p, (cells + 4) * sizeof(int)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        sm->count = cells + 4;
    }
    return sm->p;
}

extern "C" int bvx_scatter_index_bev(const float* src, int32_t b, int32_t c, int32_t s0, int32_t s1,
                                     const int32_t* src_idx, const int32_t* dst_xy, int32_t np, int32_t ny,
                                     int32_t nx, float* out, void* stream)
{
    if (b < 1 || c < 1 || s0 < 1 || s1 < 1 || np < 0 || ny < 1 || nx < 1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad index-scatter shape");
    if (!out || (np > 0 && (!src || !dst_xy))) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if (((uintptr_t)out & 15) || ((uintptr_t)dst_xy & 7) || ((uintptr_t)src_idx & 7))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "out must be 16-byte aligned, the index arrays 8-byte aligned");
    if (!src_idx && (long long)np > (long long)s0 * s1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "implicit source mesh: np must be <= s0 * s1");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)b * ny * nx;
    int* map = scatter_map(st, cells);
    if (!map) return fail(nullptr, BVX_ERR_CUDA, "index scatter: cannot allocate the map");
    k_fill_i32<<<(int)std::min<size_t>((cells + kBlock - 1) / kBlock, 148 * 8), kBlock, 0, st>>>(map, cells, -1);
    g_launches += 1;
    if (np > 0) {
        const long long work = (long long)b * np;
        k_idx_map<<<(unsigned)((work + kBlock - 1) / kBlock), kBlock, 0, st>>>(reinterpret_cast<const int2*>(dst_xy), b, np, ny, nx, map);
        g_launches += 1;
    }
    const bool vec = (nx & 3) == 0;
    const long long groups = (long long)(cells / (vec ? 4 : 1));
    const int gb = (int)((groups + kBlock - 1) / kBlock);
    int ch_per = c;
    while (ch_per > 8 || (ch_per > 1 && (long long)gb * ((c + ch_per - 1) / ch_per) < 148 * 8)) ch_per = (ch_per + 1) / 2;
    const dim3 grid(gb, (c + ch_per - 1) / ch_per);
    if (vec) k_idx_canvas<4><<<grid, kBlock, 0, st>>>(src, reinterpret_cast<const int2*>(src_idx), c, b, s0, s1, ny, nx, map, out, ch_per);
    else k_idx_canvas<1><<<grid, kBlock, 0, st>>>(src, reinterpret_cast<const int2*>(src_idx), c, b, s0, s1, ny, nx, map, out, ch_per);
    g_launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "index scatter launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

extern "C" int bvx_vfe_augment(const float* voxels, const int32_t* num_points, int32_t m, int32_t t, int32_t f,
                               int32_t with_distance, float* out, float* mask, void* stream)
{
    if (m < 0 || t < 1 || f < 3) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad vfe_augment shape m=%d t=%d f=%d", m, t, f);
    if (m == 0) return BVX_OK;
    if (!voxels || !num_points || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    int vt = std::max(1, std::min(64, 8192 / (t * f)));
    const size_t smem = ((size_t)vt * t * f + (size_t)vt * 3) * 4;
    if (smem > 200 * 1024) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "t*f too large for the VFE tile");
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_vfe_augment, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    k_vfe_augment<<<(m + vt - 1) / vt, kBlock, smem, (cudaStream_t)stream>>>(voxels, num_points, m, t, f, with_distance ? 1 : 0,
                                                                             vt, out, mask);
    g_launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "k_vfe_augment launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

extern "C" int bvx_masked_max(const float* x, const int32_t* num_points, int32_t m, int32_t t, int32_t c, float* out,
                              int32_t* argmax, void* stream)
{
    if (m < 0 || t < 1 || c < 1) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad masked_max shape");
    if (m == 0) return BVX_OK;
    if (!x || !num_points || !out) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    const long long work = (long long)m * c;
    k_masked_max<<<(unsigned)((work + kBlock - 1) / kBlock), kBlock, 0, (cudaStream_t)stream>>>(x, num_points, m, t, c, out, argmax);
    g_launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "k_masked_max launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

// ---------------------------------------------------------------- BEV scatter without any allocation
// bvx_scatter_bev keeps its voxel-index map in a cudaMalloc'ed buffer (scatter_map above).  Callers that
// cannot allocate (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) get the
// same two kernels with the map in a file-scope device array: BVX_STATIC_SCATTER_CELLS ints = 4 MB that every
// GPU loading the library holds until exit.  In host code the array's NAME is a host-side stand-in, not a
// device address: the address is looked up once per device, with that device current, and passed to the kernels.
static __device__ __align__(256) int g_static_scatter_map[BVX_STATIC_SCATTER_CELLS];
static std::mutex g_static_map_mu;
static int* g_static_map_ptr[64] = {};

static int static_scatter_map(int** out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_static_map_mu);
    if (!g_static_map_ptr[dev]) {
        int* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_static_scatter_map);     // set-up: once per device, that device current
        if (e != cudaSuccess || !ptr)
            return fail(nullptr, BVX_ERR_CUDA, "cudaGetSymbolAddress(static scatter map): %s", cudaGetErrorString(e));
        g_static_map_ptr[dev] = ptr;
    }
    *out = g_static_map_ptr[dev];
    return BVX_OK;
}

extern "C" int bvx_scatter_bev_static(const float* feats, const int32_t* coors4, int32_t m, int32_t c, int32_t b,
                                      int32_t d, int32_t ny, int32_t nx, float* canvas, void* stream)
{
    if (m < 0 || c < 1 || b < 1 || d < 1 || ny < 1 || nx < 1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad scatter shape");
    if (!canvas || (m > 0 && (!feats || !coors4))) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if (((uintptr_t)canvas & 15) || ((uintptr_t)coors4 & 15))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "canvas/coors must be 16-byte aligned");
    const unsigned long long cells = (unsigned long long)b * d * ny * nx;
    if (cells > (unsigned long long)BVX_STATIC_SCATTER_CELLS)
        return fail(nullptr, BVX_ERR_WORKSPACE,
                    "bvx_scatter_bev_static: b*d*ny*nx = %llu cells exceed the static map of %d cells "
                    "(BVX_STATIC_SCATTER_CELLS); use bvx_scatter_bev", cells, (int)BVX_STATIC_SCATTER_CELLS);
    cudaStream_t st = (cudaStream_t)stream;
    if (m == 0) {
        // nothing to place: the canvas is all zeros
        const size_t total = (size_t)cells * c, n4 = total >> 2;
        const int fb = (int)std::min<size_t>((n4 + kBlock - 1) / kBlock + 1, 148 * 16);
        k_fill_zero<<<fb, kBlock, 0, st>>>(reinterpret_cast<float4*>(canvas), n4, canvas + (n4 << 2), (int)(total & 3));
        g_launches += 1;
    } else {
        int* map = nullptr;
        const int rc = static_scatter_map(&map);
        if (rc) return rc;
        // The map keeps the last call's contents and is NOT cleared: k_bev_canvas accepts an entry v only if
        // v < m and coors4[v] is that very cell, so stale entries are harmless (nothing must start at zero).
        k_bev_map<<<(m + kBlock - 1) / kBlock, kBlock, 0, st>>>(reinterpret_cast<const int4*>(coors4), m, nullptr, b, d, ny, nx, map);
        const bool vec = (nx & 3) == 0;
        const long long groups = (long long)(cells / (vec ? 4 : 1));
        const int gb = (int)((groups + kBlock - 1) / kBlock);
        int ch_per = c;
        while (ch_per > 8 || (ch_per > 1 && (long long)gb * ((c + ch_per - 1) / ch_per) < 148 * 8)) ch_per = (ch_per + 1) / 2;
        const dim3 grid(gb, (c + ch_per - 1) / ch_per);
        if (vec)
            k_bev_canvas<4><<<grid, kBlock, 0, st>>>(feats, reinterpret_cast<const int4*>(coors4), m, nullptr, c, b, d, ny, nx,
                                                     map, canvas, ch_per);
        else
            k_bev_canvas<1><<<grid, kBlock, 0, st>>>(feats, reinterpret_cast<const int4*>(coors4), m, nullptr, c, b, d, ny, nx,
                                                     map, canvas, ch_per);
        g_launches += 2;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "static scatter launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

// ---------------------------------------------------------------- anchors area
struct AaLayout {
    AnchorsMap mp;
    size_t off_above, off_cnt, off_fill, off_ticket, off_off, off_fbase, off_binned, bytes;   // byte offsets into the workspace
    size_t smem;
};

static size_t aa_round16(size_t v) { return (v + 15) & ~(size_t)15; }

static bool aa_layout(int32_t batch, int32_t ny, int32_t nx, int32_t max_voxels_per_frame, int32_t m, AaLayout* L)
{
    if (batch < 1 || ny < 1 || nx < 1 || max_voxels_per_frame < 0 || m < 0 || nx >= (1 << kAaYShift)) return false;
    AnchorsMap& mp = L->mp;
    mp.batch = batch; mp.ny = ny; mp.nx = nx;
    mp.packed = (max_voxels_per_frame <= 65535 && (nx % 2) == 0) ? 1 : 0;
    mp.wn = mp.packed ? nx / 2 : nx;
    // band height: as many rows as fit 48 KB of shared memory (4 blocks per SM), at most 32
    int h = kAaMaxBandRows, hs = 5;
    while (h > 1 && (size_t)h * mp.wn * 4 > 48 * 1024) { h >>= 1; --hs; }
    L->smem = (size_t)h * mp.wn * 4;
    if (L->smem > 200 * 1024) return false;            // one row must fit in shared memory
    mp.h = h; mp.hshift = hs; mp.nb = (ny + h - 1) / h;
    if ((long long)batch * mp.nb > (1 << 28)) return false;
    const size_t keys = (size_t)batch * mp.nb;
    size_t o = aa_round16((size_t)batch * ny * mp.wn * 4);
    L->off_above = o;  o += aa_round16(keys * mp.wn * 4);
    L->off_cnt = o;    o += aa_round16(keys * 4);      // cnt, fill, ticket: zeroed together every call
    L->off_fill = o;   o += aa_round16(keys * 4);
    L->off_ticket = o; o += 16;
    L->off_off = o;    o += aa_round16(keys * 4);
    L->off_fbase = o;  o += aa_round16(((size_t)batch + 1) * 4);
    L->off_binned = o; o += aa_round16((size_t)m * 4);
    L->bytes = o;
    return true;
}

extern "C" size_t bvx_anchors_area_workspace_bytes(int32_t batch, int32_t ny, int32_t nx,
                                                   int32_t max_voxels_per_frame, int32_t m)
{
    AaLayout L;
    if (!aa_layout(batch, ny, nx, max_voxels_per_frame, m, &L)) return 0;
    return L.bytes;
}

extern "C" int bvx_anchors_area(const int32_t* coors, int32_t m, int32_t coor_cols, int32_t batch,
                                const float voxel_size_xy[2], const float range_xy0[2],
                                const int32_t grid_xy[2], int32_t max_voxels_per_frame,
                                const float* anchors_bv, int32_t n_anchors, void* workspace,
                                size_t workspace_bytes, float* area, void* stream)
{
    if (!voxel_size_xy || !range_xy0 || !grid_xy) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    AaLayout L;
    if (n_anchors < 0 || (coor_cols != 3 && coor_cols != 4) ||
        !aa_layout(batch, grid_xy[1], grid_xy[0], max_voxels_per_frame, m, &L))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad or unsupported anchors-area shape");
    if (coor_cols == 3 && batch != 1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "[M,3] coordinates carry no batch index: batch must be 1");
    if (!(voxel_size_xy[0] > 0.f) || !(voxel_size_xy[1] > 0.f))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "voxel size must be positive");
    if (n_anchors == 0) return BVX_OK;
    if (!workspace || !area || !anchors_bv || (m > 0 && !coors))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if (workspace_bytes < L.bytes)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "workspace too small: need %zu bytes", L.bytes);
    if (((uintptr_t)workspace & 15) || ((uintptr_t)anchors_bv & 15) || (coor_cols == 4 && ((uintptr_t)coors & 15)))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "workspace/anchors/coors must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    AnchorsMap& mp = L.mp;
    char* ws = static_cast<char*>(workspace);
    mp.words = reinterpret_cast<unsigned int*>(ws);
    mp.above = reinterpret_cast<unsigned int*>(ws + L.off_above);
    mp.cnt = reinterpret_cast<int*>(ws + L.off_cnt);
    mp.fill = reinterpret_cast<int*>(ws + L.off_fill);
    mp.ticket = reinterpret_cast<int*>(ws + L.off_ticket);
    mp.off = reinterpret_cast<int*>(ws + L.off_off);
    mp.frame_base = reinterpret_cast<int*>(ws + L.off_fbase);
    mp.binned = reinterpret_cast<unsigned int*>(ws + L.off_binned);
    // counting sort of the voxels by (frame, band)
    const size_t n4 = (L.off_off - L.off_cnt) >> 4;
    const int fb = (int)std::min<size_t>((n4 + kBlock - 1) / kBlock, 148 * 8);
    k_fill_zero<<<std::max(fb, 1), kBlock, 0, st>>>(reinterpret_cast<float4*>(ws + L.off_cnt), n4, nullptr, 0);
    g_launches += 1;
    if (m > 0) {
        k_aa_count<<<(m + 1023) / 1024, 256, 0, st>>>(coors, m, coor_cols, mp);
        g_launches += 1;
    }
    k_aa_offsets<<<batch, 256, 0, st>>>(mp);
    g_launches += 1;
    if (m > 0) {
        k_aa_bin<<<(m + 1023) / 1024, 256, 0, st>>>(coors, m, coor_cols, mp);
        g_launches += 1;
    }
    static std::atomic<int> smem_set{0};
    if (L.smem > 48 * 1024 && !smem_set.load()) {
        cudaFuncSetAttribute(k_aa_band<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(k_aa_band<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        smem_set.store(1);
    }
    if ((mp.wn & 3) == 0) k_aa_band<4><<<dim3(mp.nb, batch), 256, L.smem, st>>>(mp);
    else k_aa_band<1><<<dim3(mp.nb, batch), 256, L.smem, st>>>(mp);
    k_aa_above<<<dim3((mp.wn + 255) / 256, batch), 256, 0, st>>>(mp);
    AnchorsQuery q;
    q.stride[0] = voxel_size_xy[0]; q.stride[1] = voxel_size_xy[1];
    q.offset[0] = range_xy0[0]; q.offset[1] = range_xy0[1];
    q.gx = mp.nx; q.gy = mp.ny;
    k_aa_query<<<dim3((n_anchors + 255) / 256, batch), 256, 0, st>>>(mp, anchors_bv, n_anchors, q, area);
    g_launches += 3;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "anchors-area launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

extern "C" int bvx_sparse_sum_for_anchors_mask(const int32_t* coors, int32_t m, int32_t coor_cols, int32_t ny,
                                               int32_t nx, float* ret, void* stream)
{
    if (m < 0 || ny < 1 || nx < 1 || (coor_cols != 3 && coor_cols != 4))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad sparse-sum shape");
    if (!ret || (m > 0 && !coors)) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if ((uintptr_t)ret & 15) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "ret must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t total = (size_t)ny * nx, n4 = total >> 2;
    const int fb = (int)std::min<size_t>((n4 + kBlock - 1) / kBlock + 1, 148 * 16);
    k_fill_zero<<<fb, kBlock, 0, st>>>(reinterpret_cast<float4*>(ret), n4, ret + (n4 << 2), (int)(total & 3));
    g_launches += 1;
    if (m > 0) {
        k_aa_hist_f32<<<(m + 255) / 256, 256, 0, st>>>(coors, m, coor_cols, coor_cols - 2, ny, nx, ret);
        g_launches += 1;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "sparse-sum launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}

extern "C" int bvx_fused_get_anchors_area(const float* dense_map, int32_t ny, int32_t nx, const float* anchors_bv,
                                          int32_t n_anchors, const float stride_xy[2], const float offset_xy[2],
                                          const int32_t grid_xy[2], float* ret, void* stream)
{
    if (ny < 1 || nx < 1 || n_anchors < 0 || !stride_xy || !offset_xy || !grid_xy || grid_xy[0] < 1 || grid_xy[1] < 1)
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "bad anchors-area shape");
    if (!(stride_xy[0] > 0.f) || !(stride_xy[1] > 0.f))
        return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "stride must be positive");
    if (n_anchors == 0) return BVX_OK;
    if (!dense_map || !anchors_bv || !ret) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "null argument");
    if ((uintptr_t)anchors_bv & 15) return fail(nullptr, BVX_ERR_BAD_ARGUMENT, "anchors must be 16-byte aligned");
    AnchorsQuery q;
    q.stride[0] = stride_xy[0]; q.stride[1] = stride_xy[1];
    q.offset[0] = offset_xy[0]; q.offset[1] = offset_xy[1];
    q.gx = grid_xy[0]; q.gy = grid_xy[1];
    k_aa_query_f32<<<(n_anchors + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dense_map, ny, nx, anchors_bv,
                                                                                n_anchors, q, ret);
    g_launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, BVX_ERR_CUDA, "anchors-area launch: %s", cudaGetErrorString(e));
    return BVX_OK;
}
