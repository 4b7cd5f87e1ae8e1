/*
 * bvx.h -- C ABI of the B200-native voxelization hot path.
 *
 * Drop-in boundary for jhkohpav/3dobject_detection_temporal (a SECOND fork).
 * The reference reaches this path through a Python object protocol, not a C
 * ABI; each entry point below names the reference interface it replaces
 * (paths relative to the reference root).  The Python mirror of that protocol
 * (VoxelGenerator / VoxelFeatureExtractorV3 / PointPillarsScatter) lives in
 * 3dobject_detection_temporal_b200/ and binds these symbols with ctypes; the
 * stub a reference maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / pybind types.
 *   - every `*_dev` / unqualified data pointer is a DEVICE pointer owned by the
 *     caller (torch allocates all inputs and outputs); the library owns only
 *     the per-handle workspace.  `*_host` entry points take HOST pointers.
 *   - all work is enqueued on the given cudaStream_t (passed as void*); the
 *     device entry points never synchronise.
 *   - return value: 0 = BVX_OK, otherwise a bvx_status; the message is
 *     available from bvx_last_error().  Nothing throws across the ABI.
 *   - one handle per (device, stream); not thread-safe.
 *   - there is NO CPU fallback: every entry point needs a CUDA device.
 */
#ifndef BVX_H_
#define BVX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BVX_ABI_VERSION 5

typedef enum bvx_status {
    BVX_OK = 0,
    BVX_ERR_BAD_ARGUMENT = 1,   /* Python side raises ValueError   */
    BVX_ERR_CUDA = 2,           /* Python side raises RuntimeError */
    BVX_ERR_WORKSPACE = 3,      /* capacity given at create time exceeded */
    BVX_ERR_NO_DEVICE = 4
} bvx_status;

/* What happens when voxel number `max_voxels` would be created:
 * BREAK    = stop scanning, every later point is dropped (spconv v1.0 /
 *            second.pytorch v1.5, the era this fork derives from) -- default.
 * CONTINUE = skip only points that would open a new voxel (later upstream). */
#define BVX_OVERFLOW_BREAK 0
#define BVX_OVERFLOW_CONTINUE 1

/* Cell -> voxel lookup structure.  HASHED: open-addressing table sized by the
 * number of points (L2 resident).  DENSE: one entry per grid cell, the layout
 * of the reference's cached `coor_to_voxelidx` grid. */
#define BVX_GRID_HASHED 0
#define BVX_GRID_DENSE 1

typedef struct bvx_config {
    float voxel_size[3];         /* x, y, z                                   */
    float point_cloud_range[6];  /* x0 y0 z0 x1 y1 z1                         */
    int32_t max_num_points;      /* T: points kept per voxel                  */
    int32_t max_voxels;          /* capacity AND default per-frame cap        */
    int32_t overflow_mode;       /* BVX_OVERFLOW_*                            */
    int32_t grid_kind;           /* BVX_GRID_*                                */
    int32_t max_points;          /* capacity: points per call (all frames)    */
    int32_t max_frames;          /* capacity: frames per call                 */
    int32_t max_point_features;  /* capacity: output channels F_out per point */
    int32_t device;              /* CUDA device ordinal                       */
} bvx_config;

typedef struct bvx_handle bvx_handle;

/* One sweep of a temporal frame (new functionality named by the
 * north star -- the reference voxelizes current and previous frame separately,
 * second/pytorch/train_2st_spatio.py:182-195).  Row i of the virtual merged
 * cloud is (R @ xyz + t, feat[3:f_in], dt) with
 *   x' = ((r00*x + r01*y) + r02*z) + tx   in fp32, no FMA contraction.
 * has_transform == 0 copies xyz verbatim (key frame). */
typedef struct bvx_sweep {
    const float* points;   /* device, [n, f_in] row-major */
    int32_t n;
    int32_t has_transform;
    float R[9];            /* row-major 3x3, sweep -> key frame */
    float t[3];
    float dt;              /* seconds, appended as channel f_in */
    const int32_t* perm;   /* optional, device, [n]: row j of the sweep is points[perm[j]] -- the reference's
                            * shuffle_points (np.random.shuffle before voxelization; input_reader.proto:10,
                            * second/configs/.../server_lidaronly.fhd.onestage.config:109) as a gather on the
                            * point load instead of a shuffled copy.  NULL = rows in memory order. */
} bvx_sweep;

/* Optional convex crop of a frame's points, tested on the raw coordinates before
 * the sweep transform: the reference's remove_outside_points
 * (second/core/box_np_ops.py:645-657) -> points_in_convex_polygon_3d_jit
 * (second/core/geometry.py:96-141).  planes[k] = (a, b, c, d) is one surface of
 * the polygon as surface_equ_3d returns it (normal_vec, d), float64; a point
 * with a*x + b*y + c*z + d >= 0 for any surface is dropped (evaluated in float64,
 * left to right, like the reference's loop).  Dropping a point and compacting
 * the array, as the reference does, gives the same voxels. */
typedef struct bvx_crop {
    int32_t n_planes;      /* 0 .. 8; 0 = no crop */
    int32_t reserved;
    double planes[8][4];
} bvx_crop;

/* Replaces spconv.utils.VoxelGenerator.__init__ as constructed at
 * second/builder/voxel_builder.py:23-27.  Allocates the workspace. */
int bvx_create(const bvx_config* cfg, bvx_handle** out);
int bvx_destroy(bvx_handle* h);

/* VoxelGenerator.grid_size: round((range[3:]-range[:3])/voxel_size) in fp32,
 * half-to-even, x y z (same formula: second/kittiviewer/viewer.py:468-469;
 * consumers: second/pytorch/builder/second_builder.py:17-18). */
int bvx_grid_size(const float voxel_size[3], const float point_cloud_range[6],
                  int64_t out_xyz[3]);

/* Bytes of device workspace the handle owns. */
int bvx_workspace_bytes(const bvx_handle* h, size_t* out);

/* Replaces VoxelGenerator.generate(points, max_voxels) ->
 * (voxels, coordinates, num_points_per_voxel)   [called inside the reference's
 * prep_pointcloud; wired at second/builder/dataset_builder.py:51-58 and
 * second/core/inference.py:43-48; twin call sites
 * second/kittiviewer/viewer.py:385-392, 451-457].
 *   points     [n, f]              f32
 *   voxels     [max_voxels, T, f]  f32, rows [0, M) written, zero padded
 *   coors      [max_voxels, 3]     i32 (z, y, x)
 *   num_points [max_voxels]        i32
 *   voxel_num  [1]                 i32, M written on device
 * max_voxels <= 0 selects the handle default. */
int bvx_generate(bvx_handle* h, const float* points, int32_t n, int32_t f,
                 int32_t max_voxels, float* voxels, int32_t* coors,
                 int32_t* num_points, int32_t* voxel_num, void* stream);

/* Temporal input: K sweeps merged in-kernel with the dt channel appended; the
 * merged cloud is never written to memory.  Outputs as bvx_generate with
 * f_out = f_in + 1. */
int bvx_generate_sweeps(bvx_handle* h, const bvx_sweep* sweeps_host, int32_t k,
                        int32_t f_in, int32_t max_voxels, float* voxels,
                        int32_t* coors, int32_t* num_points, int32_t* voxel_num,
                        void* stream);

/* B frames per call; replaces B calls to generate() plus the voxel part of the
 * reference's collate merge_second_batch (imported second/pytorch/train.py:17):
 * outputs are concatenated in frame order and coors carry the batch index,
 * [sum M, 4] i32 (b, z, y, x) (second/pytorch/models/voxelnet.py:285).
 *   points               [sum n_b, f], frame b = rows [off[b], off[b+1])
 *   frame_offsets_host   [b + 1] HOST int32
 *   voxels               [b * max_voxels, T, f]   (capacity; [0, sum M) written)
 *   coors4               [b * max_voxels, 4]
 *   num_points           [b * max_voxels]
 *   frame_voxel_offsets  [b + 1] DEVICE int32, exclusive prefix of M_b
 *   mean_features        optional [b * max_voxels, mean_channels] f32: the mean
 *                        VFE fused into the gather (NULL to skip).  With it,
 *                        voxels may be NULL: the padded rows are not written. */
int bvx_generate_batched(bvx_handle* h, const float* points,
                         const int32_t* frame_offsets_host, int32_t b, int32_t f,
                         int32_t max_voxels, float* voxels, int32_t* coors4,
                         int32_t* num_points, int32_t* frame_voxel_offsets,
                         float* mean_features, int32_t mean_channels,
                         void* stream);

/* B frames of k_b sweeps each: the temporal input of the north star, batched.
 * Frame i is sweeps [sweep_offsets_host[i], sweep_offsets_host[i+1]) of
 * sweeps_host, key frame first; rows of one frame are the concatenation of its
 * sweeps, each transformed in registers (bvx_sweep above), never written to
 * memory.  append_dt != 0 appends the sweep's dt as channel f_in (f_out =
 * f_in + 1); append_dt == 0 keeps f_out = f_in -- with one transformed sweep
 * per frame this is the global rotation / scaling / flip of the reference's
 * augmentation (second/core/preprocess.py:756-804) fused into the point load.
 * crops_host: NULL, or one bvx_crop per frame (the camera-frustum crop of
 * remove_outside_points rides the same load).
 * Outputs as bvx_generate_batched with f = f_out. */
int bvx_generate_batched_sweeps(bvx_handle* h, const bvx_sweep* sweeps_host,
                                const int32_t* sweep_offsets_host, int32_t b,
                                int32_t f_in, int32_t append_dt,
                                const bvx_crop* crops_host,
                                int32_t max_voxels, float* voxels,
                                int32_t* coors4, int32_t* num_points,
                                int32_t* frame_voxel_offsets,
                                float* mean_features, int32_t mean_channels,
                                void* stream);

/* Host-buffer forms (numpy in, numpy out -- the exact shape of the reference's
 * CPU call).  Copies H2D, runs the device path, copies the M result rows D2H,
 * synchronises the stream and returns M in *voxel_num_host.  Pinned buffers
 * (bvx_host_alloc) are used in place; pageable ones go through a pinned
 * staging buffer owned by the handle. */
int bvx_generate_host(bvx_handle* h, const float* points_host, int32_t n,
                      int32_t f, int32_t max_voxels, float* voxels_host,
                      int32_t* coors_host, int32_t* num_points_host,
                      int32_t* voxel_num_host, void* stream);
int bvx_generate_batched_host(bvx_handle* h, const float* points_host,
                              const int32_t* frame_offsets_host, int32_t b,
                              int32_t f, int32_t max_voxels, float* voxels_host,
                              int32_t* coors4_host, int32_t* num_points_host,
                              int32_t* frame_voxel_offsets_host, void* stream);
int bvx_host_alloc(void** out, size_t bytes);  /* pinned host memory */
int bvx_host_free(void* p);

/* Pinned batched host form, BVX_OVERFLOW_BREAK: nothing after a frame's cut-off
 * point (the point that would open voxel number max_voxels) can reach the
 * output -- the reference's loop stops reading there -- so only a prefix of
 * each frame is copied to the device: 1.25 x the largest cut-off fraction seen
 * in the previous call on this handle, plus two chunks (the first call copies
 * everything).  Each frame's own result tells whether its prefix contained the
 * cut-off; if not, the rest of the frame is copied and its sub-batch is run
 * again, so results never depend on the guess.  On by default.
 * bvx_last_transfer: bytes copied each way by the last *_host call on the
 * handle and how many sub-batches had to be run again. */
int bvx_set_prefix_upload(bvx_handle* h, int32_t on);
/* Pinned batched host form: the padded rows voxels[M, T, f] are 4-5x their
 * information (cfg4: 1.2 points per voxel in T = 5 slots) and the PCIe link is
 * what bounds the call.  With host assembly on (default) the padding does not
 * cross the link: for T <= 8 and 4 features the kept rows come home compact
 * (16 bytes per kept point, in voxel order, plus one int per 256 voxels), else
 * the frame-local point index of every kept (voxel, slot) does (4*T bytes per
 * voxel) and the rows are read from the caller's own points.  A pool of host
 * threads owned by the handle (one per host core, at most 32) writes the padded
 * rows with streaming stores, sub-batch k-1 while the copies of sub-batch k are
 * in flight.  Every row of the reference's voxels is a verbatim copy of an
 * input row, so the result is bit-identical; voxel numbering, slots, counts and
 * coordinates are still computed on the device.  Off: the padded rows are
 * gathered on the device and copied home. */
int bvx_set_host_assemble(bvx_handle* h, int32_t on);
/* Host threads of the assembly: 0 (default) = one per core the process may run
 * on (sched_getaffinity), at most 32.  With several ranks per host give each
 * its share (cores / ranks on the host). */
int bvx_set_host_threads(bvx_handle* h, int32_t n);
/* The host half of the assembly on its own: host memory only, no device, no
 * handle (what the pipelined host form runs per sub-batch; also how it is
 * tested without a GPU).
 * rows: voxels[v, s, :] = points[frame_offsets[frame(v)] + idx[v, s], :] for
 *   s < num_points[v], zero for the other slots; frame(v) from
 *   frame_voxel_offsets [b + 1]; idx [M, t] frame-local point indices.
 * compact: rows = the kept rows (4 floats each) tile after tile; tile j of
 *   frame i (256 voxels, ceil(max_voxels / 256) tiles per frame) starts at
 *   row tile_base[i * tiles + j]; inside a tile voxel after voxel, slot after
 *   slot.  rows and voxels 16-byte aligned. */
int bvx_host_assemble_rows(const float* points, const int32_t* frame_offsets,
                           const int32_t* frame_voxel_offsets, int32_t b, int32_t t,
                           int32_t f, const int32_t* num_points, const int32_t* idx,
                           float* voxels, int32_t threads);
int bvx_host_assemble_compact(const float* rows, const int32_t* tile_base,
                              const int32_t* frame_voxel_offsets, int32_t b,
                              int32_t max_voxels, int32_t t, const int32_t* num_points,
                              float* voxels, int32_t threads);
int bvx_last_transfer(const bvx_handle* h, uint64_t* h2d_bytes,
                      uint64_t* d2h_bytes, int32_t* redone);

/* Per-frame facts of the last device- or host-form call on the handle (frames
 * 0 .. b-1 of that call; for the pipelined host form: of its last sub-batch),
 * copied to HOST arrays (each may be NULL); synchronises the stream.
 *   cells     distinct in-range cells among the points that were looked at
 *   voxels    M_b = min(cells, max_voxels)
 *   cutoff    BREAK mode, cells > max_voxels: index of the point that would have
 *             opened voxel number max_voxels (the reference's loop stops there);
 *             else = processed
 *   processed points of the frame that went through the insert kernel: in BREAK
 *             mode the kernel stops handing out chunks of a frame once more than
 *             max_voxels cells are claimed, so this is a prefix of the frame.
 *             bench.py uses it for the "bytes actually read" roofline figure. */
int bvx_frame_stats(bvx_handle* h, int32_t b, int32_t* cells, int32_t* voxels,
                    int32_t* cutoff, int32_t* processed, void* stream);

/* Replaces VoxelFeatureExtractorV3.forward(features, num_voxels)
 * (second/pytorch/models/voxel_encoder.py:201-206): out[v, c] =
 * sum_t voxels[v, t, c] / num_points[v] for c < c_out.  The reference
 * hard-codes c_out = 4; c_out = f keeps the dt channel of temporal input.
 * simple_voxel != 0 emits SimpleVoxel's [ |xy|_2, z, refl ] instead
 * (voxel_encoder.py:221-228; needs c_out >= 4, output has 3 channels). */
int bvx_mean_vfe(const float* voxels, const int32_t* num_points, int32_t m,
                 int32_t t, int32_t f, int32_t c_out, int32_t simple_voxel,
                 float* out, void* stream);

/* PointPillarsScatter / SparseConvTensor.dense() replacement behind the middle
 * extractor protocol forward(voxel_features, coors, batch_size) ->
 * [N, C*D, H, W] (second/pytorch/models/middle.py:513-526); math of
 * torchplus.scatter_nd (torchplus/ops/array_ops.py:7-20): zeros, then
 * canvas[b, c*D + z, y, x] = feats[i, c] for coors4[i] = (b, z, y, x).
 * The zero fill is part of the call. */
int bvx_scatter_bev(const float* feats, const int32_t* coors4, int32_t m,
                    int32_t c, int32_t b, int32_t d, int32_t ny, int32_t nx,
                    float* canvas, void* stream);
/* Same, with the number of voxels read from device memory (frame_voxel_offsets + B
 * of the bvx_generate_batched call that produced them; m_cap = the capacity the
 * arrays were sized for): points -> canvas without a host synchronisation.
 * Together with bvx_generate_batched(voxels = NULL, mean_features = feats) this is
 * the fused pillar path: the padded [M, T, f] rows are never written (SURVEY 8d:
 * algorithmic bytes 4*(N*f + B*C*ny*nx)). */
int bvx_scatter_bev_dev(const float* feats, const int32_t* coors4,
                        const int32_t* m_dev, int32_t m_cap, int32_t c, int32_t b,
                        int32_t d, int32_t ny, int32_t nx, float* canvas,
                        void* stream);

/* bvx_scatter_bev for callers that cannot allocate.  NO COUNTERPART IN THE
 * REFERENCE: added for code under CUDA
 * graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same arguments, same result as bvx_scatter_bev, but the voxel-index map the
 * two kernels use as scratch lives in a file-scope device array of
 * BVX_STATIC_SCATTER_CELLS int32 (4 MB, held by every GPU that loads the
 * library until exit) instead of a cudaMalloc'ed buffer: the call allocates and
 * frees nothing, on the device or through any library.
 *   limit   b * d * ny * nx <= BVX_STATIC_SCATTER_CELLS, else BVX_ERR_WORKSPACE
 *           (cfg3's 496 x 432 canvas: up to 4 frames per call).
 *   sharing every call on a device shares the one array; it keeps the last call's
 *           contents and is not cleared -- an entry only counts if coors4 of the
 *           voxel it names is that very cell, so stale entries are harmless.
 *           Calls that OVERLAP on the GPU (different streams, no ordering between
 *           them) race on the array: order them, or use bvx_scatter_bev, whose
 *           map is per stream. */
#define BVX_STATIC_SCATTER_CELLS (1 << 20)
int bvx_scatter_bev_static(const float* feats, const int32_t* coors4, int32_t m,
                           int32_t c, int32_t b, int32_t d, int32_t ny,
                           int32_t nx, float* canvas, void* stream);

/* The bandwidth part of the learned VFEs VoxelFeatureExtractor / V2
 * (second/pytorch/models/voxel_encoder.py:68-189; their linear + batch-norm
 * layers stay torch modules, see modules.py).
 * bvx_vfe_augment = what their forward does before the first linear layer
 * (voxel_encoder.py:100-112, 165-176): points_mean = sum_t xyz / num;
 *   out[v, t, :] = (voxels[v, t, :f], xyz - points_mean [, |xyz|_2])   [m, t, f + 3 (+ 1)]
 * for all t slots -- the reference does not mask its input, a padded slot gets
 * -points_mean --, and mask[v, t] = (t < num[v]) as float, the padding indicator
 * of get_paddings_indicator (voxel_encoder.py:12-33); mask may be NULL.
 * bvx_masked_max = their tail (voxel_encoder.py:124-129, 184-189): x *= mask,
 * voxelwise = max over t:  out[v, c] = max_t x[v, t, c] * (t < num[v]);
 * argmax (optional, [m, c] int32) = the winning t, for a backward pass. */
int bvx_vfe_augment(const float* voxels, const int32_t* num_points, int32_t m,
                    int32_t t, int32_t f, int32_t with_distance, float* out,
                    float* mask, void* stream);
int bvx_masked_max(const float* x, const int32_t* num_points, int32_t m,
                   int32_t t, int32_t c, float* out, int32_t* argmax,
                   void* stream);

/* Camera -> BEV projection scatter of the fusion RPN
 * (second/pytorch/models/rpn.py:957-962), for the whole batch in one call:
 *   out[b, :, dst_xy[b, j].y, dst_xy[b, j].x] = src[b, :, src_idx[j].0, src_idx[j].1]
 * for j = 0 .. np-1, zeros elsewhere; where several j fall into one cell the last
 * one wins (what the reference's sequential CPU assignment does).
 *   src      [b, c, s0, s1] f32      out [b, c, ny, nx] f32 (16-byte aligned)
 *   src_idx  [np, 2] int32 (first, second index into s0 x s1), shared by the batch;
 *            NULL = the implicit mesh j -> (j / s1, j % s1)
 *   dst_xy   [b, np, 2] int32 (x, y); out-of-range targets are skipped. */
int bvx_scatter_index_bev(const float* src, int32_t b, int32_t c, int32_t s0,
                          int32_t s1, const int32_t* src_idx,
                          const int32_t* dst_xy, int32_t np, int32_t ny,
                          int32_t nx, float* out, void* stream);

/* Anchors area from voxel coordinates, the computation
 * behind the reference's anchors_mask:
 *   dense = sparse_sum_for_anchors_mask(coors, (ny, nx))      second/core/box_np_ops.py:776-781
 *   dense = dense.cumsum(0).cumsum(1)                         second/kittiviewer/viewer.py:397-398
 *   area  = fused_get_anchors_area(dense, anchors_bv,         second/core/box_np_ops.py:784-810
 *                                  voxel_size, range, grid)
 * in one call, for every frame of a batch, without leaving the device.
 * coors: [m, coor_cols] int32, rows (z, y, x) (coor_cols 3, batch must be 1) or
 * (b, z, y, x) (coor_cols 4) -- exactly what bvx_generate / bvx_generate_batched
 * wrote.  grid_xy = grid_size x, y (the map is [grid_xy[1], grid_xy[0]]).
 * anchors_bv: [n_anchors, 4] float32 rows (x0, y0, x1, y1), shared by the frames.
 * area: [batch, n_anchors] float32 voxel counts; the caller thresholds it
 * (anchor_area_threshold, second/protos/input_reader.proto:23).
 * max_voxels_per_frame bounds the rows any one frame contributes (selects 16- or
 * 32-bit table cells; a frame with more rows than this is undefined behaviour).
 * workspace: device scratch of at least bvx_anchors_area_workspace_bytes(...,
 * m) bytes, 16-byte aligned, contents irrelevant (0 is returned for shapes the
 * library cannot serve: empty, or one map row larger than 200 KB).
 * Anchors wholly outside the range index out of bounds in the reference; here
 * their corners are clamped into the map. */
size_t bvx_anchors_area_workspace_bytes(int32_t batch, int32_t ny, int32_t nx,
                                        int32_t max_voxels_per_frame,
                                        int32_t m);
int bvx_anchors_area(const int32_t* coors, int32_t m, int32_t coor_cols,
                     int32_t batch, const float voxel_size_xy[2],
                     const float range_xy0[2], const int32_t grid_xy[2],
                     int32_t max_voxels_per_frame, const float* anchors_bv,
                     int32_t n_anchors, void* workspace, size_t workspace_bytes,
                     float* area, void* stream);

/* The same two reference functions under their own names, for callers that keep
 * the reference's decomposition (float32 maps, the caller does the two cumsums).
 * ret: [ny, nx] float32, zero filled by the call; the y, x columns of coors are
 * the last two (box_np_ops.py:779 reads columns 1, 2 of [M, 3]). */
int bvx_sparse_sum_for_anchors_mask(const int32_t* coors, int32_t m,
                                    int32_t coor_cols, int32_t ny, int32_t nx,
                                    float* ret, void* stream);
int bvx_fused_get_anchors_area(const float* dense_map, int32_t ny, int32_t nx,
                               const float* anchors_bv, int32_t n_anchors,
                               const float stride_xy[2],
                               const float offset_xy[2],
                               const int32_t grid_xy[2], float* ret,
                               void* stream);

/* Optional per-kernel device timing: CUDA events are recorded on the call's
 * stream around every kernel of the pipeline while enabled.  bvx_profile_read
 * synchronises those events and returns accumulated milliseconds and launch
 * counts per kernel slot: 0 k_insert, 1 k_mid (frame scan, voxel numbers,
 * slots of the non-first points), 4 k_overflow, 5 k_gather (slots 2, 3, 6, 7
 * unused). */
#define BVX_PROFILE_SLOTS 8
int bvx_profile_enable(bvx_handle* h, int32_t on);
int bvx_profile_read(bvx_handle* h, double ms[BVX_PROFILE_SLOTS],
                     int64_t launches[BVX_PROFILE_SLOTS], int32_t reset);

/* Message of the last failing call on this handle (or, with h == NULL, of the
 * last failing handle-less call on this thread). */
const char* bvx_last_error(const bvx_handle* h);

/* Kernel launches enqueued by this library since load (for gpu_launches). */
int64_t bvx_launch_count(void);
int bvx_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BVX_H_ */
