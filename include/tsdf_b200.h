/*
 * tsdf_b200.h — C ABI of the B200-native TSDF volumetric fusion engine.
 *
 * This is the drop-in boundary for semantic-slam's TSDF integration hot path.
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns an
 * `int` status (0 = TSDF_OK) and never throws, exits or resets the device.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   class TSDF          include/tsdf.hpp:22-98      (ctor src/tsdf.cu:62-96,
 *                        Integrate :135-168, dtor/readback :98-133,
 *                        SaveVoxelGrid2SurfacePointCloud :170-218)
 *   class TSDFfusion    include/TSDFfusion.hpp:25-65 (src/TSDFfusion.cpp:74-105 ->
 *                        src/TSDFfusion.py.in:21-53 -> third-party fusion.TSDFVolume)
 *   Object::Object      src/Object.cpp:23-49  (per-instance base pose + origin)
 *   Engine::Run         src/Engine.cpp:192-193 (per-instance masked depth)
 *
 * Conventions:
 *   depth  : fp32 metres, row-major H x W, 0 (or <0, or > max_depth) = invalid
 *   rgb    : 8-bit x 3 interleaved, row-major H x W x 3, channel order "as given";
 *            folded as c2*65536 + c1*256 + c0 (src/TSDFfusion.py.in:43 -> upstream)
 *   poses  : row-major 4x4 camera-to-world (Twc = KF->GetPoseInverse(),
 *            src/Object.cpp:23,162)
 *   K      : row-major 3x3, only [0],[2],[4],[5] are used (src/tsdf.cu:41-42)
 *
 * Threading: a handle is single-writer; different handles may be driven from
 * different host threads (each owns its non-default stream).
 * Pointers passed to *_async calls are borrowed until tsdf_sync() unless the
 * call documents a copy.  Host pointers may be pageable or pinned; device
 * pointers (cudaMalloc) are accepted wherever a frame is taken and skip the
 * upload.
 */
#ifndef TSDF_B200_H
#define TSDF_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TSDF_B200_VERSION 100 /* 0.1.0 */

/* ---- status codes ------------------------------------------------------ */
enum {
  TSDF_OK = 0,
  TSDF_E_INVALID = 1,    /* bad argument / config                           */
  TSDF_E_CUDA = 2,       /* CUDA runtime error (text in tsdf_last_error())  */
  TSDF_E_NOMEM = 3,      /* host or device allocation failed                */
  TSDF_E_CAPACITY = 4,   /* caller output buffer too small (count returned) */
  TSDF_E_IO = 5,         /* file open/read/write failed                     */
  TSDF_E_NCCL = 6,       /* NCCL missing or returned an error               */
  TSDF_E_SINGULAR = 7,   /* base2world not invertible (src/tsdf.cu:394-395) */
  TSDF_E_NODEVICE = 8    /* no CUDA device: there is NO CPU fallback        */
};

/* ---- semantics / layout ------------------------------------------------ */
/* TSDF_SEM_TSDFCU  : exactly src/tsdf.cu:15-60 — depth only, weight += 1,
 *                    skip diff <= -trunc, linear index z*Dy*Dx + y*Dx + x
 *                    (x fastest), extraction/points in the BASE frame.
 * TSDF_SEM_FUSIONPY: the TSDFfusion plugin surface (src/TSDFfusion.py.in:43):
 *                    colour + obs_weight, skip diff < -trunc, linear index
 *                    x*Dy*Dz + y*Dz + z (z fastest, numpy C order of
 *                    [Dx,Dy,Dz]); volume lives in the world frame.         */
enum { TSDF_SEM_TSDFCU = 0, TSDF_SEM_FUSIONPY = 1 };

/* Extraction predicate (K5):
 * TSDF_PRED_NONZERO : fabs(tsdf) != 0 && weight > weight_thresh  (src/tsdf.cu:179,198)
 * TSDF_PRED_ABS_LT  : fabs(tsdf) <  tsdf_thresh && weight > weight_thresh
 *                     (Zeng's original, src/obselete/tsdf.cpp:16,33)       */
enum { TSDF_PRED_NONZERO = 0, TSDF_PRED_ABS_LT = 1 };

/* Raw depth sample formats accepted by tsdf_integrate_raw.
 * The RANGE formats are lidar images whose pixels hold the distance along the ray; the depth is
 * d = range*scale / sqrt(x*x + y*y + 1), x = (col-cx)/fx, y = (row-cy)/fy
 * (src/DoN.cpp:86-100 and src/Object.cpp:611-619, mSensor == 1), computed on the device. */
enum { TSDF_DEPTH_F32 = 0, TSDF_DEPTH_U16 = 1, TSDF_DEPTH_RANGE_F32 = 2, TSDF_DEPTH_RANGE_U16 = 3 };

typedef struct tsdf_vol tsdf_vol;     /* one volume (or one z-slab of it) on one GPU */
typedef struct tsdf_comm tsdf_comm;   /* one NCCL rank                              */
typedef struct tsdf_batch tsdf_batch; /* N per-instance volumes, one launch          */
typedef struct tsdf_group tsdf_group; /* one volume as z-slabs over several GPUs, ONE process */

typedef struct tsdf_config {
  int32_t dim_x, dim_y, dim_z; /* GLOBAL grid dims (include/tsdf.hpp:65-67: 200^3)   */
  float voxel_size;            /* include/tsdf.hpp:63: 0.004                         */
  float trunc_margin;          /* include/tsdf.hpp:64: 5*voxel_size                  */
  float origin[3];             /* grid origin in the base frame (src/tsdf.cu:66-68)  */
  float base2world[16];        /* row-major; identity => volume in the world frame   */
  int32_t im_h, im_w;          /* frame shape (ctor h,w — src/tsdf.cu:62-63)         */
  float K[9];                  /* include/tsdf.hpp:96 (TUM3)                         */
  float max_depth;             /* src/tsdf.cu:46: 6.0; +inf disables                 */
  int32_t semantics;           /* TSDF_SEM_*                                         */
  int32_t with_color;          /* allocate + integrate the packed-RGB array          */
  int32_t device;              /* CUDA device ordinal                                */
  int32_t z_begin, z_end;      /* slab [z_begin,z_end) owned by this handle;
                                  0,0 => whole grid                                  */
  int32_t id;                  /* mnId: names tsdf<id>.ply / tsdf<id>.bin            */
  int32_t reserved[7];
} tsdf_config;

/* Scope: the last integrate CALL on the handle.  After tsdf_integrate_multi_async(k frames) every field
 * covers all k frames: voxels_visited = k * slab voxels, voxels_written = the sum, last_kernel_ms = from
 * the start of the call's first integrate launch to the end of its last one (per-launch times: use
 * tsdf_kernel_timing / tsdf_kernel_timing_read), and bytes_algorithmic counts every update as a pass of
 * its own (24|16 B) although two dense frames that share one pass move half of that per update.     */
typedef struct tsdf_stats {
  int64_t voxels_visited;  /* voxels evaluated by the last integrate on this handle  */
  int64_t voxels_written;  /* voxels whose weight/tsdf were updated (exact)          */
  int64_t frames;          /* frames integrated since create/reset                   */
  float last_kernel_ms;    /* CUDA-event time of the last integrate launch(es)       */
  float last_upload_ms;    /* CUDA-event time of the last frame upload + prep        */
  int64_t bytes_algorithmic; /* 24|16 B * written + frame bytes          */
} tsdf_stats;

/* ---- library ----------------------------------------------------------- */
int tsdf_version(void);
const char* tsdf_last_error(void);       /* thread-local text of the last failure */
int tsdf_device_count(int* n);           /* TSDF_E_NODEVICE when none             */

/* ---- configuration ----------------------------------------------------- */
/* Defaults = the reference's in-class initialisers (include/tsdf.hpp:63-67,96):
 * 200^3, 4 mm, trunc 5*vs, TUM3 K, 480x640, max_depth 6, identity base pose. */
void tsdf_config_default(tsdf_config* cfg);
/* fusion.py rule (src/TSDFfusion.py.in:23-29 -> upstream ctor):
 * dims = ceil((max-min)/voxel_size) in float64, origin = mins (as fp32),
 * trunc = 5*voxel_size, FUSIONPY semantics, colour on, max_depth = +inf.
 * vol_bnds is row-major 3x2 {xmin,xmax, ymin,ymax, zmin,zmax}.              */
int tsdf_config_from_bounds(const double vol_bnds[6], double voxel_size, tsdf_config* cfg);

/* ---- lifecycle (K3, K4) -------------------------------------------------- */
int tsdf_create(const tsdf_config* cfg, tsdf_vol** out);
/* Frees everything the handle owns.  A volume whose frame slots a tsdf_comm has exported to /
 * mapped from the other ranks (after the first tsdf_integrate_bcast_async) must outlive that
 * communicator: tsdf_destroy returns TSDF_E_INVALID and frees nothing until tsdf_comm_destroy
 * has been called.                                                              */
int tsdf_destroy(tsdf_vol* v);
int tsdf_reset(tsdf_vol* v);             /* tsdf=1, weight=0, colour=0 (src/tsdf.cu:79-81) */
int tsdf_get_config(const tsdf_vol* v, tsdf_config* cfg);

/* ---- integration (K1, K2, K8, K10) ---------------------------------------- */
/* One frame.  depth: H*W fp32; rgb: H*W*3 u8 or NULL; K: 9 floats or NULL
 * (=> config K); T_wc: 16 floats; obs_weight: 1.0f for the TSDF class.
 * cam2base = inv(base2world) * T_wc is composed on the host exactly as
 * src/tsdf.cu:142,253-403 does (skipped when base2world is the identity).
 * tsdf_integrate = tsdf_integrate_async + tsdf_sync.                       */
int tsdf_integrate(tsdf_vol* v, const float* depth, const uint8_t* rgb,
                   const float* K, const float* T_wc, float obs_weight);
int tsdf_integrate_async(tsdf_vol* v, const float* depth, const uint8_t* rgb,
                         const float* K, const float* T_wc, float obs_weight);
/* Raw sensor samples: depth_fmt TSDF_DEPTH_U16 applies d = (float)u16 * scale
 * on the device (examples/label_instance_rgbd.cpp:99-100: scale = 1.0f/5000.0f);
 * subsample_cols/rows > 1 keep only pixels with col%sc==0 && row%sr==0 and zero
 * the rest (examples/label_instance_rgbd.cpp:89-97: 3,4).                    */
int tsdf_integrate_raw_async(tsdf_vol* v, const void* depth, int depth_fmt, float depth_scale,
                             int subsample_cols, int subsample_rows, const uint8_t* rgb,
                             const float* K, const float* T_wc, float obs_weight);
/* k frames in one call (the off-line keyframe replay of
 * examples/label_instance_rgbd.cpp:79-110): the frames are staged as a group (uploads of the
 * whole group overlap the integration of the previous one) and integrated in frame order;
 * results are bit-identical to k successive tsdf_integrate calls.  Frames are staged and submitted
 * two at a time: when the slab lies inside both frames' frusta (dense frames) the pair shares ONE
 * pass over the volume (k_integrate2: 12 B of voxel traffic per update instead of 24, 3-11 % faster
 * per frame on B200 — the per-voxel arithmetic, not HBM, bounds it); sparse frames
 * and an odd last frame are one launch each.  Pair i+1 is staged while pair i integrates.
 * depth: k*H*W, rgb: k*H*W*3 or NULL, K: 9 or NULL, T_wc: k*16.                 */
int tsdf_integrate_multi_async(tsdf_vol* v, int k, const float* depth, const uint8_t* rgb,
                               const float* K, const float* T_wc, float obs_weight);
/* Waits for everything enqueued on the handle.  On a volume attached to a tsdf_comm it also
 * surfaces a failed frame distribution (a peer-pull wait that timed out because a rank is gone):
 * TSDF_E_NCCL, and the frames behind the failure were NOT integrated on this rank (the kernel
 * skips them; the volume holds exactly the frames before it).  tsdf_get_stats, tsdf_download*,
 * tsdf_extract_* and tsdf_gather_* sync first and report the same way.            */
int tsdf_sync(tsdf_vol* v);
int tsdf_get_stats(tsdf_vol* v, tsdf_stats* s); /* syncs the handle's stream */

/* ---- readback (K4, K5, K7) ------------------------------------------------ */
/* Copies this handle's slab to host arrays in the semantics' reference order.
 * global_layout=0: pointers address a dense LOCAL slab array;
 * global_layout=1: pointers address the GLOBAL Dx*Dy*Dz arrays and the slab
 * is scattered into place (lets G handles fill one host mirror).
 * Any pointer may be NULL.                                                  */
int tsdf_download(tsdf_vol* v, float* tsdf, float* weight, float* color, int global_layout);
/* Chunked readback for volumes larger than a comfortable host buffer (a 1024^3 volume is
 * 12.9 GB; the reference's destructor copies its 200^3 arrays whole, src/tsdf.cu:101-104):
 * copies n_voxels voxels starting at linear index first_voxel of this handle's LOCAL slab, in
 * the semantics' storage order (the order tsdf_download(..., 0) uses), to the start of each
 * non-NULL host array.  TSDF_E_INVALID if the range leaves the slab.                       */
int tsdf_download_range(tsdf_vol* v, int64_t first_voxel, int64_t n_voxels, float* tsdf,
                        float* weight, float* color);
int tsdf_upload(tsdf_vol* v, const float* tsdf, const float* weight, const float* color,
                int global_layout);
/* Surface/point extraction on the device (warp-ballot compaction), ascending
 * reference linear index.  xyz: 3*cap floats (origin + (float)i*vs,
 * src/tsdf.cu:206-208); rgb: 3*cap u8 {c0,c1,c2} or NULL; idx: cap int64
 * GLOBAL reference linear indices or NULL.  *count receives the number of
 * matching voxels even when cap is too small (then TSDF_E_CAPACITY and
 * nothing is written).  Call with cap=0 to size the buffers.                */
int tsdf_extract_points(tsdf_vol* v, int predicate, float tsdf_thresh, float weight_thresh,
                        float* xyz, uint8_t* rgb, int64_t* idx, int64_t cap, int64_t* count);
/* Same, outputs stay on the device (pointers are device pointers).          */
int tsdf_extract_points_device(tsdf_vol* v, int predicate, float tsdf_thresh, float weight_thresh,
                               float* d_xyz, uint8_t* d_rgb, int64_t* d_idx, int64_t cap,
                               int64_t* count);
/* Point extraction that allocates nothing.  (No counterpart in the reference: its only extraction
 * is the host loop of src/tsdf.cu:170-218.)
 *
 * Who needs it: a host that must not — or cannot — let this library allocate while it runs.  Examples:
 * a tracking / mapping loop with a frame deadline (cudaMalloc and cudaFree stall every stream of the
 * device, so a readback that grows its scratch would hiccup the integration running beside it); an
 * application that gave all device memory to its own pool or to torch's caching allocator up front, so
 * a cudaMalloc from a library simply fails; and a caller whose voxel arrays are its own device buffers
 * rather than a tsdf_vol (every other extraction call needs a handle, and tsdf_create allocates).
 *
 * How: the ballot / count / offset scratch of the compaction sits in one fixed array compiled into the
 * library's device image (about 16.2 MiB; every GPU that loads the library holds it until the process
 * exits).  The voxel arrays and the outputs are the caller's device buffers.  The call makes no
 * cudaMalloc / cudaFree / cudaHostAlloc and creates no stream or event.
 *
 *   cfg       dims, origin, voxel_size, semantics, device and (optionally) z_begin/z_end describe the
 *             arrays; the other fields are ignored.
 *   d_tsdf, d_weight, d_color   device arrays of the slab in the semantics' storage order (what
 *             tsdf_download(..., global_layout = 0) returns); d_color may be NULL (rgb = 0 then).
 *   d_xyz, d_rgb, d_idx, cap, count   as tsdf_extract_points_device; all three NULL = sizing call.
 *   stream    a cudaStream_t of cfg->device, or NULL for the default stream.  The call returns after
 *             the stream has drained (results are in the output buffers).
 *
 * Limits and rules:
 *   - at most 2^27 voxels (512^3) per call; a larger slab is refused with TSDF_E_INVALID and a message
 *     that names the limit (use a handle and tsdf_extract_points_device for those);
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, checked,
 *     cached per device).  Make that first call during set-up, before any stream capture is begun;
 *   - all calls on a device share the one array, and it keeps whatever the previous call wrote.  No
 *     part of it has to be zero on entry (each pass overwrites what the next one reads), but two calls
 *     that overlap on the same GPU — two host threads, or two streams — trample each other's scratch:
 *     serialise them.  Calls on different GPUs are independent.                                    */
int tsdf_extract_points_static(const tsdf_config* cfg, const float* d_tsdf, const float* d_weight,
                               const float* d_color, int predicate, float tsdf_thresh,
                               float weight_thresh, float* d_xyz, uint8_t* d_rgb, int64_t* d_idx,
                               int64_t cap, int64_t* count, void* stream);
/* Triangle-mesh readback: replaces fusion.TSDFVolume.get_mesh(), called by the plugin's
 * SaveMesh (src/TSDFfusion.py.in:48-53).  Level-0 marching cubes of the handle's tsdf array
 * (no weight mask, as upstream), on the device: indexed mesh, one vertex per sign-crossing grid
 * edge in ascending (storage index, axis) order, faces by ascending cell index.
 *   verts  3*vcap floats  origin + (index + t)*voxel_size, t = a/(a-b) along the edge
 *   norms  3*vcap floats  unit tsdf gradient (towards free space), or NULL
 *   colors 3*vcap u8      {c0,c1,c2} of the voxel nearest to the vertex, or NULL
 *   faces  3*fcap int32   vertex indices, triangles wound so that their normal points to free
 *                         space (tsdf > 0)
 * *n_verts / *n_faces always receive the counts; all four arrays NULL = sizing call (its sweep
 * is reused by the next call while the volume is unchanged).  TSDF_E_CAPACITY if a given array
 * is too small (nothing written).  A z-slab handle meshes its own slab; cells across a slab
 * face belong to neither side.  Vertex/face ORDER is this library's, not skimage's.          */
int tsdf_extract_mesh(tsdf_vol* v, float* verts, float* norms, uint8_t* colors, int64_t vcap,
                      int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces);
/* Same, outputs stay on the device (pointers are device pointers).          */
int tsdf_extract_mesh_device(tsdf_vol* v, float* d_verts, float* d_norms, uint8_t* d_colors,
                             int64_t vcap, int32_t* d_faces, int64_t fcap, int64_t* n_verts,
                             int64_t* n_faces);
/* fusion.meshwrite(filename, verts, faces, norms, colors) (src/TSDFfusion.py.in:53): ascii PLY
 * with x y z nx ny nz red green blue per vertex and "3 i j k" per face.  Whole-grid handles
 * only.                                                                                      */
int tsdf_save_mesh_ply(tsdf_vol* v, const char* path);
/* Byte-compatible writers: PLY header src/tsdf.cu:186-192 + float x,y,z;
 * BIN = 8-float header + tsdf array (src/tsdf.cu:116-132).  Whole-grid
 * handles only.  tsdf_load_bin restores tsdf (weights are not in the file). */
int tsdf_save_ply(tsdf_vol* v, const char* path, int predicate, float tsdf_thresh,
                  float weight_thresh);
int tsdf_save_bin(tsdf_vol* v, const char* path);
int tsdf_load_bin(tsdf_vol* v, const char* path);

/* ---- per-instance setup on the device (K9, K11) ------------- */
/* origin[k] = min over pixels with mask!=0 && depth>0 of the back-projected
 * x=(c-cx)*z*(1/fx), y=(r-cy)*z*(1/fy), z  (src/Object.cpp:37-49), initial 1000.
 * mask: H*W u8 (0 / non-zero) or NULL (= all ones).  depth/mask may be host or
 * device pointers.                                                          */
int tsdf_instance_origin(int device, const float* depth, const uint8_t* mask, int im_h, int im_w,
                         const float K[9], float origin_out[3]);

/* ---- batched per-instance volumes (config 5) ------------------------------ */
/* n volumes that share one frame per integrate call; each has its own base
 * pose, origin and 8-bit mask (masked depth = depth * (mask/255),
 * src/Engine.cpp:192-193).  One launch covers all active instances.         */
int tsdf_batch_create(const tsdf_config* cfgs, int n, tsdf_batch** out);
int tsdf_batch_destroy(tsdf_batch* b);
int tsdf_batch_size(const tsdf_batch* b, int* n);
/* active: n_active instance ids; masks: n_active * H*W u8 (host or device).  */
int tsdf_batch_integrate_async(tsdf_batch* b, const float* depth, const uint8_t* rgb,
                               const int32_t* active, int n_active, const uint8_t* masks,
                               const float* K, const float* T_wc, float obs_weight);
/* Same, with the masks cropped to their Mask-RCNN bounding boxes (boxes[i], reference
 * src/Engine.cpp:127,197): boxes = n_active x {x, y, w, h} pixels inside the image; mask_crops =
 * the n_active crops concatenated, crop i is h_i rows of w_i bytes.  Pixels outside a box count
 * as mask 0.  Uploads sum(w*h) bytes instead of n_active*H*W; results are identical to the
 * full-frame call.                                                              */
int tsdf_batch_integrate_boxes_async(tsdf_batch* b, const float* depth, const uint8_t* rgb,
                                     const int32_t* active, int n_active, const int32_t* boxes,
                                     const uint8_t* mask_crops, const float* K, const float* T_wc,
                                     float obs_weight);
int tsdf_batch_sync(tsdf_batch* b);
int tsdf_batch_get_stats(tsdf_batch* b, tsdf_stats* s);
int tsdf_batch_download(tsdf_batch* b, int instance, float* tsdf, float* weight, float* color);
int tsdf_batch_extract_points(tsdf_batch* b, int instance, int predicate, float tsdf_thresh,
                              float weight_thresh, float* xyz, uint8_t* rgb, int64_t* idx,
                              int64_t cap, int64_t* count);
/* Mesh of one instance volume (tsdf_extract_mesh on it): the per-object surface of Engine/Object. */
int tsdf_batch_extract_mesh(tsdf_batch* b, int instance, float* verts, float* norms, uint8_t* colors,
                            int64_t vcap, int32_t* faces, int64_t fcap, int64_t* n_verts,
                            int64_t* n_faces);

/* ---- multi-GPU: z-slabs + NCCL -------------------------------- */
/* One tsdf_comm per rank (one process per GPU, or G handles in one process).
 * uid: 128 bytes from tsdf_comm_unique_id on one rank, shipped to the others
 * by the caller (torch.distributed store, MPI, a file...).                  */
int tsdf_comm_unique_id(void* uid128);
int tsdf_comm_create(const void* uid128, int nranks, int rank, int device, tsdf_comm** out);
int tsdf_comm_destroy(tsdf_comm* c);
/* Slab of `rank` out of `nranks` for a grid with dim_z slices: contiguous,
 * sizes differ by at most one 4-slice group (keeps 128-bit alignment).      */
void tsdf_slab_range(int dim_z, int nranks, int rank, int* z_begin, int* z_end);
/* Collective over all ranks of `c` (SPMD: every rank makes the same sequence of calls on its
 * handle): the root's images (depth, rgb; ignored elsewhere, may be NULL there) reach every
 * rank over NVLink — by peer-pull fused into the prep kernel, else by a grouped ncclBroadcast
 * into every rank's frame slot — then every rank integrates its own slab.  The ~100 bytes of metadata (K, T_wc,
 * obs_weight) travel on the caller's control plane: pass the SAME values on every rank.
 * Colour is broadcast iff the volume was created with_color (then the root must pass rgb). */
int tsdf_integrate_bcast_async(tsdf_vol* v, tsdf_comm* c, int root, const float* depth,
                               const uint8_t* rgb, const float* K, const float* T_wc,
                               float obs_weight);
/* *active = 1 when frames are distributed by peer-pull (every non-root rank's prep kernel reads
 * the root's raw frame slot through NVLink, cudaIpc-mapped; set up by the first
 * tsdf_integrate_bcast_async call), 0 when ncclBroadcast is used (fallback, or
 * TSDF_B200_BCAST=nccl in the environment).  Results are identical either way.  Once peer-pull is
 * active the volume is ATTACHED to the communicator: destroy the communicator first (tsdf_destroy
 * refuses until then).  Calls of this function and plain tsdf_integrate* calls may be interleaved
 * on the handle (the distribution has its own frame ring, indexed by its own sequence number).  */
int tsdf_comm_peer_pull(const tsdf_comm* c, int* active);
/* Collective: every rank extracts its slab; the root receives all points,
 * merged to ascending GLOBAL reference index.  Non-root ranks pass NULL/0
 * buffers and get their LOCAL count in *count; the root gets the total.
 * cap < 0 on EVERY rank = count-only (sizing) call: no records are moved.    */
int tsdf_gather_points(tsdf_vol* v, tsdf_comm* c, int root, int predicate, float tsdf_thresh,
                       float weight_thresh, float* xyz, uint8_t* rgb, int64_t* idx, int64_t cap,
                       int64_t* count);

/* Collective mesh readback of a sharded volume (plugin SaveMesh on the multi-GPU configuration,
 * src/TSDFfusion.py.in:48-53): cells across a slab face need both sides, so every rank sends its
 * tsdf (and colour) slab to the root over NVLink, the root assembles a temporary WHOLE-grid volume
 * on its device (12 B per voxel of the whole grid) and meshes it there — vertices, faces and
 * their order are exactly those of tsdf_extract_mesh on a single-GPU volume.  The assembly is
 * reused while no rank's slab changes (sizing call + real call move the volume once) and freed by
 * tsdf_comm_destroy.  Arguments as tsdf_extract_mesh; non-root ranks pass NULL/0 and get 0 counts. */
int tsdf_gather_mesh(tsdf_vol* v, tsdf_comm* c, int root, float* verts, float* norms, uint8_t* colors,
                     int64_t vcap, int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces);

/* ---- multi-GPU in ONE process: a device list behind one handle --------------- */
/* The reference is a single process (Engine::Run, src/Engine.cpp:105).  tsdf_group cuts the grid of
 * `cfg` (whole-grid config; cfg->device is ignored) into n_devices z-slabs, slab i on devices[i]
 * (tsdf_slab_range), and drives them from the calling thread: the frame is staged once on
 * devices[0]; every other slab's prep kernel reads it in place over NVLink
 * (cudaDeviceEnablePeerAccess; cudaMemcpyPeerAsync where two devices have no peer access), ordered by
 * CUDA events — no NCCL, no second process.  Results are bit-identical to one whole-grid volume.
 * A device may be listed more than once (several slabs on one GPU).               */
int tsdf_group_create(const tsdf_config* cfg, const int32_t* devices, int n_devices, tsdf_group** out);
int tsdf_group_destroy(tsdf_group* g);
int tsdf_group_size(const tsdf_group* g, int* n_slabs);
int tsdf_group_slab(tsdf_group* g, int i, tsdf_vol** slab);        /* borrowed: do not destroy */
int tsdf_group_peer_access(const tsdf_group* g, int* n_direct);    /* slabs 1.. that read devices[0] in place */
/* Arguments as tsdf_integrate / tsdf_integrate_async (host pointers pageable or pinned; device
 * pointers must be accessible from devices[0]).                                  */
int tsdf_group_integrate(tsdf_group* g, const float* depth, const uint8_t* rgb, const float* K,
                         const float* T_wc, float obs_weight);
int tsdf_group_integrate_async(tsdf_group* g, const float* depth, const uint8_t* rgb, const float* K,
                               const float* T_wc, float obs_weight);
int tsdf_group_sync(tsdf_group* g);
int tsdf_group_reset(tsdf_group* g);
int tsdf_group_get_stats(tsdf_group* g, tsdf_stats* s);            /* sums over the slabs (ms: max) */
/* GLOBAL Dx*Dy*Dz host arrays in the semantics' reference order; any pointer may be NULL. */
int tsdf_group_download(tsdf_group* g, float* tsdf, float* weight, float* color);
/* As tsdf_extract_points on one whole-grid volume: ascending GLOBAL reference index (the slabs'
 * runs are merged on devices[0]).                                                 */
int tsdf_group_extract_points(tsdf_group* g, int predicate, float tsdf_thresh, float weight_thresh,
                              float* xyz, uint8_t* rgb, int64_t* idx, int64_t cap, int64_t* count);
/* As tsdf_extract_mesh on one whole-grid volume (assembled on devices[0], see tsdf_gather_mesh;
 * kept until a slab changes or tsdf_group_release_scratch).                       */
int tsdf_group_extract_mesh(tsdf_group* g, float* verts, float* norms, uint8_t* colors, int64_t vcap,
                            int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces);
int tsdf_group_release_scratch(tsdf_group* g);
int tsdf_group_save_ply(tsdf_group* g, const char* path, int predicate, float tsdf_thresh, float weight_thresh);
int tsdf_group_save_bin(tsdf_group* g, const char* path);
int tsdf_group_save_mesh_ply(tsdf_group* g, const char* path);

/* ---- pinned host memory helpers (frame upload path, readback) -------------- */
/* cudaHostAlloc: uploads from such memory skip the staging copy, and the readback calls
 * (tsdf_extract_points / _mesh, tsdf_gather_*, tsdf_group_extract_*) copy straight into output
 * arrays allocated with it (pageable outputs go through a pinned staging ring).   */
int tsdf_host_alloc(void** p, size_t bytes);
int tsdf_host_free(void* p);
int tsdf_device_alloc(int device, void** p, size_t bytes);
int tsdf_device_free(int device, void* p);
int tsdf_memcpy_h2d(int device, void* dst, const void* src, size_t bytes);
int tsdf_memcpy_d2h(int device, void* dst, const void* src, size_t bytes);

/* ---- host-side helpers restated from the reference (K6) -------------------- */
void tsdf_multiply_matrix(const float m1[16], const float m2[16], float out[16]); /* src/tsdf.cu:253-273 */
int tsdf_invert_matrix(const float m[16], float out[16]); /* src/tsdf.cu:276-403; TSDF_E_SINGULAR */

/* Pose ingest for the off-line replay: reads the keyframe poses of a saved run
 * from `bundle.txt` — header "<n_cams> <n_pts>", then per camera one intrinsics line, three
 * rows of Rcw and one row of tcw (written by src/ORB_SLAM2/System.cc:884-917, parsed by
 * examples/label_instance_rgbd.cpp:289-330) — and returns what Engine::Run feeds the TSDF:
 * Twc = KF->GetPoseInverse() (src/Engine.cpp:222), i.e. [Rcw^T | -Rcw^T tcw], row-major 4x4,
 * fp32.  Twc: cap*16 floats or NULL; *n_cams receives the camera count of the file (call with
 * cap = 0 to size).  TSDF_E_IO on open/parse failure, TSDF_E_CAPACITY if cap is too small.  */
int tsdf_read_bundle_poses(const char* path, float* Twc, int cap, int* n_cams);

/* Pose ingest for the stereo / lidar run: `CameraTrajectory.txt` (SaveTrajectoryKITTI,
 * src/ORB_SLAM2/System.cc:455-508, written by examples/stereo_kitti.cpp:113) — one line per tracked
 * frame, 12 values = the 3x4 row-major [Rwc | twc] relative to the first keyframe, i.e. already the
 * camera-to-world pose Integrate takes.  Twc: cap*16 floats (fourth row 0 0 0 1 added) or NULL;
 * *n_poses receives the line count (cap = 0 to size).  TSDF_E_IO / TSDF_E_CAPACITY as above.     */
int tsdf_read_trajectory_kitti(const char* path, float* Twc, int cap, int* n_poses);
/* `frame_id.txt` (SaveFrameId, src/ORB_SLAM2/System.cc:1104-1121, examples/stereo_kitti.cpp:117):
 * the sequence frame number of every good keyframe, one per line, in bundle.txt's camera order. */
int tsdf_read_frame_ids(const char* path, int64_t* ids, int cap, int* n_ids);

/* Image-name ingest for the same replay: `associations.txt` of a saved run (written by
 * src/ORB_SLAM2/System.cc:981; one line per keyframe "<t_rgb> <rgb file> <t_depth> <depth file>",
 * parsed by LoadImageNames, examples/label_instance_rgbd.cpp:239-268, which pairs line i with
 * keyframe i of bundle.txt, :79-87).  Empty lines are skipped like the reference does.
 * t_rgb / t_depth: cap doubles each or NULL (the reference reads the stamps into a float and
 * drops them; here they are kept, in double).  rgb_names / depth_names: cap * name_stride
 * bytes each or NULL, entry i NUL-terminated at i * name_stride.  *n_lines receives the number
 * of non-empty lines (call with cap = 0 to size).  TSDF_E_IO on open/parse failure,
 * TSDF_E_CAPACITY if cap is too small or a name does not fit name_stride.  Decoding the PNGs
 * stays with the caller (cv::imread in the reference) — this library takes raw pixels.  */
int tsdf_read_associations(const char* path, double* t_rgb, double* t_depth, char* rgb_names,
                           char* depth_names, int name_stride, int cap, int* n_lines);

/* ---- measurement hooks (used by bench.py; no effect on results) ------------ */
/* Device-side stopwatch on the handle's compute stream (CUDA events): start
 * records now; stop records after everything enqueued so far, waits for it
 * and returns the elapsed ms.  bench.py brackets its timed region with it.  */
int tsdf_timer_start(tsdf_vol* v);
int tsdf_timer_stop(tsdf_vol* v, float* ms);
/* Number of kernels this library has launched in this process so far (every <<<>>> of
 * libtsdf_b200.so, all handles, all devices): the bench's `gpu_launches` is a difference of two
 * reads around its timed window.                                              */
int tsdf_launch_count(int64_t* n);
/* Per-frame timeline of a handle: when enabled every integrate call (single, distributed, group)
 * records four CUDA events {copy stream: first operation of the frame enqueued, prepared frame
 * ready; compute stream: integrate kernel begins, ends}.  tsdf_trace_read returns them as ms
 * relative to the first event since enable (ms4: 4 floats per frame), the evidence behind the
 * multi-GPU step analysis in DESIGN.md §5.                                        */
int tsdf_trace_enable(tsdf_vol* v, int enable);
int tsdf_trace_read(tsdf_vol* v, float* ms4, int cap_frames, int* n_frames);
/* Per-launch timing of the integrate kernel: when enabled every launch is
 * bracketed by its own CUDA event pair on the launching stream; read returns
 * the summed kernel ms and the number of launches since the last reset.     */
int tsdf_kernel_timing(tsdf_vol* v, int enable);
int tsdf_kernel_timing_read(tsdf_vol* v, double* sum_ms, int64_t* launches, int reset);

#ifdef TSDF_B200_DEV
/* ---- development hooks: ONLY in libtsdf_b200_dev.so (built with -DTSDF_B200_DEV) ------------
 * The product library does not export them and carries no A/B kernel instantiations.          */
/* Times `iters` back-to-back launches of the integrate kernel on the frame
 * last uploaded to the handle (volume state keeps evolving) with CUDA events
 * on the handle's stream; returns average ms per launch.                    */
int tsdf_bench_kernel(tsdf_vol* v, int iters, float* avg_ms);
/* Selects a kernel variant for A/B evidence (0 = what the product runs); bits in
 * csrc/integrate.cu launch_integrate and csrc/mesh.cu.                     */
int tsdf_set_variant(tsdf_vol* v, int variant);
/* Validation hook: brute-forces the exact-arithmetic forms of the hot loop (shared
 * reciprocal + Markstein division, round-down magic rounding, colour blends) against the
 * IEEE operators on the device.  fails[0..4] = mismatches of {reciprocal, quotient, rounding,
 * general colour blend, integer-weight colour blend (exhaustive)}; all must be 0.  */
int tsdf_selftest_exactmath(int device, int64_t iters_per_thread, uint64_t seed, int64_t fails[5]);
/* Work counters of the sparse integrate kernel since the last enable: out[0..6] = {blocks, blocks
 * skipped by the block-level hull test, warp row segments visited, segments culled by the warp test,
 * runs (4 voxels) projected, runs gathered, runs with an update}; enable != 0 (re)arms them.   */
int tsdf_debug_counters(tsdf_vol* v, int enable, uint64_t* out);
#endif /* TSDF_B200_DEV */

#ifdef __cplusplus
}
#endif
#endif /* TSDF_B200_H */
