// common.cuh — internal types shared by the translation units of libtsdf_b200.so.
// Nothing here is part of the ABI; the ABI is include/tsdf_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/tsdf_b200.h"

// ---- error plumbing -----------------------------------------------------------------
void tsdf_set_error(const char* fmt, ...);  // thread-local, printf style
#define CU_CHECK(expr)                                                                    \
  do {                                                                                    \
    cudaError_t e__ = (expr);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      cudaGetLastError(); /* clear it: a later launch check must not report this failure */ \
      tsdf_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return e__ == cudaErrorMemoryAllocation ? TSDF_E_NOMEM : TSDF_E_CUDA;               \
    }                                                                                     \
  } while (0)
// every kernel launch site ends with this: counts the launch (tsdf_launch_count, the bench's
// `gpu_launches`) and picks up a launch-configuration error
extern std::atomic<long long> g_tsdf_launches;
#define CU_LAUNCHED()                                           \
  do {                                                          \
    g_tsdf_launches.fetch_add(1, std::memory_order_relaxed);    \
    CU_CHECK(cudaGetLastError());                               \
  } while (0)
#define REQUIRE(cond, msg)                                  \
  do {                                                      \
    if (!(cond)) {                                          \
      tsdf_set_error("%s: %s", __func__, msg);              \
      return TSDF_E_INVALID;                                \
    }                                                       \
  } while (0)

struct DeviceGuard {  // restores the caller's current device
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// ---- kernel-side parameter blocks (passed by value => constant bank) ------------------
constexpr int kMaxFusedFrames = 8;

struct FrameP {       // one camera frame
  float c[12];        // cam2base rows 0..2 (row-major 3x4), src/tsdf.cu:33-38
  float fx, cx, fy, cy;  // K[0],K[2],K[4],K[5], src/tsdf.cu:41-42
  float obs_weight;
  int pad_;
  const int* skip;    // device word that is non-zero when this frame must NOT be integrated (a peer-pull
                      // wait timed out: the prepared frame is stale), or null — see comm.cu
  int mb_x0, mb_y0, mb_x1, mb_y1;  // pixel rectangle (inclusive) outside which the frame is known to be
                                   // invalid for this launch: the instance's mask box, else the image
  const void* h_tmap; // HOST pointer: CUtensorMap of the prepared frame (128 B) or null; read by the launcher only
  const void* img;    // prepared frame: float depth[H*W]  or  float4 {depth, c0, c1, c2}[H*W]
  const float* dmax;  // largest depth of the prepared frame (+inf when it holds a NaN)
};

struct GridP {        // one volume slab
  int n_fast;         // voxels along the fast (contiguous) axis of the slab
  int n_mid;          // = dim_y
  int n_slow;         // slow axis count of the slab
  int z0;             // global z of local slice 0
  int nx_, nz_;       // dim_x and local slices (host-side use)
  float ox, oy, oz;   // origin (base frame)
  float vs, trunc;
  float inv_trunc;    // RN(1/trunc), host IEEE division
  float w_lim, h_lim; // W-0.5, H-0.5: (int)roundf(u) in [0,W) <=> -0.5 < u < W-0.5
  int W, H;
  float* tsdf;
  float* weight;
  float* color;                  // may be null
  unsigned long long* written;   // device counter (exact N_written)
  unsigned long long* dbg;       // dev build only: 8 work counters of the sparse kernel (tsdf_debug_counters), else null
};

struct BatchDesc {    // one active instance of a batched launch (lives in global memory)
  GridP g;
  FrameP f;
  const uint8_t* mask;  // H*W 8-bit instance mask on the device, or null
};

// Prepared-frame pixel when colour is on: float4 {depth (invalid canonicalised to 0.0f),
// c0, c1, c2 as exact floats}.
// Depth-only frames are a plain float image.

// ---- host-side handle -------------------------------------------------------------------
// Grow-only buffers of the readback calls (points, mesh, gather): device arrays the emit kernels
// write and one pinned host staging area for pageable destinations; kept between calls by their owner
// (the reference-style "count, allocate, emit" sequence must not pay cudaMalloc/cudaFree per call).
struct OutBuffers {
  void* d[4] = {nullptr, nullptr, nullptr, nullptr};
  size_t dcap[4] = {0, 0, 0, 0};
  void* h = nullptr;       // pinned
  size_t hcap = 0;
  cudaEvent_t ev[2] = {nullptr, nullptr};   // chunk-landed events of the staged copy (device of first use)
  int ev_dev = -1;
  int need_dev(int i, size_t bytes);    // TSDF status; contents are NOT preserved when it grows
  int need_host(size_t bytes);
  void release();
};
// Device -> caller's host array on `st`: straight cudaMemcpyAsync when dst is pinned, else through
// the pinned staging area in double-buffered chunks (then the data is in dst on return).  A pinned
// destination is only ENQUEUED: the caller synchronises `st` after its last copy.
int copy_to_host(OutBuffers* ob, void* dst, const void* d_src, size_t bytes, cudaStream_t st);

struct FrameSlot {
  void* h_pinned = nullptr;   // staging for pageable inputs (depth + rgb)
  void* d_raw = nullptr;      // raw upload target (depth fp32|u16 + rgb u8)
  void* d_frame = nullptr;    // prepared frame consumed by the integrate kernel
  alignas(64) unsigned char tmap[128];  // CUtensorMap over d_frame as [H][4W] floats, box 64 x 32 pixels (colour frames)
  bool tmap_ok = false;
  cudaEvent_t uploaded = nullptr;  // recorded on the copy stream after prep
  cudaEvent_t consumed = nullptr;  // recorded on the compute stream after integrate
  bool in_flight = false;
};

constexpr int kWrittenRing = 4;
constexpr int kBcastSlots = 6;  // frame ring of the distributed (multi-GPU) integrate calls

struct tsdf_comm;
struct tsdf_vol {
  tsdf_config cfg;
  int dzl = 0;                 // local slices
  int64_t n_local = 0;         // voxels in this slab
  bool zfast = false;          // FUSIONPY layout [x][y][zl]; else TSDFCU layout [zl][y][x]
  bool identity_base = false;  // skip the host pose composition
  float base2world_inv[16];
  float* d_tsdf = nullptr;
  float* d_weight = nullptr;
  float* d_color = nullptr;
  // exact N_written: a ring of kWrittenRing device counters, launch n adds into counter n % ring; the
  // publish kernel of launch n runs on the AUX stream behind it (moves the count to the pinned mirror,
  // re-arms the counter, zeroes the consumed slot's max-depth tiles), so consecutive integrate
  // launches follow each other on the compute stream with nothing between them
  unsigned long long* d_written = nullptr;
  uint64_t lseq = 0;                        // integrate launch groups submitted
  cudaStream_t s_aux = nullptr;
  cudaEvent_t ev_int_done[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_pub_done[4] = {nullptr, nullptr, nullptr, nullptr};
  unsigned long long* h_written = nullptr;  // pinned mirror (mapped: the kernel writes it)
  unsigned long long* hd_written = nullptr; // its device address
  cudaStream_t s_compute = nullptr;
  cudaStream_t s_copy = nullptr;
  cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_u0 = nullptr, ev_u1 = nullptr;
  // slots 0..2: single-frame ring; 3..3+2*kMaxFusedFrames: two groups for fused launches
  FrameSlot slot[3 + kMaxFusedFrames * 2];
  // ring of the distributed integrate calls (tsdf_integrate_bcast_async, tsdf_group_*): indexed by
  // the distribution sequence number, never by `seq`, so single-frame calls may be interleaved
  FrameSlot bslot[kBcastSlots];
  // set while a communicator has this volume's frame slots exported / mapped: tsdf_destroy refuses
  // (the peers may still read them), tsdf_sync surfaces *comm_err (a timed-out peer-pull wait)
  tsdf_comm* attached = nullptr;
  const int* comm_err = nullptr;   // pinned host word of the attached communicator
  const int* comm_skip = nullptr;  // its device twin (FrameP::skip)
  uint64_t seq = 0;            // single-frame submissions
  uint64_t mseq = 0;           // fused submissions
  int64_t frames = 0;
  int64_t last_written = 0;
  int64_t last_frames_in_launch = 1;
  int variant = 0;
  unsigned long long* d_dbg = nullptr;   // dev build: work counters of the sparse kernel
  bool dev_hold = false;       // set by tsdf_set_variant: prepared frames stay intact after their launch
  // integer-weight bookkeeping: true while every stored weight is provably an integer
  // (only obs_weight == 1 integrations since the last reset); w_bound >= max stored weight
  bool intw = true;
  int64_t w_bound = 0;
  bool wmode_ok = false;       // decided per launch by submit()
  // state of the last launch, for tsdf_bench_kernel
  FrameP last_frames[kMaxFusedFrames];
  int last_nf = 0;
  bool last_color = false;
  // measurement
  cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
  bool ktiming = false;
  std::vector<cudaEvent_t> kev;   // pairs
  size_t kev_used = 0;            // events recorded since reset (2 per launch)
  // per-frame timeline (tsdf_trace_*): 4 events per frame {copy stream: first op enqueued, prepared
  // frame ready; compute stream: integrate begins, integrate ends}
  bool tracing = false;
  std::vector<cudaEvent_t> tev;
  size_t tev_used = 0;
  OutBuffers out;              // readback outputs (tsdf_extract_points / tsdf_extract_mesh)
  // extraction scratch
  void* d_scratch = nullptr;
  size_t scratch_bytes = 0;
  // every operation that can change the voxel arrays bumps `epoch`; the mark pass of the last
  // extraction is reused while epoch and predicate are unchanged
  uint64_t epoch = 0;
  bool mark_valid = false;
  uint64_t mark_epoch = 0;
  int mark_pred = 0;
  float mark_tt = 0, mark_wt = 0;
  int64_t mark_total = 0;
  // mesh readback scratch (mesh.cu): records, case bytes and offsets of the last mark pass, reused
  // by the call that follows a sizing call while `epoch` is unchanged
  void* d_mesh = nullptr;
  size_t mesh_bytes = 0;
  bool mesh_valid = false;
  uint64_t mesh_epoch = 0;
  int64_t mesh_nv = 0, mesh_nf = 0;
  uint32_t mesh_vtiles = 0, mesh_ftiles = 0;  // tiles that hold vertices / triangles
};

inline size_t frame_pixels(const tsdf_config& c) { return (size_t)c.im_h * c.im_w; }
// A prepared frame is px*16 bytes of pixels followed by a tail written by k_frame_prep:
//   word 0            the frame's largest depth (float bits; +inf when the frame holds a NaN)
//   words 16 ...      max-depth tiles, level 0: 16x16 px tiles, row-major tw0 x th0
//   then              max-depth tiles, level 1: 64x64 px tiles, row-major tw1 x th1
// The integrate kernel's segment cull reads them ("everything this segment can project to is
// nearer than the segment, or invalid").
constexpr int kTile0Log2 = 4, kTile1Log2 = 6, kTailHeadWords = 16;
__host__ __device__ inline int tiles_across(int n, int log2) { return (n + (1 << log2) - 1) >> log2; }
inline size_t frame_tail_bytes(int H, int W) {
  size_t words = kTailHeadWords + (size_t)tiles_across(W, kTile0Log2) * tiles_across(H, kTile0Log2) +
                 (size_t)tiles_across(W, kTile1Log2) * tiles_across(H, kTile1Log2);
  return (words * 4 + 63) & ~(size_t)63;
}
inline size_t frame_alloc_bytes(int H, int W) { return (size_t)H * W * 16 + frame_tail_bytes(H, W); }
inline float* frame_dmax_ptr(const void* d_frame, size_t px) {
  return reinterpret_cast<float*>(reinterpret_cast<char*>(const_cast<void*>(d_frame)) + px * 16);
}

// Peer-pull distribution fused into the prep kernel (comm.cu); all zero = plain prep.
struct PrepSync {
  const unsigned* wait_flag = nullptr;  // peer: spin until *wait_flag >= wait_target before reading the source
  int wait_target = 0;
  unsigned* ready_flag = nullptr;       // root: store ready_value at the head (its source is already in place:
  unsigned ready_value = 0;             //       the H2D copies precede this kernel in stream order)
  unsigned* done_flag = nullptr;        // peer: the last block to finish stores done_value (source consumed)
  unsigned done_value = 0;
  int ring = 0;                         // peer: slots in the root's ring; `ready` >= wait_target + ring at the
                                        //       tail means the slot was overwritten while being read
  unsigned* arrive = nullptr;           // device counter for the last-block detection (re-arms itself)
  int* d_err = nullptr;                 // set to 1 on timeout (device word: FrameP::skip of this rank) ...
  int* h_err = nullptr;                 // ... and its pinned host twin, read by the host API
  unsigned long long timeout_ns = 0;
};

// ---- launchers implemented in the .cu files -----------------------------------------------
// integrate.cu
int launch_integrate(tsdf_vol* v, const FrameP* frames, int nf, bool color, cudaStream_t st);
// N_written -> pinned mirror, counter = 0; optionally zeroes the tail (max-depth tiles) of the frame
// slot the launch before it consumed, so that the next prep into that slot needs no memset
int launch_publish(tsdf_vol* v, bool accumulate, cudaStream_t st, void* d_frame_to_clear = nullptr);  // counter lseq % ring
int launch_integrate_batch(const BatchDesc* d_descs, int n_active, int max_fast, int max_rows,
                           bool zfast, bool color, bool vec4, int wmode, cudaStream_t st);
void fill_grid_params(const tsdf_vol* v, bool color, GridP* g);
bool account_weights(tsdf_vol* v, const FrameP* fp, int nf);  // updates intw/w_bound, returns wmode_ok
// frame.cu
int launch_frame_prep(const void* d_depth, int depth_fmt, float depth_scale, int sub_c, int sub_r,
                      const uint8_t* d_rgb, const uint8_t* d_mask, int H, int W, float max_depth,
                      bool with_color, const float intr[4] /* fx,cx,fy,cy: RANGE formats only */,
                      void* d_frame, cudaStream_t st, bool clear_tail = true, const PrepSync* sync = nullptr);
int launch_fill(float* p, float val, int64_t n, cudaStream_t st);
int launch_instance_origin(const float* d_depth, const uint8_t* d_mask, int H, int W, float fx,
                           float cx, float fy, float cy, float* d_out3, cudaStream_t st);
// extract.cu
int extract_device(tsdf_vol* v, int predicate, float tsdf_thresh, float weight_thresh, float* d_xyz,
                   uint8_t* d_rgb, int64_t* d_idx, int64_t cap, int64_t* count);
int ensure_scratch(tsdf_vol* v, size_t bytes);
// runs of records (run r = [offs[r], offs[r+1]), each ascending by index, indices distinct) -> one
// ascending sequence; d_offs: nranks + 1 device entries
int launch_merge_scatter(int64_t total, int nranks, const int64_t* d_offs, const float* g_xyz, const int64_t* g_idx,
                         const uint8_t* g_rgb, float* o_xyz, int64_t* o_idx, uint8_t* o_rgb, cudaStream_t st);
int launch_scan(const uint32_t* d_cnt, int64_t m, int64_t* d_off, int64_t* d_tot, cudaStream_t st,  // off[m] = total
                const uint32_t* d_cnt_b = nullptr, int64_t* d_off_b = nullptr,                   // second triple,
                int64_t* d_tot_b = nullptr,                                                       // same m
                uint32_t* d_list = nullptr, uint32_t* d_list_b = nullptr,  // optional: indices of the non-zero counts
                uint32_t* d_nlist2 = nullptr);                             // their two lengths (zeroed by the caller)
size_t scan_scratch_bytes(int64_t m);  // size of one d_tot
// mesh.cu
int mesh_device(tsdf_vol* v, float* d_verts, float* d_norms, uint8_t* d_cols, int64_t vcap,
                int32_t* d_faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces);

// api.cpp helpers used elsewhere
int write_points_ply(const char* path, const float* xyz, int64_t n);                       // src/tsdf.cu:186-216
int write_bin(const char* path, const tsdf_config& c, const float* tsdf, size_t n);        // src/tsdf.cu:116-132
int write_mesh_ply(const char* path, const float* verts, const float* norms, const uint8_t* cols, int64_t nv,
                   const int32_t* faces, int64_t nf);                                      // fusion.meshwrite
int trace_mark(tsdf_vol* v, int which, cudaStream_t st);   // no-op unless tsdf_trace_enable(v, 1)
int check_obs_weight(float obs_weight);  // one validation for every integrate entry point
void host_copy(void* dst, const void* src, size_t bytes);   // memcpy, split over a few helper threads when large
int ensure_bcast_slot(tsdf_vol* v, int index);
// moves `bytes` of a frame (pageable / pinned host or device memory) to where the prep kernel reads it
int frame_to_device(tsdf_vol* v, FrameSlot& s, const void* src, size_t bytes, size_t raw_off,
                    const void** d_out, bool into_slot);
// waits for the slots' uploads, launches the integrate kernel(s) + publish, records the events
int submit_frames(tsdf_vol* v, FrameSlot** slots, const FrameP* fp, int nf, bool color, bool accumulate = false);
int compose_cam2base(const tsdf_vol* v, const float* T_wc, float* c2b);
int make_frame_params(const tsdf_vol* v, const float* K, const float* T_wc, float obs_weight,
                      const void* d_frame, FrameP* out);
int stage_frame(tsdf_vol* v, int slot_index, const void* depth, int depth_fmt, float depth_scale,
                int sub_c, int sub_r, const uint8_t* rgb, bool want_color, FrameSlot** out_slot,
                const float* K = nullptr, bool clear_tail = true);
