// batch.cu — batched per-instance object volumes (BASELINE config 5).
//
// In the reference every Mask-RCNN instance owns a `TSDF*` (include/Object.hpp:216-218) and
// Engine::Run would, per mask, build a full-frame masked depth image
// (src/Engine.cpp:192-193) and call Object::Integrate -> TSDF::Integrate, i.e. one blocking
// 1.2 MB upload and one <<<Dz,Dy>>> launch per instance per keyframe (src/tsdf.cu:161-167).
// Here the frame is uploaded ONCE, the 8-bit masks of the active instances are uploaded as
// one block, and ONE launch (blockIdx.z = instance) integrates all of them; the mask test
// `depth * (mask/255)` is applied at gather time inside the kernel.
#include <new>

#include "common.cuh"

struct MaskBox { int x, y, w, h; long long off; };  // bounding box of a cropped mask + its offset in the crop block

namespace {
// Boxes path: the host uploads only the mask pixels inside each instance's bounding box
// (Mask-RCNN gives boxes[i] next to masks[i], reference src/Engine.cpp:127,197); this kernel
// rebuilds the full-frame 8-bit masks the integrate kernel gathers from (zero outside the box).
__global__ void k_expand_masks(const uint8_t* __restrict__ crops, const MaskBox* __restrict__ boxes,
                               int H, int W, uint8_t* __restrict__ masks) {
  const MaskBox bx = boxes[blockIdx.y];
  uint8_t* out = masks + (size_t)blockIdx.y * H * W;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < H * W; p += gridDim.x * blockDim.x) {
    const int r = p / W, c = p - r * W;
    const bool in = r >= bx.y && r < bx.y + bx.h && c >= bx.x && c < bx.x + bx.w;
    out[p] = in ? crops[bx.off + (long long)(r - bx.y) * bx.w + (c - bx.x)] : (uint8_t)0;
  }
}
}  // namespace

struct tsdf_batch {
  int n = 0;
  std::vector<char> seen;        // scratch of the duplicate-id check
  std::vector<tsdf_vol*> vols;   // per-instance state (arrays, config, base pose); own streams unused
  tsdf_vol* frame_owner = nullptr;  // vols[0]: its slots/streams carry the shared frame
  BatchDesc* d_descs = nullptr;
  BatchDesc* h_descs = nullptr;  // pinned
  uint8_t* d_masks = nullptr;    // n * H*W
  uint8_t* h_masks = nullptr;    // pinned staging, n * H*W
  uint8_t* d_crops = nullptr;    // cropped masks as uploaded (boxes path), <= n * H*W
  MaskBox* d_boxes = nullptr;
  MaskBox* h_boxes = nullptr;    // pinned
  unsigned long long* d_written = nullptr;
  unsigned long long* h_written = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, masks_free = nullptr;
  bool masks_in_flight = false;
  int last_active = 0;
  int64_t last_visited = 0;
  bool last_color = false;
};

extern "C" int tsdf_batch_destroy(tsdf_batch* b) {
  if (!b) return TSDF_OK;
  if (b->frame_owner) {
    DeviceGuard g(b->frame_owner->cfg.device);
    cudaStreamSynchronize(b->frame_owner->s_compute);
    cudaStreamSynchronize(b->frame_owner->s_copy);
    cudaFree(b->d_descs); cudaFree(b->d_masks); cudaFree(b->d_written); cudaFree(b->d_crops); cudaFree(b->d_boxes);
    if (b->h_boxes) cudaFreeHost(b->h_boxes);
    if (b->h_descs) cudaFreeHost(b->h_descs);
    if (b->h_masks) cudaFreeHost(b->h_masks);
    if (b->h_written) cudaFreeHost(b->h_written);
    for (cudaEvent_t e : {b->ev0, b->ev1, b->masks_free}) if (e) cudaEventDestroy(e);
  }
  for (tsdf_vol* v : b->vols) tsdf_destroy(v);
  delete b;
  return TSDF_OK;
}

extern "C" int tsdf_batch_create(const tsdf_config* cfgs, int n, tsdf_batch** out) {
  REQUIRE(cfgs && out && n >= 1, "bad argument");
  *out = nullptr;
  for (int i = 1; i < n; ++i) {
    REQUIRE(cfgs[i].semantics == cfgs[0].semantics && cfgs[i].with_color == cfgs[0].with_color &&
                cfgs[i].device == cfgs[0].device && cfgs[i].im_h == cfgs[0].im_h &&
                cfgs[i].im_w == cfgs[0].im_w && cfgs[i].max_depth == cfgs[0].max_depth,
            "instances of a batch must share semantics, colour mode, device, frame shape, max_depth");
  }
  tsdf_batch* b = new (std::nothrow) tsdf_batch();
  if (!b) { tsdf_set_error("tsdf_batch_create: host allocation failed"); return TSDF_E_NOMEM; }
  b->n = n;
  for (int i = 0; i < n; ++i) {
    tsdf_vol* v = nullptr;
    int rc = tsdf_create(&cfgs[i], &v);
    if (rc) { tsdf_batch_destroy(b); return rc; }
    b->vols.push_back(v);   // owned by the batch from here on: tsdf_batch_destroy frees it
    if (v->dzl != v->cfg.dim_z) {
      tsdf_batch_destroy(b);
      tsdf_set_error("tsdf_batch_create: batched instances must be whole grids");
      return TSDF_E_INVALID;
    }
  }
  b->frame_owner = b->vols[0];
  DeviceGuard g(cfgs[0].device);
  const size_t px = frame_pixels(cfgs[0]);
  auto fail = [&](const char* what) { tsdf_set_error("tsdf_batch_create: %s failed", what); tsdf_batch_destroy(b); return TSDF_E_NOMEM; };
  if (cudaMalloc(&b->d_descs, sizeof(BatchDesc) * (size_t)n)) return fail("cudaMalloc descs");
  if (cudaHostAlloc(&b->h_descs, sizeof(BatchDesc) * (size_t)n, cudaHostAllocDefault)) return fail("cudaHostAlloc descs");
  if (cudaMalloc(&b->d_masks, px * (size_t)n)) return fail("cudaMalloc masks");
  if (cudaHostAlloc(&b->h_masks, px * (size_t)n, cudaHostAllocDefault)) return fail("cudaHostAlloc masks");
  if (cudaMalloc(&b->d_boxes, sizeof(MaskBox) * (size_t)n)) return fail("cudaMalloc boxes");
  if (cudaHostAlloc(&b->h_boxes, sizeof(MaskBox) * (size_t)n, cudaHostAllocDefault)) return fail("cudaHostAlloc boxes");
  if (cudaMalloc(&b->d_written, 8)) return fail("cudaMalloc counter");
  if (cudaHostAlloc(&b->h_written, 8, cudaHostAllocDefault)) return fail("cudaHostAlloc counter");
  *b->h_written = 0;
  cudaEventCreate(&b->ev0); cudaEventCreate(&b->ev1);
  cudaEventCreateWithFlags(&b->masks_free, cudaEventDisableTiming);
  *out = b;
  return TSDF_OK;
}

extern "C" int tsdf_batch_size(const tsdf_batch* b, int* n) {
  REQUIRE(b && n, "null argument");
  *n = b->n;
  return TSDF_OK;
}

static int batch_integrate(tsdf_batch* b, const float* depth, const uint8_t* rgb, const int32_t* active,
                           int n_active, const uint8_t* masks, const int32_t* boxes, const uint8_t* crops,
                           const float* K, const float* T_wc, float obs_weight) {
  REQUIRE(b && depth && T_wc && active && n_active >= 1 && n_active <= b->n, "bad argument");
  int rcw = check_obs_weight(obs_weight);
  if (rcw) return rcw;
  // every id in range and at most once: two blockIdx.z slices on one instance would race on its
  // voxels (and count its weights twice)
  b->seen.assign((size_t)b->n, 0);
  for (int i = 0; i < n_active; ++i) {
    REQUIRE(active[i] >= 0 && active[i] < b->n, "instance id out of range");
    REQUIRE(!b->seen[(size_t)active[i]], "duplicate instance id in the active list");
    b->seen[(size_t)active[i]] = 1;
  }
  tsdf_vol* fo = b->frame_owner;
  DeviceGuard g(fo->cfg.device);
  const size_t px = frame_pixels(fo->cfg);
  const bool color = fo->d_color != nullptr && rgb != nullptr;
  // the previous launch may still read descs/masks
  if (b->masks_in_flight) { CU_CHECK(cudaEventSynchronize(b->masks_free)); b->masks_in_flight = false; }
  // 1. shared frame: one upload + prep on the copy stream
  FrameSlot* slot = nullptr;
  int rc = stage_frame(fo, (int)(fo->seq % 3), depth, TSDF_DEPTH_F32, 1.0f, 1, 1, rgb, color, &slot);
  if (rc) return rc;
  ++fo->seq;
  // 2. masks: one block copy (device masks are used in place)
  const uint8_t* d_masks = nullptr;
  if (boxes) {  // cropped masks: upload sum(w*h) bytes, expand to full frames on the device
    const int H = fo->cfg.im_h, W = fo->cfg.im_w;
    long long total = 0;
    for (int i = 0; i < n_active; ++i) {
      const int32_t* q = boxes + 4 * i;
      REQUIRE(q[0] >= 0 && q[1] >= 0 && q[2] >= 0 && q[3] >= 0 && q[0] + q[2] <= W && q[1] + q[3] <= H,
              "mask box outside the image");
      b->h_boxes[i] = MaskBox{q[0], q[1], q[2], q[3], total};
      total += (long long)q[2] * q[3];
    }
    if (!b->d_crops) CU_CHECK(cudaMalloc(&b->d_crops, px * (size_t)b->n));
    if (total > 0) {
      REQUIRE(crops, "boxes given without mask crops");
      cudaPointerAttributes a;
      const bool known = cudaPointerGetAttributes(&a, crops) == cudaSuccess;
      cudaGetLastError();
      const void* src = crops;
      if (!(known && (a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged))) {
        host_copy(b->h_masks, crops, (size_t)total);  // pageable -> pinned staging
        src = b->h_masks;
      }
      CU_CHECK(cudaMemcpyAsync(b->d_crops, src, (size_t)total, cudaMemcpyDefault, fo->s_copy));
    }
    CU_CHECK(cudaMemcpyAsync(b->d_boxes, b->h_boxes, sizeof(MaskBox) * (size_t)n_active, cudaMemcpyHostToDevice, fo->s_copy));
    k_expand_masks<<<dim3(32, n_active), 256, 0, fo->s_copy>>>(b->d_crops, b->d_boxes, H, W, b->d_masks);
    CU_LAUNCHED();
    d_masks = b->d_masks;
  } else if (masks) {
    cudaPointerAttributes a;
    bool on_dev = cudaPointerGetAttributes(&a, masks) == cudaSuccess &&
                  (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    if (on_dev) d_masks = masks;
    else {
      const bool pinned = a.type == cudaMemoryTypeHost;
      const void* src = masks;
      if (!pinned) { host_copy(b->h_masks, masks, px * (size_t)n_active); src = b->h_masks; }
      CU_CHECK(cudaMemcpyAsync(b->d_masks, src, px * (size_t)n_active, cudaMemcpyHostToDevice, fo->s_copy));
      d_masks = b->d_masks;
    }
  }
  // 3. descriptors
  int max_fast = 0, max_rows = 0;
  int64_t visited = 0;
  bool vec4 = true;
  bool wmode = true;
  for (int i = 0; i < n_active; ++i) {
    tsdf_vol* v = b->vols[active[i]];
    ++v->epoch;
    const tsdf_config& c = v->cfg;
    BatchDesc& d = b->h_descs[i];
    (void)c;
    fill_grid_params(v, color, &d.g);
    d.g.written = b->d_written;
    make_frame_params(v, K, T_wc, obs_weight, slot->d_frame, &d.f);
    wmode = account_weights(v, &d.f, 1) && wmode;
    d.mask = d_masks ? d_masks + px * (size_t)i : nullptr;
    if (boxes) {  // outside its box the instance's mask is 0: segments that cannot project into it are culled
      const int32_t* q = boxes + 4 * i;
      d.f.mb_x0 = q[0]; d.f.mb_y0 = q[1]; d.f.mb_x1 = q[0] + q[2] - 1; d.f.mb_y1 = q[1] + q[3] - 1;
    }
    max_fast = d.g.n_fast > max_fast ? d.g.n_fast : max_fast;
    int rows = d.g.n_mid * d.g.n_slow;
    max_rows = rows > max_rows ? rows : max_rows;
    vec4 = vec4 && (d.g.n_fast % 4 == 0);
    visited += v->n_local;
    v->frames += 1;
  }
  CU_CHECK(cudaMemcpyAsync(b->d_descs, b->h_descs, sizeof(BatchDesc) * (size_t)n_active,
                           cudaMemcpyHostToDevice, fo->s_copy));
  CU_CHECK(cudaEventRecord(slot->uploaded, fo->s_copy));  // covers frame, masks and descriptors
  // 4. one launch for all active instances
  CU_CHECK(cudaStreamWaitEvent(fo->s_compute, slot->uploaded, 0));
  CU_CHECK(cudaMemsetAsync(b->d_written, 0, 8, fo->s_compute));
  CU_CHECK(cudaEventRecord(b->ev0, fo->s_compute));
  rc = launch_integrate_batch(b->d_descs, n_active, max_fast, max_rows, fo->zfast, color, vec4, (color && wmode) ? 1 : 0, fo->s_compute);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(b->ev1, fo->s_compute));
  CU_CHECK(cudaMemcpyAsync(b->h_written, b->d_written, 8, cudaMemcpyDeviceToHost, fo->s_compute));
  CU_CHECK(cudaEventRecord(slot->consumed, fo->s_compute));
  CU_CHECK(cudaEventRecord(b->masks_free, fo->s_compute));
  slot->in_flight = true;
  b->masks_in_flight = true;
  b->last_active = n_active;
  b->last_visited = visited;
  b->last_color = color;
  return TSDF_OK;
}

extern "C" int tsdf_batch_integrate_async(tsdf_batch* b, const float* depth, const uint8_t* rgb,
                                          const int32_t* active, int n_active, const uint8_t* masks,
                                          const float* K, const float* T_wc, float obs_weight) {
  return batch_integrate(b, depth, rgb, active, n_active, masks, nullptr, nullptr, K, T_wc, obs_weight);
}

extern "C" int tsdf_batch_integrate_boxes_async(tsdf_batch* b, const float* depth, const uint8_t* rgb,
                                                const int32_t* active, int n_active, const int32_t* boxes,
                                                const uint8_t* mask_crops, const float* K, const float* T_wc,
                                                float obs_weight) {
  REQUIRE(boxes, "null boxes");
  return batch_integrate(b, depth, rgb, active, n_active, nullptr, boxes, mask_crops, K, T_wc, obs_weight);
}

extern "C" int tsdf_batch_sync(tsdf_batch* b) {
  REQUIRE(b, "null handle");
  int rc = tsdf_sync(b->frame_owner);
  b->masks_in_flight = false;
  return rc;
}

extern "C" int tsdf_batch_get_stats(tsdf_batch* b, tsdf_stats* s) {
  REQUIRE(b && s, "null argument");
  int rc = tsdf_batch_sync(b);
  if (rc) return rc;
  memset(s, 0, sizeof(*s));
  s->voxels_visited = b->last_visited;
  s->voxels_written = (int64_t)*b->h_written;
  s->frames = b->frame_owner->seq;
  if (b->last_active > 0) cudaEventElapsedTime(&s->last_kernel_ms, b->ev0, b->ev1);
  s->bytes_algorithmic = (b->last_color ? 24 : 16) * s->voxels_written +
                         (int64_t)frame_pixels(b->frame_owner->cfg) * ((b->last_color ? 8 : 4) + b->last_active);
  return TSDF_OK;
}

extern "C" int tsdf_batch_download(tsdf_batch* b, int instance, float* tsdf, float* weight, float* color) {
  REQUIRE(b && instance >= 0 && instance < b->n, "bad argument");
  int rc = tsdf_batch_sync(b);
  if (rc) return rc;
  return tsdf_download(b->vols[instance], tsdf, weight, color, 0);
}

extern "C" int tsdf_batch_extract_points(tsdf_batch* b, int instance, int predicate, float tsdf_thresh,
                                         float weight_thresh, float* xyz, uint8_t* rgb, int64_t* idx,
                                         int64_t cap, int64_t* count) {
  REQUIRE(b && instance >= 0 && instance < b->n, "bad argument");
  int rc = tsdf_batch_sync(b);
  if (rc) return rc;
  return tsdf_extract_points(b->vols[instance], predicate, tsdf_thresh, weight_thresh, xyz, rgb, idx, cap, count);
}

extern "C" int tsdf_batch_extract_mesh(tsdf_batch* b, int instance, float* verts, float* norms, uint8_t* colors,
                                       int64_t vcap, int32_t* faces, int64_t fcap, int64_t* n_verts,
                                       int64_t* n_faces) {
  REQUIRE(b && instance >= 0 && instance < b->n, "bad argument");
  int rc = tsdf_batch_sync(b);
  if (rc) return rc;
  return tsdf_extract_mesh(b->vols[instance], verts, norms, colors, vcap, faces, fcap, n_verts, n_faces);
}
