// group.cu — ONE process, several GPUs: a volume cut into z-slabs over a device list, driven by a
// single host thread through one handle (SURVEY §8b "n_gpus/device list", §5 "one process is enough").
//
// The reference is one process (Engine::Run, src/Engine.cpp:105; Object::Integrate,
// src/Object.cpp:143-166), so "drops into Engine.cpp ... z-slab sharded on 8 B200" needs a handle a
// C++ host can hold without torchrun or a second process.  The data path is the same fused
// peer-pull as the one-process-per-GPU path (comm_dist.inc) but inside one address space it needs
// neither cudaIpc nor flag polling: the root stages the frame once, every other slab's prep kernel
// reads the root's raw slot directly over NVLink (cudaDeviceEnablePeerAccess), and all ordering is
// CUDA events across the devices' copy streams — no spinning kernel anywhere.  Without peer access
// between two devices the raw frame is moved with cudaMemcpyPeerAsync first.  No NCCL.
// A device may appear more than once in the list (several slabs on one GPU): that is how the group
// logic, the merge and the whole-grid mesh are tested on a 1-GPU box.
#include <algorithm>
#include <new>

#include "common.cuh"

struct tsdf_group {
  tsdf_config cfg;                       // whole grid
  int n = 0;
  std::vector<tsdf_vol*> slab;           // slab[0] is the root (stages the frames)
  std::vector<int> dev;
  std::vector<char> direct;              // slab i reads the root's raw slot in place
  uint64_t seq = 0;
  cudaEvent_t raw_ready[kBcastSlots] = {};            // root copy stream: frame is in its raw slot
  std::vector<cudaEvent_t> prep_done[kBcastSlots];    // slab i copy stream: it has consumed the raw slot
  OutBuffers gathered, merged;           // root device: readback
  tsdf_vol* whole = nullptr;             // assembled whole-grid copy for the mesh (kept while the slabs are unchanged)
  std::vector<uint64_t> whole_epochs;
  std::vector<int64_t> counts, offs;
};

extern "C" int tsdf_group_destroy(tsdf_group* g) {
  if (!g) return TSDF_OK;
  for (int i = 0; i < (int)g->slab.size(); ++i) {
    if (!g->slab[i]) continue;
    DeviceGuard dg(g->dev[i]);
    tsdf_sync(g->slab[i]);
    for (int k = 0; k < kBcastSlots; ++k)
      if (i < (int)g->prep_done[k].size() && g->prep_done[k][i]) cudaEventDestroy(g->prep_done[k][i]);
  }
  if (!g->dev.empty()) {
    DeviceGuard dg(g->dev[0]);
    for (cudaEvent_t e : g->raw_ready) if (e) cudaEventDestroy(e);
    g->gathered.release();
    g->merged.release();
  }
  if (g->whole) tsdf_destroy(g->whole);
  for (tsdf_vol* v : g->slab) if (v) tsdf_destroy(v);
  delete g;
  return TSDF_OK;
}

extern "C" int tsdf_group_create(const tsdf_config* cfg, const int32_t* devices, int n_devices, tsdf_group** out) {
  REQUIRE(cfg && devices && out && n_devices >= 1 && n_devices <= 64, "bad argument");
  *out = nullptr;
  REQUIRE(cfg->z_begin == 0 && (cfg->z_end == 0 || cfg->z_end == cfg->dim_z), "a group takes the WHOLE grid's config");
  REQUIRE((cfg->dim_z + 3) / 4 >= n_devices, "fewer 4-slice groups than devices");
  int ndev = 0;
  int rc = tsdf_device_count(&ndev);
  if (rc) return rc;
  for (int i = 0; i < n_devices; ++i) REQUIRE(devices[i] >= 0 && devices[i] < ndev, "device ordinal out of range");
  tsdf_group* g = new (std::nothrow) tsdf_group();
  if (!g) { tsdf_set_error("tsdf_group_create: host allocation failed"); return TSDF_E_NOMEM; }
  g->cfg = *cfg;
  g->cfg.z_begin = 0;
  g->cfg.z_end = cfg->dim_z;
  g->n = n_devices;
  g->slab.assign((size_t)n_devices, nullptr);
  g->dev.assign(devices, devices + n_devices);
  g->direct.assign((size_t)n_devices, 1);
  g->counts.assign((size_t)n_devices, 0);
  g->offs.assign((size_t)n_devices + 1, 0);
  for (int k = 0; k < kBcastSlots; ++k) g->prep_done[k].assign((size_t)n_devices, nullptr);
  auto fail = [&](int code) { tsdf_group_destroy(g); return code; };
  for (int i = 0; i < n_devices; ++i) {
    tsdf_config c = g->cfg;
    c.device = devices[i];
    tsdf_slab_range(c.dim_z, n_devices, i, &c.z_begin, &c.z_end);
    if ((rc = tsdf_create(&c, &g->slab[i]))) return fail(rc);
    DeviceGuard dg(devices[i]);
    for (int k = 0; k < kBcastSlots; ++k)
      if (cudaEventCreateWithFlags(&g->prep_done[k][i], cudaEventDisableTiming) != cudaSuccess) {
        tsdf_set_error("tsdf_group_create: event creation failed");
        return fail(TSDF_E_CUDA);
      }
    if (i > 0 && devices[i] != devices[0]) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[i], devices[0]) != cudaSuccess) can = 0;
      if (can) {
        cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
      }
      cudaGetLastError();
      if (getenv("TSDF_B200_GROUP_NO_PEER")) can = 0;   // test hook: force the staged-copy path
      g->direct[i] = can ? 1 : 0;
    }
  }
  {
    DeviceGuard dg(devices[0]);
    for (int k = 0; k < kBcastSlots; ++k)
      if (cudaEventCreateWithFlags(&g->raw_ready[k], cudaEventDisableTiming) != cudaSuccess) {
        tsdf_set_error("tsdf_group_create: event creation failed");
        return fail(TSDF_E_CUDA);
      }
  }
  *out = g;
  return TSDF_OK;
}

extern "C" int tsdf_group_size(const tsdf_group* g, int* n) {
  REQUIRE(g && n, "null argument");
  *n = g->n;
  return TSDF_OK;
}

extern "C" int tsdf_group_slab(tsdf_group* g, int i, tsdf_vol** slab) {
  REQUIRE(g && slab && i >= 0 && i < g->n, "bad argument");
  *slab = g->slab[i];
  return TSDF_OK;
}

extern "C" int tsdf_group_peer_access(const tsdf_group* g, int* n_direct) {
  REQUIRE(g && n_direct, "null argument");
  int k = 0;
  for (int i = 1; i < g->n; ++i) k += g->direct[i] ? 1 : 0;
  *n_direct = k;
  return TSDF_OK;
}

extern "C" int tsdf_group_integrate_async(tsdf_group* g, const float* depth, const uint8_t* rgb, const float* K,
                                          const float* T_wc, float obs_weight) {
  REQUIRE(g && depth && T_wc, "null argument");
  int rc = check_obs_weight(obs_weight);
  if (rc) return rc;
  tsdf_vol* root = g->slab[0];
  const bool color = root->d_color != nullptr && rgb != nullptr;
  const uint64_t seq = ++g->seq;
  const int k = (int)((seq - 1) % kBcastSlots);
  const size_t px = frame_pixels(g->cfg);
  const size_t rgb_off = ((px * 4 + 63) / 64) * 64;
  const size_t raw_bytes = px * 7 + 64;
  // every slab's slot k: the kernel that last read its prepared frame must be done (bounds the
  // host's run-ahead to the ring depth)
  for (int i = 0; i < g->n; ++i) {
    tsdf_vol* v = g->slab[i];
    DeviceGuard dg(g->dev[i]);
    if ((rc = ensure_bcast_slot(v, k))) return rc;
    FrameSlot& s = v->bslot[k];
    if (s.in_flight) { CU_CHECK(cudaEventSynchronize(s.consumed)); s.in_flight = false; }
  }
  {  // ---- root: stage the frame once -----------------------------------------------------------
    DeviceGuard dg(g->dev[0]);
    FrameSlot& s = root->bslot[k];
    if (seq > (uint64_t)kBcastSlots)   // the raw slot is still being read by a slab's prep of frame seq - ring?
      for (int i = 1; i < g->n; ++i) CU_CHECK(cudaStreamWaitEvent(root->s_copy, g->prep_done[k][i], 0));
    trace_mark(root, 0, root->s_copy);
    const void* dummy = nullptr;
    if ((rc = frame_to_device(root, s, depth, px * 4, 0, &dummy, true))) return rc;
    if (color && (rc = frame_to_device(root, s, rgb, px * 3, rgb_off, &dummy, true))) return rc;
    CU_CHECK(cudaEventRecord(g->raw_ready[k], root->s_copy));
  }
  const char* root_raw = reinterpret_cast<const char*>(root->bslot[k].d_raw);
  for (int i = 0; i < g->n; ++i) {  // ---- every slab: prep (reading the root's slot) + integrate ----
    tsdf_vol* v = g->slab[i];
    DeviceGuard dg(g->dev[i]);
    FrameSlot* s = &v->bslot[k];
    const char* src = root_raw;
    if (i > 0) trace_mark(v, 0, v->s_copy);
    if (i > 0) {
      CU_CHECK(cudaStreamWaitEvent(v->s_copy, g->raw_ready[k], 0));
      if (!g->direct[i]) {
        CU_CHECK(cudaMemcpyPeerAsync(s->d_raw, g->dev[i], root_raw, g->dev[0], raw_bytes, v->s_copy));
        src = reinterpret_cast<const char*>(s->d_raw);
      }
    }
    rc = launch_frame_prep(src, TSDF_DEPTH_F32, 1.0f, 1, 1, color ? reinterpret_cast<const uint8_t*>(src) + rgb_off : nullptr,
                           nullptr, v->cfg.im_h, v->cfg.im_w, v->cfg.max_depth, color, nullptr, s->d_frame, v->s_copy,
                           /*clear_tail=*/v->dev_hold, nullptr);
    if (rc) return rc;
    if (i > 0) CU_CHECK(cudaEventRecord(g->prep_done[k][i], v->s_copy));
    CU_CHECK(cudaEventRecord(s->uploaded, v->s_copy));
    trace_mark(v, 1, v->s_copy);
    FrameP fp;
    make_frame_params(v, K, T_wc, obs_weight, s->d_frame, &fp);
    if ((rc = submit_frames(v, &s, &fp, 1, color))) return rc;
  }
  return TSDF_OK;
}

extern "C" int tsdf_group_sync(tsdf_group* g) {
  REQUIRE(g, "null handle");
  for (tsdf_vol* v : g->slab) {
    int rc = tsdf_sync(v);
    if (rc) return rc;
  }
  return TSDF_OK;
}

extern "C" int tsdf_group_integrate(tsdf_group* g, const float* depth, const uint8_t* rgb, const float* K,
                                    const float* T_wc, float obs_weight) {
  int rc = tsdf_group_integrate_async(g, depth, rgb, K, T_wc, obs_weight);
  return rc ? rc : tsdf_group_sync(g);
}

extern "C" int tsdf_group_reset(tsdf_group* g) {
  REQUIRE(g, "null handle");
  for (tsdf_vol* v : g->slab) {
    int rc = tsdf_sync(v);
    if (!rc) rc = tsdf_reset(v);
    if (rc) return rc;
  }
  return TSDF_OK;
}

extern "C" int tsdf_group_get_stats(tsdf_group* g, tsdf_stats* s) {
  REQUIRE(g && s, "null argument");
  memset(s, 0, sizeof(*s));
  for (tsdf_vol* v : g->slab) {
    tsdf_stats t;
    int rc = tsdf_get_stats(v, &t);
    if (rc) return rc;
    s->voxels_visited += t.voxels_visited;
    s->voxels_written += t.voxels_written;
    s->bytes_algorithmic += t.bytes_algorithmic;
    s->last_kernel_ms = std::max(s->last_kernel_ms, t.last_kernel_ms);
    s->last_upload_ms = std::max(s->last_upload_ms, t.last_upload_ms);
    s->frames = t.frames;
  }
  return TSDF_OK;
}

extern "C" int tsdf_group_download(tsdf_group* g, float* tsdf, float* weight, float* color) {
  REQUIRE(g, "null handle");
  for (tsdf_vol* v : g->slab) {   // disjoint ranges of the caller's GLOBAL arrays
    int rc = tsdf_download(v, tsdf, weight, color, 1);
    if (rc) return rc;
  }
  return TSDF_OK;
}

// Every slab extracts on its own device; the runs are collected on the root device (directly when the
// slab lives there, else with a peer copy), merged to ascending GLOBAL reference index by
// k_merge_scatter and copied to the caller with one asynchronous D2H per array.
extern "C" int tsdf_group_extract_points(tsdf_group* g, int predicate, float tsdf_thresh, float weight_thresh,
                                         float* xyz, uint8_t* rgb, int64_t* idx, int64_t cap, int64_t* count) {
  REQUIRE(g && count, "null argument");
  REQUIRE(predicate == TSDF_PRED_NONZERO || predicate == TSDF_PRED_ABS_LT, "bad predicate");
  int rc;
  int64_t total = 0;
  for (int i = 0; i < g->n; ++i) {
    DeviceGuard dg(g->dev[i]);
    if ((rc = tsdf_sync(g->slab[i]))) return rc;
    if ((rc = extract_device(g->slab[i], predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &g->counts[i]))) return rc;
    g->offs[i] = total;
    total += g->counts[i];
  }
  g->offs[g->n] = total;
  *count = total;
  if (!xyz && !rgb && !idx) return TSDF_OK;
  if (cap < total) {
    tsdf_set_error("tsdf_group_extract_points: capacity %lld < %lld points", (long long)cap, (long long)total);
    return TSDF_E_CAPACITY;
  }
  if (total == 0) return TSDF_OK;
  tsdf_vol* root = g->slab[0];
  const bool want_rgb = root->d_color != nullptr;
  float* g_xyz; uint8_t* g_rgb; int64_t* g_idx;
  {
    DeviceGuard dg(g->dev[0]);
    OutBuffers& G = g->gathered;
    if ((rc = G.need_dev(0, (size_t)total * 12)) || (rc = G.need_dev(2, (size_t)total * 8)) ||
        (want_rgb && (rc = G.need_dev(1, (size_t)total * 3)))) return rc;
    g_xyz = static_cast<float*>(G.d[0]);
    g_rgb = want_rgb ? static_cast<uint8_t*>(G.d[1]) : nullptr;
    g_idx = static_cast<int64_t*>(G.d[2]);
  }
  for (int i = 0; i < g->n; ++i) {
    const int64_t n = g->counts[i], o = g->offs[i];
    if (!n) continue;
    tsdf_vol* v = g->slab[i];
    DeviceGuard dg(g->dev[i]);
    if (g->dev[i] == g->dev[0]) {
      if ((rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, g_xyz + 3 * o, g_rgb ? g_rgb + 3 * o : nullptr,
                               g_idx + o, n, &g->counts[i]))) return rc;
    } else {
      OutBuffers& L = v->out;
      if ((rc = L.need_dev(0, (size_t)n * 12)) || (rc = L.need_dev(2, (size_t)n * 8)) ||
          (want_rgb && (rc = L.need_dev(1, (size_t)n * 3)))) return rc;
      if ((rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, static_cast<float*>(L.d[0]),
                               want_rgb ? static_cast<uint8_t*>(L.d[1]) : nullptr, static_cast<int64_t*>(L.d[2]), n,
                               &g->counts[i]))) return rc;
      CU_CHECK(cudaMemcpyPeerAsync(g_xyz + 3 * o, g->dev[0], L.d[0], g->dev[i], (size_t)n * 12, v->s_compute));
      CU_CHECK(cudaMemcpyPeerAsync(g_idx + o, g->dev[0], L.d[2], g->dev[i], (size_t)n * 8, v->s_compute));
      if (want_rgb) CU_CHECK(cudaMemcpyPeerAsync(g_rgb + 3 * o, g->dev[0], L.d[1], g->dev[i], (size_t)n * 3, v->s_compute));
    }
    CU_CHECK(cudaStreamSynchronize(v->s_compute));
  }
  DeviceGuard dg(g->dev[0]);
  cudaStream_t st = root->s_compute;
  const float* o_xyz = g_xyz; const uint8_t* o_rgb = g_rgb; const int64_t* o_idx = g_idx;
  if (root->zfast && g->n > 1) {   // z-fast slabs interleave per (x, y) row; x-fast slabs concatenate
    OutBuffers& M = g->merged;
    if ((rc = M.need_dev(0, (size_t)total * 12)) || (rc = M.need_dev(2, (size_t)total * 8)) ||
        (want_rgb && (rc = M.need_dev(1, (size_t)total * 3))) || (rc = M.need_dev(3, 8 * ((size_t)g->n + 1)))) return rc;
    CU_CHECK(cudaMemcpyAsync(M.d[3], g->offs.data(), 8 * ((size_t)g->n + 1), cudaMemcpyHostToDevice, st));
    if ((rc = launch_merge_scatter(total, g->n, static_cast<const int64_t*>(M.d[3]), g_xyz, g_idx, g_rgb,
                                   static_cast<float*>(M.d[0]), static_cast<int64_t*>(M.d[2]),
                                   want_rgb ? static_cast<uint8_t*>(M.d[1]) : nullptr, st))) return rc;
    o_xyz = static_cast<const float*>(M.d[0]);
    o_rgb = want_rgb ? static_cast<const uint8_t*>(M.d[1]) : nullptr;
    o_idx = static_cast<const int64_t*>(M.d[2]);
  }
  if (xyz && (rc = copy_to_host(&g->gathered, xyz, o_xyz, (size_t)total * 12, st))) return rc;
  if (idx && (rc = copy_to_host(&g->gathered, idx, o_idx, (size_t)total * 8, st))) return rc;
  if (rgb) {
    if (o_rgb) { if ((rc = copy_to_host(&g->gathered, rgb, o_rgb, (size_t)total * 3, st))) return rc; }
    else memset(rgb, 0, (size_t)total * 3);
  }
  CU_CHECK(cudaStreamSynchronize(st));
  return TSDF_OK;
}

// Mesh of a sharded volume: cells across a slab face need both sides, so the slabs' tsdf (and colour)
// arrays are assembled into a temporary WHOLE-grid volume on the root device (peer copies; 8 B per
// voxel of the whole grid + the weight array tsdf_create adds: 12.9 GB at 1024^3, 7 % of a B200) and
// meshed there by the whole-grid kernels: vertex / face order and every bit are those of a
// single-GPU volume.  SaveMesh is a one-off at the end of a run; the copies run at NVLink rate.
// The assembled copy is kept while no slab changes (the sizing call and the call that follows it
// share it); tsdf_group_release_scratch() or the next change frees / refreshes it.
static int assemble_whole(tsdf_group* g, tsdf_vol** out) {
  std::vector<uint64_t> ep;
  for (tsdf_vol* v : g->slab) ep.push_back(v->epoch);
  if (g->whole && ep == g->whole_epochs) { *out = g->whole; return TSDF_OK; }
  tsdf_config c = g->cfg;
  c.device = g->dev[0];
  tsdf_vol* w = g->whole;
  int rc = w ? TSDF_OK : tsdf_create(&c, &w);
  if (rc) return rc;
  g->whole = nullptr;
  const size_t dx = (size_t)c.dim_x, dy = (size_t)c.dim_y, dz = (size_t)c.dim_z;
  for (int i = 0; i < g->n && !rc; ++i) {
    tsdf_vol* v = g->slab[i];
    if ((rc = tsdf_sync(v))) break;
    DeviceGuard dg(g->dev[0]);
    float* dst[2] = {w->d_tsdf, w->d_color};
    float* src[2] = {v->d_tsdf, v->d_color};
    for (int a = 0; a < 2 && !rc; ++a) {
      if (!dst[a] || !src[a]) continue;
      cudaMemcpy3DPeerParms p;
      memset(&p, 0, sizeof p);
      p.srcDevice = g->dev[i];
      p.dstDevice = g->dev[0];
      if (w->zfast) {   // [x][y][z]: rows of Dz floats, the slab fills z in [z_begin, z_end) of every row
        p.srcPtr = make_cudaPitchedPtr(src[a], (size_t)v->dzl * 4, (size_t)v->dzl * 4, dx * dy);
        p.dstPtr = make_cudaPitchedPtr(dst[a] + v->cfg.z_begin, dz * 4, dz * 4, dx * dy);
        p.extent = make_cudaExtent((size_t)v->dzl * 4, dx * dy, 1);
      } else {          // [z][y][x]: the slab is one contiguous range
        const size_t row = dx * 4, rows = dy * (size_t)v->dzl;
        p.srcPtr = make_cudaPitchedPtr(src[a], row, row, rows);
        p.dstPtr = make_cudaPitchedPtr(dst[a] + (size_t)v->cfg.z_begin * dy * dx, row, row, rows);
        p.extent = make_cudaExtent(row, rows, 1);
      }
      cudaError_t e = cudaMemcpy3DPeerAsync(&p, w->s_compute);
      if (e != cudaSuccess) { tsdf_set_error("group mesh: slab copy: %s", cudaGetErrorString(e)); rc = TSDF_E_CUDA; }
    }
  }
  if (!rc) {
    DeviceGuard dg(g->dev[0]);
    if (cudaStreamSynchronize(w->s_compute) != cudaSuccess) { tsdf_set_error("group mesh: copy failed"); rc = TSDF_E_CUDA; }
    ++w->epoch;
  }
  if (rc) { tsdf_destroy(w); return rc; }
  g->whole = w;
  g->whole_epochs = ep;
  *out = w;
  return TSDF_OK;
}

extern "C" int tsdf_group_release_scratch(tsdf_group* g) {
  REQUIRE(g, "null handle");
  if (g->whole) tsdf_destroy(g->whole);
  g->whole = nullptr;
  g->whole_epochs.clear();
  return TSDF_OK;
}

extern "C" int tsdf_group_extract_mesh(tsdf_group* g, float* verts, float* norms, uint8_t* colors, int64_t vcap,
                                       int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces) {
  REQUIRE(g && n_verts && n_faces, "null argument");
  tsdf_vol* w = nullptr;
  int rc = assemble_whole(g, &w);
  if (rc) return rc;
  return tsdf_extract_mesh(w, verts, norms, colors, vcap, faces, fcap, n_verts, n_faces);
}

extern "C" int tsdf_group_save_mesh_ply(tsdf_group* g, const char* path) {
  REQUIRE(g && path, "null argument");
  tsdf_vol* w = nullptr;
  int rc = assemble_whole(g, &w);
  if (rc) return rc;
  return tsdf_save_mesh_ply(w, path);
}

extern "C" int tsdf_group_save_ply(tsdf_group* g, const char* path, int predicate, float tsdf_thresh,
                                   float weight_thresh) {
  REQUIRE(g && path, "null argument");
  int64_t n = 0;
  int rc = tsdf_group_extract_points(g, predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &n);
  if (rc) return rc;
  std::vector<float> xyz((size_t)(n ? n : 1) * 3);
  if (n && (rc = tsdf_group_extract_points(g, predicate, tsdf_thresh, weight_thresh, xyz.data(), nullptr, nullptr, n, &n)))
    return rc;
  return write_points_ply(path, xyz.data(), n);
}

extern "C" int tsdf_group_save_bin(tsdf_group* g, const char* path) {
  REQUIRE(g && path, "null argument");
  const size_t n = (size_t)g->cfg.dim_x * g->cfg.dim_y * g->cfg.dim_z;
  std::vector<float> t(n);
  int rc = tsdf_group_download(g, t.data(), nullptr, nullptr);
  if (rc) return rc;
  return write_bin(path, g->cfg, t.data(), n);
}
