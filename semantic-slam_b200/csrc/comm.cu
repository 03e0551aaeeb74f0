// comm.cu — multi-GPU plumbing for z-slab sharded volumes.
//
// The reference has no multi-GPU path at all; this is new design.  The voxel
// update is embarrassingly parallel across voxels, so the only exchange steps are
//   * per frame: the bulky image payload (depth fp32 + rgb u8) is broadcast from the rank
//     that received it to all ranks over NVLink/NVSwitch with ncclBroadcast (one grouped
//     launch), into each rank's raw frame slot, then every rank preps + integrates its slab;
//     the 100-byte pose/intrinsics metadata travels on the host control plane (the caller
//     passes it on every rank);
//   * at readback: counts are all-gathered, point records are sent to the root with grouped
//     ncclSend/ncclRecv and merged to ascending GLOBAL reference index.
// NCCL is bound at run time with dlopen so that the core library has no link-time
// dependency (a process that already loaded torch's libnccl.so.2 reuses that one).
#include <dlfcn.h>
#include <nccl.h>

#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace {

struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*) = nullptr;  // optional
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi g_nccl;

int load_nccl() {
  if (g_nccl.h) return TSDF_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    tsdf_set_error("NCCL not found: %s", dlerror());
    return TSDF_E_NCCL;
  }
#define BIND(field, sym)                                              \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, sym)); \
  if (!g_nccl.field) { tsdf_set_error("NCCL symbol %s missing", sym); return TSDF_E_NCCL; }
  BIND(GetUniqueId, "ncclGetUniqueId");
  BIND(CommInitRank, "ncclCommInitRank");
  g_nccl.CommInitRankConfig = reinterpret_cast<decltype(g_nccl.CommInitRankConfig)>(dlsym(h, "ncclCommInitRankConfig"));
  BIND(CommDestroy, "ncclCommDestroy");
  BIND(Broadcast, "ncclBroadcast");
  BIND(AllGather, "ncclAllGather");
  BIND(Send, "ncclSend");
  BIND(Recv, "ncclRecv");
  BIND(GroupStart, "ncclGroupStart");
  BIND(GroupEnd, "ncclGroupEnd");
  BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
  g_nccl.h = h;
  return TSDF_OK;
}

#define NCCL_CHECK(expr)                                                              \
  do {                                                                                \
    ncclResult_t r__ = (expr);                                                        \
    if (r__ != ncclSuccess) {                                                         \
      tsdf_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(r__)); \
      return TSDF_E_NCCL;                                                             \
    }                                                                                 \
  } while (0)

}  // namespace

#include "comm_dist.inc"

// ---- readback: gather to the root, merge on the device -------------------------------------------
extern "C" int tsdf_gather_points(tsdf_vol* v, tsdf_comm* c, int root, int predicate, float tsdf_thresh,
                                  float weight_thresh, float* xyz, uint8_t* rgb, int64_t* idx,
                                  int64_t cap, int64_t* count) {
  REQUIRE(v && c && count, "null argument");
  REQUIRE(root >= 0 && root < c->nranks, "bad root");
  DeviceGuard g(v->cfg.device);
  const bool is_root = c->rank == root;
  const bool want_rgb = v->d_color != nullptr;
  int rc = tsdf_sync(v);
  if (rc) return rc;
  // 1. local extraction into this rank's persistent device buffers (idx always: it is the merge key)
  int64_t n_local = 0;
  rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &n_local);
  if (rc) return rc;
  cudaStream_t st = v->s_compute;
  // 2. counts
  std::vector<int64_t> counts((size_t)c->nranks, 0);
  if (c->nranks > 1) {
    CU_CHECK(cudaMemcpyAsync(c->d_counts + c->rank, &n_local, 8, cudaMemcpyHostToDevice, st));
    NCCL_CHECK(g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclInt64, c->comm, st));
    CU_CHECK(cudaMemcpyAsync(counts.data(), c->d_counts, 8 * (size_t)c->nranks, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaStreamSynchronize(st));
  } else {
    counts[0] = n_local;
  }
  int64_t total = 0;
  std::vector<int64_t> offs((size_t)c->nranks + 1, 0);
  for (int r = 0; r < c->nranks; ++r) { offs[r] = total; total += counts[r]; }
  offs[c->nranks] = total;
  *count = is_root ? total : n_local;
  if (cap < 0) return TSDF_OK;  // collective count-only call (every rank passes cap < 0)
  // 3. records: every rank emits into its own buffers — the root straight into its place in the
  //    gathered arrays — and sends them to the root (every rank takes part even if the root's
  //    caller buffers turn out too small: the root reports TSDF_E_CAPACITY afterwards)
  OutBuffers& loc = is_root ? c->gathered : c->out;
  const int64_t n_mine = is_root ? total : n_local;
  const size_t nn = (size_t)std::max<int64_t>(n_mine, 1);
  if ((rc = loc.need_dev(0, nn * 12)) || (rc = loc.need_dev(2, nn * 8)) || (want_rgb && (rc = loc.need_dev(1, nn * 3)))) return rc;
  float* g_xyz = static_cast<float*>(loc.d[0]);
  uint8_t* g_rgb = want_rgb ? static_cast<uint8_t*>(loc.d[1]) : nullptr;
  int64_t* g_idx = static_cast<int64_t*>(loc.d[2]);
  const int64_t my_off = is_root ? offs[root] : 0;
  if (n_local) {
    rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, g_xyz + 3 * my_off, g_rgb ? g_rgb + 3 * my_off : nullptr,
                        g_idx + my_off, n_local, &n_local);
    if (rc) return rc;
  }
  if (c->nranks > 1) {
    NCCL_CHECK(g_nccl.GroupStart());
    if (!is_root) {
      if (n_local) {
        NCCL_CHECK(g_nccl.Send(g_xyz, (size_t)n_local * 3, ncclFloat32, root, c->comm, st));
        NCCL_CHECK(g_nccl.Send(g_idx, (size_t)n_local, ncclInt64, root, c->comm, st));
        if (want_rgb) NCCL_CHECK(g_nccl.Send(g_rgb, (size_t)n_local * 3, ncclUint8, root, c->comm, st));
      }
    } else {
      for (int r = 0; r < c->nranks; ++r) {
        if (r == root || counts[r] == 0) continue;
        NCCL_CHECK(g_nccl.Recv(g_xyz + offs[r] * 3, (size_t)counts[r] * 3, ncclFloat32, r, c->comm, st));
        NCCL_CHECK(g_nccl.Recv(g_idx + offs[r], (size_t)counts[r], ncclInt64, r, c->comm, st));
        if (want_rgb) NCCL_CHECK(g_nccl.Recv(g_rgb + offs[r] * 3, (size_t)counts[r] * 3, ncclUint8, r, c->comm, st));
      }
    }
    NCCL_CHECK(g_nccl.GroupEnd());
  }
  if (!is_root || !(xyz || rgb || idx)) {
    CU_CHECK(cudaStreamSynchronize(st));
    return TSDF_OK;
  }
  if (cap < total) {
    CU_CHECK(cudaStreamSynchronize(st));
    tsdf_set_error("tsdf_gather_points: capacity %lld < %lld points", (long long)cap, (long long)total);
    return TSDF_E_CAPACITY;
  }
  if (total == 0) { CU_CHECK(cudaStreamSynchronize(st)); return TSDF_OK; }
  // 4. merge to ascending GLOBAL reference index on the device.  x-fast slabs are contiguous index
  //    ranges (the concatenation is already sorted); z-fast slabs interleave per (x, y) row.
  const float* o_xyz = g_xyz; const uint8_t* o_rgb = g_rgb; const int64_t* o_idx = g_idx;
  if (v->zfast && c->nranks > 1) {
    OutBuffers& m = c->out;   // the root's `out` is otherwise unused
    if ((rc = m.need_dev(0, (size_t)total * 12)) || (rc = m.need_dev(2, (size_t)total * 8)) ||
        (want_rgb && (rc = m.need_dev(1, (size_t)total * 3))) || (rc = m.need_dev(3, 8 * ((size_t)c->nranks + 1)))) return rc;
    CU_CHECK(cudaMemcpyAsync(m.d[3], offs.data(), 8 * ((size_t)c->nranks + 1), cudaMemcpyHostToDevice, st));
    if ((rc = launch_merge_scatter(total, c->nranks, static_cast<const int64_t*>(m.d[3]), g_xyz, g_idx, g_rgb,
                                   static_cast<float*>(m.d[0]), static_cast<int64_t*>(m.d[2]),
                                   want_rgb ? static_cast<uint8_t*>(m.d[1]) : nullptr, st))) return rc;
    o_xyz = static_cast<const float*>(m.d[0]);
    o_rgb = want_rgb ? static_cast<const uint8_t*>(m.d[1]) : nullptr;
    o_idx = static_cast<const int64_t*>(m.d[2]);
    CU_CHECK(cudaStreamSynchronize(st));   // `offs` (pageable) has been consumed by the H2D copy
  }
  // 5. one asynchronous D2H per array into the caller's memory
  if (xyz && (rc = copy_to_host(&c->gathered, xyz, o_xyz, (size_t)total * 12, st))) return rc;
  if (idx && (rc = copy_to_host(&c->gathered, idx, o_idx, (size_t)total * 8, st))) return rc;
  if (rgb) {
    if (o_rgb) { if ((rc = copy_to_host(&c->gathered, rgb, o_rgb, (size_t)total * 3, st))) return rc; }
    else memset(rgb, 0, (size_t)total * 3);
  }
  CU_CHECK(cudaStreamSynchronize(st));
  return TSDF_OK;
}

// ---- mesh of a sharded volume (plugin SaveMesh on the multi-GPU configuration) ----------------------
// Cells across a slab face need both sides: every rank sends its tsdf (and colour) slab to the root,
// which assembles a temporary WHOLE-grid volume on its own device and meshes it with the whole-grid
// kernels — vertex / face order and every bit are those of a single-GPU volume.  Collective; the
// ranks all-gather their slab epochs first, and an assembly whose epochs are unchanged is reused
// (the sizing call and the call that follows it move the volume once).
extern "C" int tsdf_gather_mesh(tsdf_vol* v, tsdf_comm* c, int root, float* verts, float* norms, uint8_t* colors,
                                int64_t vcap, int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces) {
  REQUIRE(v && c && n_verts && n_faces, "null argument");
  REQUIRE(root >= 0 && root < c->nranks, "bad root");
  DeviceGuard g(v->cfg.device);
  const bool is_root = c->rank == root;
  int rc = tsdf_sync(v);
  if (rc) return rc;
  cudaStream_t st = v->s_compute;
  *n_verts = *n_faces = 0;
  std::vector<int64_t> ep((size_t)c->nranks, 0);
  const int64_t mine = (int64_t)v->epoch;
  if (c->nranks > 1) {
    CU_CHECK(cudaMemcpyAsync(c->d_counts + c->rank, &mine, 8, cudaMemcpyHostToDevice, st));
    NCCL_CHECK(g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclInt64, c->comm, st));
    CU_CHECK(cudaMemcpyAsync(ep.data(), c->d_counts, 8 * (size_t)c->nranks, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaStreamSynchronize(st));
  } else {
    ep[0] = mine;
  }
  const bool fresh = ep != c->whole_epochs;   // the same decision on every rank
  if (fresh) {
    const tsdf_config& vc = v->cfg;
    const size_t dx = (size_t)vc.dim_x, dy = (size_t)vc.dim_y, dz = (size_t)vc.dim_z;
    tsdf_vol* w = nullptr;
    if (is_root) {
      w = c->whole;
      c->whole = nullptr;
      if (!w) {
        tsdf_config wc = vc;
        wc.z_begin = 0; wc.z_end = vc.dim_z;
        if ((rc = tsdf_create(&wc, &w))) return rc;   // (a failure here leaves the peers in their sends: fatal for the job)
      }
      ++w->epoch;
    }
    for (int r = 0; r < c->nranks; ++r) {
      int z0, z1;
      tsdf_slab_range(vc.dim_z, c->nranks, r, &z0, &z1);
      const size_t nvox = dx * dy * (size_t)(z1 - z0);
      for (int a = 0; a < 2; ++a) {   // tsdf, colour
        float* mine_arr = a == 0 ? v->d_tsdf : v->d_color;
        if (a == 1 && !v->d_color) continue;
        if (r != root) {
          if (c->rank == r) NCCL_CHECK(g_nccl.Send(mine_arr, nvox, ncclFloat32, root, c->comm, st));
          if (!is_root) continue;
        }
        if (!is_root) continue;
        float* warr = a == 0 ? w->d_tsdf : w->d_color;
        const float* src = mine_arr;
        if (!v->zfast) {   // [z][y][x]: a contiguous range of the whole array
          float* dst = warr + (size_t)z0 * dy * dx;
          if (r != root) NCCL_CHECK(g_nccl.Recv(dst, nvox, ncclFloat32, r, c->comm, st));
          else CU_CHECK(cudaMemcpyAsync(dst, src, nvox * 4, cudaMemcpyDeviceToDevice, st));
        } else {           // [x][y][z]: a z window of every row
          if (r != root) {
            if ((rc = c->gathered.need_dev(3, nvox * 4))) return rc;
            NCCL_CHECK(g_nccl.Recv(c->gathered.d[3], nvox, ncclFloat32, r, c->comm, st));
            src = static_cast<const float*>(c->gathered.d[3]);
          }
          CU_CHECK(cudaMemcpy2DAsync(warr + z0, dz * 4, src, (size_t)(z1 - z0) * 4, (size_t)(z1 - z0) * 4, dx * dy,
                                     cudaMemcpyDeviceToDevice, st));
        }
      }
    }
    CU_CHECK(cudaStreamSynchronize(st));
    c->whole = w;
    c->whole_epochs = ep;
  }
  if (!is_root) return TSDF_OK;
  return tsdf_extract_mesh(c->whole, verts, norms, colors, vcap, faces, fcap, n_verts, n_faces);
}
