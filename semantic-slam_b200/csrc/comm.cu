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

// ---- peer-pull frame distribution (fused broadcast + prep) ------------------------------------
// Instead of an NCCL broadcast followed by the prep kernel, every non-root rank's prep kernel
// reads the root's raw frame slot directly through NVLink (cudaIpc-mapped peer memory) and
// writes its own prepared frame: one kernel does the transfer and the conversion, and no NCCL
// CTAs spin on the SMs the integrate kernel is using.  Ordering is by two kinds of sequence
// flags in the ROOT's memory: ready (root -> peers: "frame s is in its slot") and done[r]
// (peer r -> root: "I have consumed frame s", so the slot may be overwritten three frames
// later).  The waits are single-thread kernels on the copy streams, bounded by a timeout that
// raises an error flag instead of hanging.
namespace {
constexpr int kFlagWords = 64;          // [0] ready, [8 + r] done of rank r
// a root that is slow to produce a frame (file IO, a stalled host) must not trip this; a rank that
// died must not hang the others for ever.  TSDF_B200_PEER_TIMEOUT_MS overrides.
unsigned long long g_wait_timeout_ns = 20000000000ull;

__device__ __forceinline__ unsigned ld_flag(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// waits until *flag >= target (sequence numbers; target <= 0 passes at once)
__global__ void k_wait_flag(const unsigned* flag, int target, int* err, unsigned long long timeout_ns) {
  if (target <= 0) return;
  const unsigned long long t0 = now_ns();
  while ((int)ld_flag(flag) < target) {
    __nanosleep(200);
    if (now_ns() - t0 > timeout_ns) { *err = 1; __threadfence_system(); return; }
  }
}
// waits until every done[r] >= target, r != root
__global__ void k_wait_all(const unsigned* done, int nranks, int root, int target, int* err, unsigned long long timeout_ns) {
  const int r = threadIdx.x;
  if (target <= 0 || r >= nranks || r == root) return;
  const unsigned long long t0 = now_ns();
  while ((int)ld_flag(done + r) < target) {
    __nanosleep(200);
    if (now_ns() - t0 > timeout_ns) { *err = 2; __threadfence_system(); return; }
  }
}
__global__ void k_set_flag(unsigned* flag, unsigned value) {
  __threadfence_system();   // everything this stream did before (frame bytes / prep reads) is complete
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(flag), "r"(value));
  __threadfence_system();
}
}  // namespace

struct tsdf_comm {
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0, device = 0;
  int64_t* d_counts = nullptr;  // nranks entries
  // peer-pull state
  bool p2p_tried = false, p2p_ok = false;
  int p2p_root = -1;
  tsdf_vol* p2p_vol = nullptr;     // the volume whose raw frame slots are exported / imported
  unsigned* d_flags = nullptr;     // this rank's flag words (only the root's are used)
  void* d_handles = nullptr;       // 4 cudaIpcMemHandle_t staged for the one-time exchange
  void* peer_raw[3] = {nullptr, nullptr, nullptr};  // root's raw slots (non-root ranks)
  unsigned* peer_flags = nullptr;  // root's flag words (non-root ranks)
  int* h_err = nullptr;            // pinned, device-visible: set by a wait kernel on timeout
  int* hd_err = nullptr;
  int bseq = 0;                    // frames distributed so far (same on every rank)
};

static_assert(sizeof(ncclUniqueId) == 128, "uid128 contract");

extern "C" int tsdf_comm_unique_id(void* uid128) {
  REQUIRE(uid128, "null argument");
  int rc = load_nccl();
  if (rc) return rc;
  ncclUniqueId id;
  NCCL_CHECK(g_nccl.GetUniqueId(&id));
  memcpy(uid128, &id, 128);
  return TSDF_OK;
}

extern "C" int tsdf_comm_create(const void* uid128, int nranks, int rank, int device, tsdf_comm** out) {
  REQUIRE(uid128 && out && nranks >= 1 && rank >= 0 && rank < nranks, "bad argument");
  *out = nullptr;
  int rc = load_nccl();
  if (rc) return rc;
  int ndev = 0;
  if ((rc = tsdf_device_count(&ndev))) return rc;
  REQUIRE(device >= 0 && device < ndev, "device ordinal out of range");
  // ncclCommInitRank binds to the CURRENT device: leave it set for the lifetime of the call
  DeviceGuard g(device);
  tsdf_comm* c = new (std::nothrow) tsdf_comm();
  if (!c) { tsdf_set_error("tsdf_comm_create: host allocation failed"); return TSDF_E_NOMEM; }
  c->nranks = nranks; c->rank = rank; c->device = device;
  ncclUniqueId id;
  memcpy(&id, uid128, 128);
  // The per-frame payload is ~2 MB: a couple of channels saturate nothing but are plenty, while
  // every extra NCCL CTA spins on an SM the integrate kernel wants (measured: kernel time under
  // the overlapped broadcast).  TSDF_B200_NCCL_MAX_CTAS overrides (0 = NCCL's default).
  int max_ctas = 4;
  if (const char* e = getenv("TSDF_B200_NCCL_MAX_CTAS")) max_ctas = atoi(e);
  ncclResult_t r;
  if (g_nccl.CommInitRankConfig && max_ctas > 0) {
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    cfg.minCTAs = 1;
    cfg.maxCTAs = max_ctas;
    r = g_nccl.CommInitRankConfig(&c->comm, nranks, id, rank, &cfg);
  } else {
    r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  }
  if (r != ncclSuccess) {
    tsdf_set_error("ncclCommInitRank: %s", g_nccl.GetErrorString(r));
    delete c;
    return TSDF_E_NCCL;
  }
  if (cudaMalloc(&c->d_flags, kFlagWords * 4) != cudaSuccess || cudaMemset(c->d_flags, 0, kFlagWords * 4) != cudaSuccess ||
      cudaMalloc(&c->d_handles, 4 * sizeof(cudaIpcMemHandle_t)) != cudaSuccess ||
      cudaHostAlloc(&c->h_err, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess ||
      cudaHostGetDevicePointer(reinterpret_cast<void**>(&c->hd_err), c->h_err, 0) != cudaSuccess ||
      cudaMalloc(&c->d_counts, sizeof(int64_t) * (size_t)nranks) != cudaSuccess) {
    g_nccl.CommDestroy(c->comm);
    delete c;
    tsdf_set_error("tsdf_comm_create: device allocation failed");
    return TSDF_E_NOMEM;
  }
  *c->h_err = 0;
  *out = c;
  return TSDF_OK;
}

extern "C" int tsdf_comm_destroy(tsdf_comm* c) {
  if (!c) return TSDF_OK;
  DeviceGuard g(c->device);
  cudaDeviceSynchronize();   // our use of the peers' memory is over
  for (void* p : c->peer_raw) if (p) cudaIpcCloseMemHandle(p);
  if (c->peer_flags) cudaIpcCloseMemHandle(c->peer_flags);
  if (c->p2p_ok && c->comm && c->p2p_vol) {
    // every rank agreed on p2p_ok, so every rank gets here: a barrier (tiny all-gather) keeps the
    // root from freeing its exported slots before the importers have closed their mappings
    cudaStream_t st = nullptr;
    if (g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclInt64, c->comm, st) == ncclSuccess)
      cudaStreamSynchronize(st);
  }
  cudaFree(c->d_flags); cudaFree(c->d_handles);
  if (c->h_err) cudaFreeHost(c->h_err);
  cudaFree(c->d_counts);
  if (c->comm) g_nccl.CommDestroy(c->comm);
  delete c;
  return TSDF_OK;
}

extern "C" int tsdf_comm_peer_pull(const tsdf_comm* c, int* active) {
  REQUIRE(c && active, "null argument");
  *active = c->p2p_ok ? 1 : 0;
  return TSDF_OK;
}

// defined in api.cu
int bcast_stage(tsdf_vol* v, bool is_root, const float* depth, const uint8_t* rgb, bool color,
                FrameSlot** slot, void** d_depth, void** d_rgb, const void** send_depth,
                const void** send_rgb, bool into_slot);
int bcast_finish(tsdf_vol* v, FrameSlot* slot, bool color, const float* K, const float* T_wc,
                 float obs_weight, const void* src_raw, int (*after_prep)(void*, cudaStream_t), void* after_arg);
int ensure_frame_slot(tsdf_vol* v, int index);

// One-time collective: export the root's three raw frame slots and its flag words with cudaIpc,
// ship the handles with one ncclBroadcast, open them on the other ranks, and agree (all-gather
// of a status word) whether every rank succeeded.  Any failure leaves the NCCL path in place.
static int p2p_setup(tsdf_vol* v, tsdf_comm* c, int root) {
  c->p2p_tried = true;
  if (const char* t = getenv("TSDF_B200_PEER_TIMEOUT_MS")) {
    const long long ms = atoll(t);
    if (ms > 0) g_wait_timeout_ns = (unsigned long long)ms * 1000000ull;
  }
  const char* mode = getenv("TSDF_B200_BCAST");
  bool ok = !(mode && strcmp(mode, "nccl") == 0) && c->nranks > 1;
  for (int k = 0; k < 3 && ok; ++k) ok = ensure_frame_slot(v, k) == TSDF_OK;
  cudaIpcMemHandle_t hs[4];
  memset(hs, 0, sizeof hs);
  if (ok && c->rank == root) {
    for (int k = 0; k < 3 && ok; ++k) ok = cudaIpcGetMemHandle(&hs[k], v->slot[k].d_raw) == cudaSuccess;
    ok = ok && cudaIpcGetMemHandle(&hs[3], c->d_flags) == cudaSuccess;
    cudaGetLastError();
  }
  if (c->rank == root) CU_CHECK(cudaMemcpyAsync(c->d_handles, hs, sizeof hs, cudaMemcpyHostToDevice, v->s_copy));
  NCCL_CHECK(g_nccl.Broadcast(c->d_handles, c->d_handles, sizeof hs, ncclUint8, root, c->comm, v->s_copy));
  CU_CHECK(cudaMemcpyAsync(hs, c->d_handles, sizeof hs, cudaMemcpyDeviceToHost, v->s_copy));
  CU_CHECK(cudaStreamSynchronize(v->s_copy));
  if (ok && c->rank != root) {
    void* p = nullptr;
    for (int k = 0; k < 3 && ok; ++k) {
      ok = cudaIpcOpenMemHandle(&p, hs[k], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      if (ok) c->peer_raw[k] = p;
    }
    if (ok) {
      ok = cudaIpcOpenMemHandle(&p, hs[3], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      if (ok) c->peer_flags = reinterpret_cast<unsigned*>(p);
    }
    cudaGetLastError();
  }
  // agreement: every rank must have succeeded (the root's handles are all-zero if it failed)
  int64_t mine = ok ? 1 : 0;
  CU_CHECK(cudaMemcpyAsync(c->d_counts + c->rank, &mine, 8, cudaMemcpyHostToDevice, v->s_copy));
  NCCL_CHECK(g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclInt64, c->comm, v->s_copy));
  std::vector<int64_t> all((size_t)c->nranks, 0);
  CU_CHECK(cudaMemcpyAsync(all.data(), c->d_counts, 8 * (size_t)c->nranks, cudaMemcpyDeviceToHost, v->s_copy));
  CU_CHECK(cudaStreamSynchronize(v->s_copy));
  bool all_ok = true;
  for (int64_t a : all) all_ok = all_ok && a == 1;
  c->p2p_ok = all_ok;
  c->p2p_root = root;
  c->p2p_vol = v;
  return TSDF_OK;
}

static int set_done_flag(void* arg, cudaStream_t st) {   // after_prep hook of a non-root rank
  tsdf_comm* c = static_cast<tsdf_comm*>(arg);
  k_set_flag<<<1, 1, 0, st>>>(c->peer_flags + 8 + c->rank, (unsigned)c->bseq);
  CU_LAUNCHED();
  return TSDF_OK;
}

extern "C" int tsdf_integrate_bcast_async(tsdf_vol* v, tsdf_comm* c, int root, const float* depth,
                                          const uint8_t* rgb, const float* K, const float* T_wc,
                                          float obs_weight) {
  REQUIRE(v && c && T_wc, "null argument (pose metadata is required on every rank)");
  REQUIRE(root >= 0 && root < c->nranks, "bad root");
  REQUIRE(v->cfg.device == c->device, "volume and communicator live on different devices");
  const bool is_root = c->rank == root;
  const bool color = v->d_color != nullptr;  // every rank decides the same way
  REQUIRE(!is_root || depth, "root needs a depth image");
  REQUIRE(!is_root || !color || rgb, "root needs an rgb image for a colour volume");
  DeviceGuard g(v->cfg.device);
  FrameSlot* slot = nullptr;
  void *d_depth = nullptr, *d_rgb = nullptr;
  const void *s_depth = nullptr, *s_rgb = nullptr;
  int rc;
  if (!c->p2p_tried && c->nranks > 1 && (rc = p2p_setup(v, c, root))) return rc;
  if (*c->h_err) {
    tsdf_set_error("peer-pull frame distribution: flag wait timed out (code %d): a rank is gone or stalled", *c->h_err);
    return TSDF_E_NCCL;
  }
  if (c->p2p_ok && root == c->p2p_root && v == c->p2p_vol) {
    // ---- peer-pull: fused broadcast + prep (see the comment at k_wait_flag) ----------------
    const int seq = ++c->bseq;
    const int k = (int)(v->seq % 3);   // the slot bcast_stage picks; the same on every rank (SPMD)
    if (is_root) {
      // slot k last held frame seq-3: wait until every peer has consumed it, then overwrite
      k_wait_all<<<1, 32, 0, v->s_copy>>>(c->d_flags + 8, c->nranks, root, seq - 3, c->hd_err, g_wait_timeout_ns);
      CU_LAUNCHED();
      rc = bcast_stage(v, true, depth, rgb, color, &slot, &d_depth, &d_rgb, &s_depth, &s_rgb, true);
      if (rc) return rc;
      k_set_flag<<<1, 1, 0, v->s_copy>>>(c->d_flags, (unsigned)seq);   // ready: frame seq is in slot k
      CU_LAUNCHED();
      return bcast_finish(v, slot, color, K, T_wc, obs_weight, nullptr, nullptr, nullptr);
    }
    rc = bcast_stage(v, false, nullptr, nullptr, color, &slot, &d_depth, &d_rgb, &s_depth, &s_rgb, false);
    if (rc) return rc;
    k_wait_flag<<<1, 1, 0, v->s_copy>>>(c->peer_flags, seq, c->hd_err, g_wait_timeout_ns);
    CU_LAUNCHED();
    return bcast_finish(v, slot, color, K, T_wc, obs_weight, c->peer_raw[k], set_done_flag, c);
  }
  rc = bcast_stage(v, is_root, depth, rgb, color, &slot, &d_depth, &d_rgb, &s_depth, &s_rgb, false);
  if (rc) return rc;
  const size_t px = frame_pixels(v->cfg);
  if (c->nranks > 1) {
    NCCL_CHECK(g_nccl.GroupStart());
    NCCL_CHECK(g_nccl.Broadcast(s_depth, d_depth, px * 4, ncclUint8, root, c->comm, v->s_copy));
    if (color) NCCL_CHECK(g_nccl.Broadcast(s_rgb, d_rgb, px * 3, ncclUint8, root, c->comm, v->s_copy));
    NCCL_CHECK(g_nccl.GroupEnd());
  } else {
    if (s_depth != d_depth) CU_CHECK(cudaMemcpyAsync(d_depth, s_depth, px * 4, cudaMemcpyDeviceToDevice, v->s_copy));
    if (color && s_rgb != d_rgb) CU_CHECK(cudaMemcpyAsync(d_rgb, s_rgb, px * 3, cudaMemcpyDeviceToDevice, v->s_copy));
  }
  return bcast_finish(v, slot, color, K, T_wc, obs_weight, nullptr, nullptr, nullptr);
}

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
  // 1. local extraction into device buffers (idx always: it is the merge key)
  int64_t n_local = 0;
  rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &n_local);
  if (rc) return rc;
  float* d_xyz = nullptr; uint8_t* d_rgb = nullptr; int64_t* d_idx = nullptr;
  float* r_xyz = nullptr; uint8_t* r_rgb = nullptr; int64_t* r_idx = nullptr;
  auto cleanup = [&]() { cudaFree(d_xyz); cudaFree(d_rgb); cudaFree(d_idx); cudaFree(r_xyz); cudaFree(r_rgb); cudaFree(r_idx); };
  const size_t nl = (size_t)std::max<int64_t>(n_local, 1);
  if (cudaMalloc(&d_xyz, nl * 12) || cudaMalloc(&d_idx, nl * 8) || (want_rgb && cudaMalloc(&d_rgb, nl * 3))) {
    cleanup(); tsdf_set_error("tsdf_gather_points: device allocation failed"); return TSDF_E_NOMEM;
  }
  if (n_local) {
    rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, d_xyz, d_rgb, d_idx, n_local, &n_local);
    if (rc) { cleanup(); return rc; }
  }
  // 2. counts
  cudaStream_t st = v->s_compute;
  std::vector<int64_t> counts((size_t)c->nranks, 0);
  if (c->nranks > 1) {
    cudaMemcpyAsync(c->d_counts + c->rank, &n_local, 8, cudaMemcpyHostToDevice, st);
    ncclResult_t r = g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclInt64, c->comm, st);
    if (r != ncclSuccess) { cleanup(); tsdf_set_error("ncclAllGather: %s", g_nccl.GetErrorString(r)); return TSDF_E_NCCL; }
    cudaMemcpyAsync(counts.data(), c->d_counts, 8 * (size_t)c->nranks, cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) { cleanup(); tsdf_set_error("gather: sync failed"); return TSDF_E_CUDA; }
  } else {
    counts[0] = n_local;
  }
  int64_t total = 0;
  std::vector<int64_t> offs((size_t)c->nranks + 1, 0);
  for (int r = 0; r < c->nranks; ++r) { offs[r] = total; total += counts[r]; }
  offs[c->nranks] = total;
  *count = is_root ? total : n_local;
  if (cap < 0) { cleanup(); return TSDF_OK; }  // collective count-only call (every rank passes cap < 0)
  // 3. records to the root (every rank takes part even if the root's buffers are too small:
  //    the root then receives into scratch and reports TSDF_E_CAPACITY afterwards)
  const size_t nt = (size_t)std::max<int64_t>(total, 1);
  if (is_root) {
    if (cudaMalloc(&r_xyz, nt * 12) || cudaMalloc(&r_idx, nt * 8) || (want_rgb && cudaMalloc(&r_rgb, nt * 3))) {
      cleanup(); tsdf_set_error("tsdf_gather_points: root allocation failed"); return TSDF_E_NOMEM;
    }
  }
  if (c->nranks > 1) {
    g_nccl.GroupStart();
    if (!is_root) {
      if (n_local) {
        g_nccl.Send(d_xyz, (size_t)n_local * 3, ncclFloat32, root, c->comm, st);
        g_nccl.Send(d_idx, (size_t)n_local, ncclInt64, root, c->comm, st);
        if (want_rgb) g_nccl.Send(d_rgb, (size_t)n_local * 3, ncclUint8, root, c->comm, st);
      }
    } else {
      for (int r = 0; r < c->nranks; ++r) {
        if (r == root || counts[r] == 0) continue;
        g_nccl.Recv(r_xyz + offs[r] * 3, (size_t)counts[r] * 3, ncclFloat32, r, c->comm, st);
        g_nccl.Recv(r_idx + offs[r], (size_t)counts[r], ncclInt64, r, c->comm, st);
        if (want_rgb) g_nccl.Recv(r_rgb + offs[r] * 3, (size_t)counts[r] * 3, ncclUint8, r, c->comm, st);
      }
    }
    ncclResult_t r = g_nccl.GroupEnd();
    if (r != ncclSuccess) { cleanup(); tsdf_set_error("gather send/recv: %s", g_nccl.GetErrorString(r)); return TSDF_E_NCCL; }
  }
  if (is_root && n_local) {
    cudaMemcpyAsync(r_xyz + offs[root] * 3, d_xyz, (size_t)n_local * 12, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(r_idx + offs[root], d_idx, (size_t)n_local * 8, cudaMemcpyDeviceToDevice, st);
    if (want_rgb) cudaMemcpyAsync(r_rgb + offs[root] * 3, d_rgb, (size_t)n_local * 3, cudaMemcpyDeviceToDevice, st);
  }
  if (cudaStreamSynchronize(st) != cudaSuccess) { cleanup(); tsdf_set_error("gather: %s", cudaGetErrorString(cudaGetLastError())); return TSDF_E_CUDA; }
  rc = TSDF_OK;
  if (is_root && (xyz || rgb || idx)) {
    if (cap < total) {
      tsdf_set_error("tsdf_gather_points: capacity %lld < %lld points", (long long)cap, (long long)total);
      rc = TSDF_E_CAPACITY;
    } else if (total) {
      // 4. D2H, then k-way merge by GLOBAL reference index.  x-fast slabs are contiguous index
      //    ranges (concatenation is already sorted); z-fast slabs interleave per (x,y) row.
      std::vector<float> hx(nt * 3); std::vector<int64_t> hi(nt); std::vector<uint8_t> hc(want_rgb ? nt * 3 : 0);
      cudaMemcpy(hx.data(), r_xyz, (size_t)total * 12, cudaMemcpyDeviceToHost);
      cudaMemcpy(hi.data(), r_idx, (size_t)total * 8, cudaMemcpyDeviceToHost);
      if (want_rgb) cudaMemcpy(hc.data(), r_rgb, (size_t)total * 3, cudaMemcpyDeviceToHost);
      std::vector<int64_t> cur(offs.begin(), offs.end() - 1);
      for (int64_t o = 0; o < total; ++o) {
        int best = -1; int64_t bi = 0;
        for (int r = 0; r < c->nranks; ++r)
          if (cur[r] < offs[r + 1] && (best < 0 || hi[cur[r]] < bi)) { best = r; bi = hi[cur[r]]; }
        const int64_t s = cur[best]++;
        if (xyz) { xyz[3 * o] = hx[3 * s]; xyz[3 * o + 1] = hx[3 * s + 1]; xyz[3 * o + 2] = hx[3 * s + 2]; }
        if (idx) idx[o] = bi;
        if (rgb) {
          if (want_rgb) { rgb[3 * o] = hc[3 * s]; rgb[3 * o + 1] = hc[3 * s + 1]; rgb[3 * o + 2] = hc[3 * s + 2]; }
          else { rgb[3 * o] = rgb[3 * o + 1] = rgb[3 * o + 2] = 0; }
        }
      }
    }
  }
  cleanup();
  return rc;
}
