// api.cu — the C ABI of include/tsdf_b200.h: lifecycle, frame upload path, integrate
// entry points, readback, file writers, host-side pose helpers (K2-K4, K6-K8, K10).
// Host code is plain C++ calling the CUDA runtime; there is NO CPU fallback: every entry
// point that needs a device fails with TSDF_E_NODEVICE / TSDF_E_CUDA when none is present.
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <limits>
#include <mutex>
#include <thread>

#include <cuda.h>   // CUtensorMap types only: the encoder is fetched with cudaGetDriverEntryPoint (no -lcuda)

#include "common.cuh"

// ---- errors ------------------------------------------------------------------------------
static thread_local char g_err[768] = "";
void tsdf_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* tsdf_last_error(void) { return g_err; }
std::atomic<long long> g_tsdf_launches{0};
extern "C" int tsdf_launch_count(int64_t* n) {
  REQUIRE(n, "null argument");
  *n = (int64_t)g_tsdf_launches.load(std::memory_order_relaxed);
  return TSDF_OK;
}
extern "C" int tsdf_version(void) { return TSDF_B200_VERSION; }

extern "C" int tsdf_device_count(int* n) {
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (n) *n = (e == cudaSuccess) ? c : 0;
  if (e != cudaSuccess || c == 0) {
    cudaGetLastError();
    tsdf_set_error("no CUDA device (%s); this library has no CPU fallback",
                   e == cudaSuccess ? "count=0" : cudaGetErrorString(e));
    return TSDF_E_NODEVICE;
  }
  return TSDF_OK;
}

// ---- host pose helpers (K6) ----------------------------------------------------------------
// Restated from the algorithm of reference src/tsdf.cu:253-273 / :276-403 with the same
// operation order (left-to-right fp32 sums; det reciprocal in double, narrowed to float).
// This translation unit is compiled with -fmad=false and host code is never contracted.
extern "C" void tsdf_multiply_matrix(const float a[16], const float b[16], float o[16]) {
  float t[16];
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) {
      volatile float s = a[r * 4 + 0] * b[0 * 4 + c];
      s = s + a[r * 4 + 1] * b[1 * 4 + c];
      s = s + a[r * 4 + 2] * b[2 * 4 + c];
      s = s + a[r * 4 + 3] * b[3 * 4 + c];
      t[r * 4 + c] = s;
    }
  memcpy(o, t, sizeof(t));
}

namespace {
// signed sum of six triple products, evaluated in source order
struct Tri { int a, b, c; };
float cof(const float* m, bool first_neg, const Tri (&t)[6]) {
  // sign pattern of the reference: (+ - - + + -) or (- + + - - +)
  static const int P[6] = {+1, -1, -1, +1, +1, -1};
  volatile float acc = (first_neg ? -m[t[0].a] : m[t[0].a]) * m[t[0].b] * m[t[0].c];
  for (int i = 1; i < 6; ++i) {
    volatile float u = m[t[i].a] * m[t[i].b];
    u = u * m[t[i].c];
    int sg = first_neg ? -P[i] : P[i];
    acc = sg > 0 ? acc + u : acc - u;
  }
  return acc;
}
}  // namespace

extern "C" int tsdf_invert_matrix(const float m[16], float out[16]) {
  float inv[16];
  inv[0]  = cof(m, false, {{5,10,15},{5,11,14},{9,6,15},{9,7,14},{13,6,11},{13,7,10}});
  inv[4]  = cof(m, true,  {{4,10,15},{4,11,14},{8,6,15},{8,7,14},{12,6,11},{12,7,10}});
  inv[8]  = cof(m, false, {{4,9,15},{4,11,13},{8,5,15},{8,7,13},{12,5,11},{12,7,9}});
  inv[12] = cof(m, true,  {{4,9,14},{4,10,13},{8,5,14},{8,6,13},{12,5,10},{12,6,9}});
  inv[1]  = cof(m, true,  {{1,10,15},{1,11,14},{9,2,15},{9,3,14},{13,2,11},{13,3,10}});
  inv[5]  = cof(m, false, {{0,10,15},{0,11,14},{8,2,15},{8,3,14},{12,2,11},{12,3,10}});
  inv[9]  = cof(m, true,  {{0,9,15},{0,11,13},{8,1,15},{8,3,13},{12,1,11},{12,3,9}});
  inv[13] = cof(m, false, {{0,9,14},{0,10,13},{8,1,14},{8,2,13},{12,1,10},{12,2,9}});
  inv[2]  = cof(m, false, {{1,6,15},{1,7,14},{5,2,15},{5,3,14},{13,2,7},{13,3,6}});
  inv[6]  = cof(m, true,  {{0,6,15},{0,7,14},{4,2,15},{4,3,14},{12,2,7},{12,3,6}});
  inv[10] = cof(m, false, {{0,5,15},{0,7,13},{4,1,15},{4,3,13},{12,1,7},{12,3,5}});
  inv[14] = cof(m, true,  {{0,5,14},{0,6,13},{4,1,14},{4,2,13},{12,1,6},{12,2,5}});
  inv[3]  = cof(m, true,  {{1,6,11},{1,7,10},{5,2,11},{5,3,10},{9,2,7},{9,3,6}});
  inv[7]  = cof(m, false, {{0,6,11},{0,7,10},{4,2,11},{4,3,10},{8,2,7},{8,3,6}});
  inv[11] = cof(m, true,  {{0,5,11},{0,7,9},{4,1,11},{4,3,9},{8,1,7},{8,3,5}});
  inv[15] = cof(m, false, {{0,5,10},{0,6,9},{4,1,10},{4,2,9},{8,1,6},{8,2,5}});
  volatile float det = m[0] * inv[0];
  det = det + m[1] * inv[4];
  det = det + m[2] * inv[8];
  det = det + m[3] * inv[12];
  if (det == 0) {
    tsdf_set_error("tsdf_invert_matrix: singular matrix");
    return TSDF_E_SINGULAR;
  }
  float rdet = (float)(1.0 / (double)det);
  for (int i = 0; i < 16; ++i) out[i] = inv[i] * rdet;
  return TSDF_OK;
}

// ---- configuration ---------------------------------------------------------------------------
extern "C" void tsdf_config_default(tsdf_config* c) {
  memset(c, 0, sizeof(*c));
  c->dim_x = c->dim_y = c->dim_z = 200;      // include/tsdf.hpp:65-67
  c->voxel_size = 0.004f;                    // include/tsdf.hpp:63
  c->trunc_margin = c->voxel_size * 5;       // include/tsdf.hpp:64
  for (int i = 0; i < 4; ++i) c->base2world[i * 5] = 1.0f;
  c->im_h = 480;
  c->im_w = 640;
  const float K[9] = {535.4f, 0, 320.1f, 0, 539.2f, 247.6f, 0, 0, 1};  // include/tsdf.hpp:96
  memcpy(c->K, K, sizeof(K));
  c->max_depth = 6.0f;                       // src/tsdf.cu:46
  c->semantics = TSDF_SEM_TSDFCU;
  c->with_color = 0;
  c->device = 0;
}

extern "C" int tsdf_config_from_bounds(const double b[6], double voxel_size, tsdf_config* c) {
  REQUIRE(b && c && voxel_size > 0, "null argument or voxel_size <= 0");
  tsdf_config_default(c);
  int dims[3];
  for (int k = 0; k < 3; ++k) {
    double ext = b[2 * k + 1] - b[2 * k];
    REQUIRE(ext > 0, "empty bounds");
    dims[k] = (int)ceil(ext / voxel_size);   // upstream: np.ceil((max-min)/voxel_size)
    c->origin[k] = (float)b[2 * k];          // upstream: vol_bnds[:,0].astype(np.float32)
  }
  c->dim_x = dims[0]; c->dim_y = dims[1]; c->dim_z = dims[2];
  c->voxel_size = (float)voxel_size;
  c->trunc_margin = (float)(voxel_size * 5.0);  // upstream: trunc = 5*voxel_size (float64), used as fp32
  c->max_depth = std::numeric_limits<float>::infinity();
  c->semantics = TSDF_SEM_FUSIONPY;
  c->with_color = 1;
  return TSDF_OK;
}

extern "C" void tsdf_slab_range(int dim_z, int nranks, int rank, int* z_begin, int* z_end) {
  // split in units of 4 slices so every slab keeps 128-bit alignment on a z-fast layout
  int groups = (dim_z + 3) / 4;
  int base = groups / nranks, rem = groups % nranks;
  int g0 = rank * base + std::min(rank, rem);
  int g1 = g0 + base + (rank < rem ? 1 : 0);
  *z_begin = std::min(g0 * 4, dim_z);
  *z_end = std::min(g1 * 4, dim_z);
}

// ---- lifecycle (K3, K4) -----------------------------------------------------------------------
static bool is_identity(const float* m) {
  for (int i = 0; i < 16; ++i)
    if (m[i] != ((i % 5 == 0) ? 1.0f : 0.0f)) return false;
  return true;
}

static void free_vol(tsdf_vol* v) {
  if (!v) return;
  DeviceGuard g(v->cfg.device);
  if (v->s_compute) cudaStreamSynchronize(v->s_compute);
  if (v->s_copy) cudaStreamSynchronize(v->s_copy);
  if (v->s_aux) cudaStreamSynchronize(v->s_aux);
  for (cudaEvent_t e : v->ev_int_done) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : v->ev_pub_done) if (e) cudaEventDestroy(e);
  auto free_slot = [](FrameSlot& s) {
    if (s.h_pinned) cudaFreeHost(s.h_pinned);
    if (s.d_raw) cudaFree(s.d_raw);
    if (s.d_frame) cudaFree(s.d_frame);
    if (s.uploaded) cudaEventDestroy(s.uploaded);
    if (s.consumed) cudaEventDestroy(s.consumed);
  };
  for (FrameSlot& s : v->slot) free_slot(s);
  for (FrameSlot& s : v->bslot) free_slot(s);
  cudaFree(v->d_tsdf); cudaFree(v->d_weight); cudaFree(v->d_color);
  cudaFree(v->d_written); cudaFree(v->d_scratch); cudaFree(v->d_mesh);
  v->out.release();
  cudaFree(v->d_dbg);
  if (v->h_written) cudaFreeHost(v->h_written);
  for (cudaEvent_t e : {v->ev_k0, v->ev_k1, v->ev_u0, v->ev_u1, v->ev_t0, v->ev_t1}) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : v->kev) cudaEventDestroy(e);
  for (cudaEvent_t e : v->tev) cudaEventDestroy(e);
  if (v->s_compute) cudaStreamDestroy(v->s_compute);
  if (v->s_copy) cudaStreamDestroy(v->s_copy);
  if (v->s_aux) cudaStreamDestroy(v->s_aux);
  delete v;
}

static int create_impl(const tsdf_config* cfg, tsdf_vol* v) {
  v->cfg = *cfg;
  tsdf_config& c = v->cfg;
  if (c.z_begin == 0 && c.z_end == 0) c.z_end = c.dim_z;
  REQUIRE(c.dim_x > 0 && c.dim_y > 0 && c.dim_z > 0, "non-positive grid dims");
  REQUIRE(c.z_begin >= 0 && c.z_end > c.z_begin && c.z_end <= c.dim_z, "bad slab range");
  REQUIRE(c.voxel_size > 0 && c.trunc_margin > 0, "voxel_size/trunc_margin must be > 0");
  REQUIRE(c.im_h > 0 && c.im_w > 0, "bad image shape");
  REQUIRE(c.semantics == TSDF_SEM_TSDFCU || c.semantics == TSDF_SEM_FUSIONPY, "bad semantics");
  v->dzl = c.z_end - c.z_begin;
  v->n_local = (int64_t)c.dim_x * c.dim_y * v->dzl;
  v->zfast = c.semantics == TSDF_SEM_FUSIONPY;
  v->identity_base = is_identity(c.base2world);
  int rc = tsdf_invert_matrix(c.base2world, v->base2world_inv);  // src/tsdf.cu:74 (result checked here)
  if (rc) return rc;
  int ndev = 0;
  rc = tsdf_device_count(&ndev);
  if (rc) return rc;
  REQUIRE(c.device >= 0 && c.device < ndev, "device ordinal out of range");
  DeviceGuard g(c.device);
  size_t bytes = (size_t)v->n_local * sizeof(float);
  CU_CHECK(cudaMalloc(&v->d_tsdf, bytes));
  CU_CHECK(cudaMalloc(&v->d_weight, bytes));
  if (c.with_color) CU_CHECK(cudaMalloc(&v->d_color, bytes));
  CU_CHECK(cudaMalloc(&v->d_written, sizeof(unsigned long long) * kWrittenRing));
  CU_CHECK(cudaHostAlloc(&v->h_written, sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable));
  CU_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void**>(&v->hd_written), v->h_written, 0));
  *v->h_written = 0;
  // The copy stream (H2D, NCCL broadcast, frame prep) gets the highest priority: its few small
  // CTAs must be dispatched ahead of the tens of thousands of integrate blocks still queued on
  // the compute stream, otherwise frame i+1 only arrives in the tail of integrate i and every
  // step pays the broadcast latency.
  int prio_least = 0, prio_greatest = 0;
  CU_CHECK(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
  CU_CHECK(cudaStreamCreateWithPriority(&v->s_compute, cudaStreamNonBlocking, prio_least));
  CU_CHECK(cudaStreamCreateWithPriority(&v->s_copy, cudaStreamNonBlocking, prio_greatest));
  CU_CHECK(cudaStreamCreateWithPriority(&v->s_aux, cudaStreamNonBlocking, prio_greatest));
  // on the stream whose kernels add into the counters: a plain cudaMemset runs on the legacy default
  // stream, which a cudaStreamNonBlocking stream does not wait for
  CU_CHECK(cudaMemsetAsync(v->d_written, 0, sizeof(unsigned long long) * kWrittenRing, v->s_compute));
  for (int i = 0; i < kWrittenRing; ++i) {
    CU_CHECK(cudaEventCreateWithFlags(&v->ev_int_done[i], cudaEventDisableTiming));
    CU_CHECK(cudaEventCreateWithFlags(&v->ev_pub_done[i], cudaEventDisableTiming));
  }
  CU_CHECK(cudaEventCreate(&v->ev_k0));
  CU_CHECK(cudaEventCreate(&v->ev_k1));
  CU_CHECK(cudaEventCreate(&v->ev_u0));
  CU_CHECK(cudaEventCreate(&v->ev_u1));
  CU_CHECK(cudaEventCreate(&v->ev_t0));
  CU_CHECK(cudaEventCreate(&v->ev_t1));
  return tsdf_reset(v);
}

extern "C" int tsdf_create(const tsdf_config* cfg, tsdf_vol** out) {
  REQUIRE(cfg && out, "null argument");
  *out = nullptr;
  tsdf_vol* v = new (std::nothrow) tsdf_vol();
  if (!v) { tsdf_set_error("tsdf_create: host allocation failed"); return TSDF_E_NOMEM; }
  int rc = create_impl(cfg, v);
  if (rc) { free_vol(v); return rc; }
  *out = v;
  return TSDF_OK;
}

extern "C" int tsdf_destroy(tsdf_vol* v) {
  if (v && v->attached) {
    // its frame slots are exported to / mapped by the other ranks of a communicator
    tsdf_set_error("tsdf_destroy: the volume is attached to a tsdf_comm; call tsdf_comm_destroy first");
    return TSDF_E_INVALID;
  }
  free_vol(v);
  return TSDF_OK;
}

extern "C" int tsdf_reset(tsdf_vol* v) {
  if (v) ++v->epoch;  // voxel arrays change: invalidates the cached extraction mark pass
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  int rc = launch_fill(v->d_tsdf, 1.0f, v->n_local, v->s_compute);  // src/tsdf.cu:79-80
  if (rc) return rc;
  CU_CHECK(cudaMemsetAsync(v->d_weight, 0, (size_t)v->n_local * 4, v->s_compute));  // :81
  if (v->d_color) CU_CHECK(cudaMemsetAsync(v->d_color, 0, (size_t)v->n_local * 4, v->s_compute));
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  v->frames = 0;
  v->last_written = 0;
  v->intw = true;
  v->w_bound = 0;
  return TSDF_OK;
}

extern "C" int tsdf_get_config(const tsdf_vol* v, tsdf_config* cfg) {
  REQUIRE(v && cfg, "null argument");
  *cfg = v->cfg;
  return TSDF_OK;
}

// ---- frame upload path (K2) ----------------------------------------------------------------
// Pinned, multi-buffered staging; H2D + prep run on the copy stream and are chained to the
// compute stream by events, so the upload of frame i+1 overlaps the integrate of frame i.
// (The reference does two blocking cudaMemcpy from pageable memory per frame per object,
// src/tsdf.cu:161-162.)
// Tensor map of a prepared colour frame for the staged-tile variant of the sparse kernel
// (integrate.cu, `Stage`): the float4 image as a 2-D [H][4W] fp32 tensor, box = 64 x 32 pixels.
static bool encode_frame_tmap(const tsdf_vol* v, FrameSlot& s) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) { cudaGetLastError(); p = nullptr; }
    return reinterpret_cast<EncodeFn>(p);
  }();
  const int H = v->cfg.im_h, W = v->cfg.im_w;
  if (!fn || !v->d_color || W < 64 || H < 32) return false;
  static_assert(sizeof(CUtensorMap) == 128, "tensor map size");
  const cuuint64_t gdim[2] = {(cuuint64_t)W * 4, (cuuint64_t)H};
  const cuuint64_t gstride[1] = {(cuuint64_t)W * 16};
  const cuuint32_t box[2] = {256, 32}, estr[2] = {1, 1};
  CUtensorMap tm;
  if (fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, s.d_frame, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  memcpy(s.tmap, &tm, 128);
  return true;
}

static int ensure_slot(tsdf_vol* v, FrameSlot& s) {
  const size_t px = frame_pixels(v->cfg);
  if (!s.d_raw) {
    CU_CHECK(cudaMalloc(&s.d_raw, px * 7 + 64));        // depth (<=4 B) + rgb (3 B) per pixel
    CU_CHECK(cudaMalloc(&s.d_frame, frame_alloc_bytes(v->cfg.im_h, v->cfg.im_w)));  // float4 per pixel (float when no colour) + tail
    // tail = 0 from the start: preps that skip the memset rely on k_publish having cleared it.  Issued on
    // the copy stream, where every prep into this slot runs (a plain cudaMemset would go to the legacy
    // default stream, which the non-blocking copy stream is not ordered behind)
    CU_CHECK(cudaMemsetAsync(frame_dmax_ptr(s.d_frame, px), 0, frame_tail_bytes(v->cfg.im_h, v->cfg.im_w), v->s_copy));
    CU_CHECK(cudaEventCreateWithFlags(&s.uploaded, cudaEventDisableTiming));
    CU_CHECK(cudaEventCreateWithFlags(&s.consumed, cudaEventDisableTiming));
    s.tmap_ok = encode_frame_tmap(v, s);
  }
  return TSDF_OK;
}

int ensure_bcast_slot(tsdf_vol* v, int index) { return ensure_slot(v, v->bslot[index]); }

enum PtrKind { PTR_PAGEABLE, PTR_PINNED, PTR_DEVICE };
static PtrKind ptr_kind(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return PTR_PAGEABLE; }
  if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return PTR_DEVICE;
  if (a.type == cudaMemoryTypeHost) return PTR_PINNED;
  return PTR_PAGEABLE;
}

// ---- readback buffers ------------------------------------------------------------------------
int OutBuffers::need_dev(int i, size_t bytes) {
  if (bytes <= dcap[i]) return TSDF_OK;
  if (d[i]) { cudaFree(d[i]); d[i] = nullptr; dcap[i] = 0; }
  const size_t want = bytes + bytes / 4 + 4096;   // grow with slack: successive calls rarely re-allocate
  CU_CHECK(cudaMalloc(&d[i], want));
  dcap[i] = want;
  return TSDF_OK;
}
int OutBuffers::need_host(size_t bytes) {
  if (bytes <= hcap) return TSDF_OK;
  if (h) { cudaFreeHost(h); h = nullptr; hcap = 0; }
  CU_CHECK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
  hcap = bytes;
  return TSDF_OK;
}
void OutBuffers::release() {
  for (int i = 0; i < 2; ++i) { if (ev[i]) cudaEventDestroy(ev[i]); ev[i] = nullptr; }
  ev_dev = -1;
  for (int i = 0; i < 4; ++i) { if (d[i]) cudaFree(d[i]); d[i] = nullptr; dcap[i] = 0; }
  if (h) cudaFreeHost(h);
  h = nullptr; hcap = 0;
}

int copy_to_host(OutBuffers* ob, void* dst, const void* d_src, size_t bytes, cudaStream_t st) {
  if (!bytes) return TSDF_OK;
  if (ptr_kind(dst) == PTR_PINNED) {
    CU_CHECK(cudaMemcpyAsync(dst, d_src, bytes, cudaMemcpyDeviceToHost, st));   // the CALLER syncs `st`
    return TSDF_OK;
  }
  // pageable destination: D2H into pinned staging at full PCIe rate, chunk i+1 in flight while the
  // host copies chunk i out (cudaMemcpy to pageable memory does the same at a fraction of the rate)
  constexpr size_t kChunk = 4u << 20;
  int rc = ob->need_host(2 * kChunk);
  if (rc) return rc;
  // events belong to a device: (re)create them on the one `st` lives on (= the current device)
  int cur = 0;
  CU_CHECK(cudaGetDevice(&cur));
  if (ob->ev_dev != cur) {
    for (int i = 0; i < 2; ++i) {
      if (ob->ev[i]) cudaEventDestroy(ob->ev[i]);
      ob->ev[i] = nullptr;
      CU_CHECK(cudaEventCreateWithFlags(&ob->ev[i], cudaEventDisableTiming));
    }
    ob->ev_dev = cur;
  }
  cudaEvent_t* ev = ob->ev;
  const size_t n_chunks = (bytes + kChunk - 1) / kChunk;
  auto issue = [&](size_t c) -> cudaError_t {
    const size_t off = c * kChunk, len = std::min(kChunk, bytes - off);
    cudaError_t e = cudaMemcpyAsync(static_cast<char*>(ob->h) + (c & 1) * kChunk, static_cast<const char*>(d_src) + off, len,
                                    cudaMemcpyDeviceToHost, st);
    return e != cudaSuccess ? e : cudaEventRecord(ev[c & 1], st);
  };
  CU_CHECK(issue(0));
  for (size_t c = 0; c < n_chunks; ++c) {
    CU_CHECK(cudaEventSynchronize(ev[c & 1]));
    if (c + 1 < n_chunks) CU_CHECK(issue(c + 1));   // lands in the other half
    const size_t off = c * kChunk, len = std::min(kChunk, bytes - off);
    memcpy(static_cast<char*>(dst) + off, static_cast<const char*>(ob->h) + (c & 1) * kChunk, len);
  }
  return TSDF_OK;
}

// ---- pageable -> pinned staging copy on several host threads -----------------------------------
// The reference hands over pageable memory (cv::Mat data, src/Object.cpp:160-164), so the drop-in
// path starts with a host copy into the pinned ring; on one thread that copy (~12 GB/s) costs more
// than the sparse kernels' whole frame (2.15 MB: 0.10-0.12 ms on the GPU box against 0.09-0.14 ms).
// Helper threads (TSDF_B200_COPY_THREADS = n, OFF by default) split copies of 256 KB and more with the
// caller: 0.022 ms for the same 2.15 MB with 3 helpers.  Measured on the B200 box, three alternating
// rounds (profiles/r2_v5_dropin_ab.jsonl): an asynchronous C2 stream gains (5400 -> 6700 frames/s),
// but the helpers' sleep / wake-up latency makes the blocking 200^3 call erratic (0.15-0.34 ms against
// a steady 0.19) and costs the batched-instance path 18 % (1640 -> 1330 keyframes/s) — hence opt-in.
// The pool is created on first use, leaked on purpose (detached threads, nothing to destroy at exit)
// and re-created in a forked child; a second caller that finds it busy copies on its own thread.
namespace {
struct CopyPool {
  std::mutex m;
  std::condition_variable cv_work, cv_done;
  const char* src = nullptr;
  char* dst = nullptr;
  size_t bytes = 0, part = 0;
  int parts = 0, next = 0, pending = 0;
  uint64_t gen = 0;
  int n_threads = 0;
  std::mutex busy;   // one parallel copy at a time
};
CopyPool* g_copy_pool = nullptr;
long g_copy_pool_pid = 0;
std::mutex g_copy_pool_mu;

void copy_worker(CopyPool* p) {
  uint64_t seen = 0;
  std::unique_lock<std::mutex> lk(p->m);
  for (;;) {
    p->cv_work.wait(lk, [&] { return p->gen != seen; });
    seen = p->gen;
    while (p->next < p->parts) {
      const int i = p->next++;
      const size_t off = (size_t)i * p->part, len = std::min(p->part, p->bytes - off);
      const char* s = p->src;
      char* d = p->dst;
      lk.unlock();
      memcpy(d + off, s + off, len);
      lk.lock();
      if (--p->pending == 0) p->cv_done.notify_all();
    }
  }
}

CopyPool* copy_pool() {
  std::lock_guard<std::mutex> g(g_copy_pool_mu);
  const long pid = (long)getpid();
  if (g_copy_pool && g_copy_pool_pid == pid) return g_copy_pool->n_threads > 0 ? g_copy_pool : nullptr;
  int n = 0;
  if (const char* e = getenv("TSDF_B200_COPY_THREADS")) n = atoi(e);
  const int hw = (int)std::thread::hardware_concurrency();
  if (hw > 0 && n > hw - 1) n = hw - 1;
  if (n > 15) n = 15;
  CopyPool* p = new (std::nothrow) CopyPool();   // (a pool inherited through fork() has no threads: dropped)
  if (!p) return nullptr;
  p->n_threads = n > 0 ? n : 0;
  try {
    for (int i = 0; i < p->n_threads; ++i) std::thread(copy_worker, p).detach();
  } catch (...) {
    p->n_threads = 0;
  }
  g_copy_pool = p;
  g_copy_pool_pid = pid;
  return p->n_threads > 0 ? p : nullptr;
}
}  // namespace

void host_copy(void* dst, const void* src, size_t bytes) {
  CopyPool* p = bytes >= (256u << 10) ? copy_pool() : nullptr;
  if (!p || !p->busy.try_lock()) {
    memcpy(dst, src, bytes);
    return;
  }
  {
    std::unique_lock<std::mutex> lk(p->m);
    p->src = static_cast<const char*>(src);
    p->dst = static_cast<char*>(dst);
    p->bytes = bytes;
    p->parts = p->n_threads + 1;
    p->part = ((bytes + p->parts - 1) / p->parts + 4095) & ~(size_t)4095;
    p->parts = (int)((bytes + p->part - 1) / p->part);
    p->next = 0;
    p->pending = p->parts;
    ++p->gen;
    p->cv_work.notify_all();
    while (p->next < p->parts) {   // the caller takes parts too
      const int i = p->next++;
      const size_t off = (size_t)i * p->part, len = std::min(p->part, bytes - off);
      lk.unlock();
      memcpy(p->dst + off, p->src + off, len);
      lk.lock();
      --p->pending;
    }
    p->cv_done.wait(lk, [&] { return p->pending == 0; });
  }
  p->busy.unlock();
}

// Moves `bytes` from src (any kind) to where the prep kernel can read it; returns the device
// pointer to read from.  Pageable data goes through the slot's pinned buffer.
int frame_to_device(tsdf_vol* v, FrameSlot& s, const void* src, size_t bytes, size_t raw_off,
                    const void** d_out, bool into_slot) {
  PtrKind k = ptr_kind(src);
  char* dst = reinterpret_cast<char*>(s.d_raw) + raw_off;
  if (k == PTR_DEVICE) {
    if (!into_slot) { *d_out = src; return TSDF_OK; }
    // peers read the frame from this slot's raw buffer: device inputs are copied into it
    CU_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, v->s_copy));
    *d_out = dst;
    return TSDF_OK;
  }
  if (k == PTR_PAGEABLE) {
    if (!s.h_pinned) CU_CHECK(cudaHostAlloc(&s.h_pinned, frame_pixels(v->cfg) * 7 + 64, cudaHostAllocDefault));
    char* stage = reinterpret_cast<char*>(s.h_pinned) + raw_off;
    host_copy(stage, src, bytes);
    src = stage;
  }
  CU_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, v->s_copy));
  *d_out = dst;
  return TSDF_OK;
}

int stage_frame(tsdf_vol* v, int slot_index, const void* depth, int depth_fmt, float depth_scale,
                int sub_c, int sub_r, const uint8_t* rgb, bool want_color, FrameSlot** out_slot,
                const float* K, bool clear_tail) {
  FrameSlot& s = v->slot[slot_index];
  int rc = ensure_slot(v, s);
  if (rc) return rc;
  if (s.in_flight) {  // the kernel that last read this slot must be done before we overwrite it
    CU_CHECK(cudaEventSynchronize(s.consumed));
    s.in_flight = false;
  }
  const size_t px = frame_pixels(v->cfg);
  const size_t dbytes = px * ((depth_fmt == TSDF_DEPTH_U16 || depth_fmt == TSDF_DEPTH_RANGE_U16) ? 2 : 4);
  const float* k = K ? K : v->cfg.K;
  const float intr[4] = {k[0], k[2], k[4], k[5]};
  const void* d_depth = nullptr;
  const void* d_rgb = nullptr;
  rc = frame_to_device(v, s, depth, dbytes, 0, &d_depth, false);
  if (rc) return rc;
  if (want_color && rgb) {
    rc = frame_to_device(v, s, rgb, px * 3, ((px * 4 + 63) / 64) * 64, &d_rgb, false);
    if (rc) return rc;
  }
  rc = launch_frame_prep(d_depth, depth_fmt, depth_scale, sub_c, sub_r,
                         reinterpret_cast<const uint8_t*>(d_rgb), nullptr, v->cfg.im_h, v->cfg.im_w,
                         v->cfg.max_depth, want_color, intr, s.d_frame, v->s_copy, clear_tail);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(s.uploaded, v->s_copy));
  *out_slot = &s;
  return TSDF_OK;
}

int compose_cam2base(const tsdf_vol* v, const float* T_wc, float* c2b) {
  if (v->identity_base) memcpy(c2b, T_wc, 16 * sizeof(float));  // inv(I)*T == T exactly
  else tsdf_multiply_matrix(v->base2world_inv, T_wc, c2b);       // src/tsdf.cu:142
  return TSDF_OK;
}

int make_frame_params(const tsdf_vol* v, const float* K, const float* T_wc, float obs_weight,
                      const void* d_frame, FrameP* out) {
  float c2b[16];
  compose_cam2base(v, T_wc, c2b);
  memcpy(out->c, c2b, 12 * sizeof(float));
  const float* k = K ? K : v->cfg.K;
  out->fx = k[0]; out->cx = k[2]; out->fy = k[4]; out->cy = k[5];
  out->obs_weight = obs_weight;
  out->pad_ = 0;
  out->skip = v->comm_skip;
  out->h_tmap = nullptr;
  for (const FrameSlot& fs : v->slot) if (fs.d_frame == d_frame && fs.tmap_ok) out->h_tmap = fs.tmap;
  for (const FrameSlot& fs : v->bslot) if (fs.d_frame == d_frame && fs.tmap_ok) out->h_tmap = fs.tmap;
  out->mb_x0 = 0; out->mb_y0 = 0; out->mb_x1 = v->cfg.im_w - 1; out->mb_y1 = v->cfg.im_h - 1;
  out->img = d_frame;
  out->dmax = frame_dmax_ptr(d_frame, frame_pixels(v->cfg));
  return TSDF_OK;
}

// Integer-weight bookkeeping (see blend_intw in integrate.cu).  Must be called once per
// launch, before it: returns whether the launch may use the integer-weight colour path.
bool account_weights(tsdf_vol* v, const FrameP* fp, int nf) {
  bool all_one = true;
  for (int i = 0; i < nf; ++i) all_one = all_one && fp[i].obs_weight == 1.0f;
  if (!all_one) v->intw = false;
  v->w_bound += nf;  // only meaningful while intw holds
  return v->intw && v->w_bound <= 4094;
}

// rcp_rn / div_rn of the hot loop are exact for operands in [2^-60, 2^60]; the weights this bound
// admits stay far inside it.  Shared by every integrate entry point (single, multi, bcast, batch).
int check_obs_weight(float obs_weight) {
  if (!(obs_weight >= 1e-6f && obs_weight <= 1e6f)) {   // also rejects NaN
    tsdf_set_error("obs_weight must be in [1e-6, 1e6] (got %g)", (double)obs_weight);
    return TSDF_E_INVALID;
  }
  return TSDF_OK;
}

// Launches the integrate kernel for nf prepared frames and records stats/events.
int submit_frames(tsdf_vol* v, FrameSlot** slots, const FrameP* fp, int nf, bool color, bool accumulate) {
  if (v) ++v->epoch;  // voxel arrays change: invalidates the cached extraction mark pass
  for (int i = 0; i < nf; ++i) CU_CHECK(cudaStreamWaitEvent(v->s_compute, slots[i]->uploaded, 0));
  cudaEvent_t k0 = v->ev_k0, k1 = v->ev_k1;
  if (v->ktiming) {  // a private event pair per launch
    if (v->kev_used + 2 > v->kev.size()) {
      for (int i = 0; i < 2; ++i) {
        cudaEvent_t e;
        CU_CHECK(cudaEventCreate(&e));
        v->kev.push_back(e);
      }
    }
    k0 = v->kev[v->kev_used];
    k1 = v->kev[v->kev_used + 1];
    v->kev_used += 2;
  }
  v->wmode_ok = account_weights(v, fp, nf);
  const int wi = (int)(v->lseq % kWrittenRing);
  // this launch adds into counter wi: the publish of the launch that last used it must have re-armed it
  if (v->lseq >= (uint64_t)kWrittenRing) CU_CHECK(cudaStreamWaitEvent(v->s_compute, v->ev_pub_done[wi], 0));
  // a later submission of one tsdf_integrate_multi_async call keeps the call's first start event, so that
  // tsdf_get_stats' last_kernel_ms spans the same k frames as its voxels_written / bytes_algorithmic
  if (v->ktiming || !accumulate) CU_CHECK(cudaEventRecord(k0, v->s_compute));
  trace_mark(v, 2, v->s_compute);
  int rc = launch_integrate(v, fp, nf, color, v->s_compute);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(k1, v->s_compute));
  trace_mark(v, 3, v->s_compute);
  // publish on the aux stream, behind this launch: N_written -> pinned mirror (the step's 8-byte
  // device->host result), counter re-armed; single-frame submissions also zero the consumed slot's
  // max-depth tiles there, so the next prep into it runs without a memset (handles under A/B tooling
  // keep the frame intact: tsdf_bench_kernel re-launches on it).  The compute stream carries
  // integrate launches only.
  CU_CHECK(cudaEventRecord(v->ev_int_done[wi], v->s_compute));
  CU_CHECK(cudaStreamWaitEvent(v->s_aux, v->ev_int_done[wi], 0));
  rc = launch_publish(v, accumulate, v->s_aux, (nf == 1 && !v->dev_hold) ? slots[0]->d_frame : nullptr);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(v->ev_pub_done[wi], v->s_aux));
  for (int i = 0; i < nf; ++i) {
    CU_CHECK(cudaEventRecord(slots[i]->consumed, v->s_aux));
    slots[i]->in_flight = true;
  }
  ++v->lseq;
  memcpy(v->last_frames, fp, sizeof(FrameP) * nf);
  v->last_nf = nf;
  v->last_color = color;
  v->last_frames_in_launch = nf;
  v->frames += nf;
  return TSDF_OK;
}

extern "C" int tsdf_integrate_raw_async(tsdf_vol* v, const void* depth, int depth_fmt,
                                        float depth_scale, int sub_c, int sub_r, const uint8_t* rgb,
                                        const float* K, const float* T_wc, float obs_weight) {
  REQUIRE(v && depth && T_wc, "null argument");
  REQUIRE(depth_fmt >= TSDF_DEPTH_F32 && depth_fmt <= TSDF_DEPTH_RANGE_U16, "bad depth format");
  if (depth_fmt >= TSDF_DEPTH_RANGE_F32) {
    const float* k = K ? K : v->cfg.K;
    REQUIRE(k[0] != 0.0f && k[4] != 0.0f, "range formats need non-zero fx, fy");
  }
  int rcw = check_obs_weight(obs_weight);
  if (rcw) return rcw;
  DeviceGuard g(v->cfg.device);
  const bool color = v->d_color != nullptr && rgb != nullptr;
  FrameSlot* s = nullptr;
  trace_mark(v, 0, v->s_copy);
  CU_CHECK(cudaEventRecord(v->ev_u0, v->s_copy));
  int rc = stage_frame(v, (int)(v->seq % 3), depth, depth_fmt, depth_scale, sub_c < 1 ? 1 : sub_c,
                       sub_r < 1 ? 1 : sub_r, rgb, color, &s, K, /*clear_tail=*/v->dev_hold);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(v->ev_u1, v->s_copy));
  trace_mark(v, 1, v->s_copy);
  ++v->seq;
  FrameP fp;
  make_frame_params(v, K, T_wc, obs_weight, s->d_frame, &fp);
  return submit_frames(v, &s, &fp, 1, color);
}

extern "C" int tsdf_integrate_async(tsdf_vol* v, const float* depth, const uint8_t* rgb,
                                    const float* K, const float* T_wc, float obs_weight) {
  return tsdf_integrate_raw_async(v, depth, TSDF_DEPTH_F32, 1.0f, 1, 1, rgb, K, T_wc, obs_weight);
}

extern "C" int tsdf_integrate(tsdf_vol* v, const float* depth, const uint8_t* rgb, const float* K,
                              const float* T_wc, float obs_weight) {
  int rc = tsdf_integrate_async(v, depth, rgb, K, T_wc, obs_weight);
  return rc ? rc : tsdf_sync(v);
}

extern "C" int tsdf_integrate_multi_async(tsdf_vol* v, int k, const float* depth, const uint8_t* rgb,
                                          const float* K, const float* T_wc, float obs_weight) {
  REQUIRE(v && depth && T_wc && k >= 1, "null argument or k < 1");
  int rcw = check_obs_weight(obs_weight);
  if (rcw) return rcw;
  DeviceGuard g(v->cfg.device);
  const bool color = v->d_color != nullptr && rgb != nullptr;
  const size_t px = frame_pixels(v->cfg);
  // Frames are staged and submitted two at a time, in frame order: two consecutive dense frames share
  // one pass over the volume (k_integrate2, launch_integrate), and the host stages pair i + 1 (copy
  // into the pinned ring, H2D, prep) while the GPU integrates pair i.  (Round 1 staged groups of up to
  // 8 frames on the caller's thread before the group's first launch: 5.5 ms per frame at k = 8.)
  int done = 0;
  while (done < k) {
    const int nf = (k - done >= 2) ? 2 : 1;
    FrameSlot* slots[2];
    FrameP fp[2];
    for (int i = 0; i < nf; ++i) {
      const int f = done + i;
      const int si = 3 + (int)(v->mseq++ % (2 * kMaxFusedFrames));
      int rc = stage_frame(v, si, depth + (size_t)f * px, TSDF_DEPTH_F32, 1.0f, 1, 1,
                           rgb ? rgb + (size_t)f * px * 3 : nullptr, color, &slots[i]);
      if (rc) return rc;
      make_frame_params(v, K, T_wc + 16 * f, obs_weight, slots[i]->d_frame, &fp[i]);
    }
    int rc = submit_frames(v, slots, fp, nf, color, /*accumulate=*/done > 0);
    if (rc) return rc;
    done += nf;
  }
  v->last_frames_in_launch = k;
  return TSDF_OK;
}

extern "C" int tsdf_sync(tsdf_vol* v) {
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  CU_CHECK(cudaStreamSynchronize(v->s_copy));
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  CU_CHECK(cudaStreamSynchronize(v->s_aux));
  for (FrameSlot& s : v->slot) s.in_flight = false;
  for (FrameSlot& s : v->bslot) s.in_flight = false;
  v->last_written = (int64_t)*v->h_written;
  if (v->comm_err && *v->comm_err) {
    tsdf_set_error("frame distribution: a peer-pull wait timed out (code %d): a rank is gone or stalled; the "
                   "frames behind it were NOT integrated on this rank", *v->comm_err);
    return TSDF_E_NCCL;
  }
  return TSDF_OK;
}

extern "C" int tsdf_get_stats(tsdf_vol* v, tsdf_stats* s) {
  REQUIRE(v && s, "null argument");
  int rc = tsdf_sync(v);
  if (rc) return rc;
  memset(s, 0, sizeof(*s));
  s->voxels_visited = v->n_local * v->last_frames_in_launch;
  s->voxels_written = v->last_written;
  s->frames = v->frames;
  if (v->last_nf > 0) {
    if (v->ktiming && v->kev_used >= 2) cudaEventElapsedTime(&s->last_kernel_ms, v->kev[v->kev_used - 2], v->kev[v->kev_used - 1]);
    else if (cudaEventElapsedTime(&s->last_kernel_ms, v->ev_k0, v->ev_k1) != cudaSuccess) { cudaGetLastError(); s->last_kernel_ms = 0; }
    if (cudaEventElapsedTime(&s->last_upload_ms, v->ev_u0, v->ev_u1) != cudaSuccess) {
      cudaGetLastError();
      s->last_upload_ms = 0;
    }
  }
  const int64_t bvox = v->last_color ? 24 : 16;
  const int64_t bpix = v->last_color ? 8 : 4;
  s->bytes_algorithmic = bvox * v->last_written +
                         bpix * (int64_t)frame_pixels(v->cfg) * v->last_frames_in_launch;
  return TSDF_OK;
}

#ifdef TSDF_B200_DEV
// Work counters of the sparse kernel: {blocks, blocks skipped by the hull test, warp row segments,
// segments culled by the warp test, runs projected, runs gathered, runs with an update, -}.
// enable allocates / zeroes them; out (8 values) may be null.
extern "C" int tsdf_debug_counters(tsdf_vol* v, int enable, uint64_t* out) {
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  if (out) {
    if (v->d_dbg) CU_CHECK(cudaMemcpy(out, v->d_dbg, 64, cudaMemcpyDeviceToHost));
    else memset(out, 0, 64);
  }
  if (enable) {
    if (!v->d_dbg) CU_CHECK(cudaMalloc(&v->d_dbg, 64));
    CU_CHECK(cudaMemset(v->d_dbg, 0, 64));
  } else if (v->d_dbg) {
    cudaFree(v->d_dbg);
    v->d_dbg = nullptr;
  }
  return TSDF_OK;
}

extern "C" int tsdf_set_variant(tsdf_vol* v, int variant) {
  REQUIRE(v, "null handle");
  v->variant = variant;
  v->dev_hold = true;   // A/B tooling re-launches on the last prepared frame: keep its tail intact
  return TSDF_OK;
}

extern "C" int tsdf_bench_kernel(tsdf_vol* v, int iters, float* avg_ms) {
  if (v) ++v->epoch;  // voxel arrays change: invalidates the cached extraction mark pass
  REQUIRE(v && avg_ms && iters > 0, "bad argument");
  REQUIRE(v->last_nf > 0, "no frame integrated yet");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  CU_CHECK(cudaEventRecord(v->ev_k0, v->s_compute));
  for (int i = 0; i < iters; ++i) {
    v->wmode_ok = account_weights(v, v->last_frames, v->last_nf);
    rc = launch_integrate(v, v->last_frames, v->last_nf, v->last_color, v->s_compute);
    if (rc) return rc;
  }
  CU_CHECK(cudaEventRecord(v->ev_k1, v->s_compute));
  rc = launch_publish(v, false, v->s_compute);
  if (rc) return rc;
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  ++v->lseq;
  float ms = 0;
  CU_CHECK(cudaEventElapsedTime(&ms, v->ev_k0, v->ev_k1));
  *avg_ms = ms / iters;
  v->frames += (int64_t)iters * v->last_nf;
  *v->h_written /= (unsigned long long)iters;  // per-launch average (tsdf_sync re-reads the mirror)
  v->last_written = (int64_t)*v->h_written;
  return TSDF_OK;
}
#endif  // TSDF_B200_DEV

// ---- readback (K4, K5, K7) -------------------------------------------------------------------
// Slab <-> host arrays in the semantics' reference order.  TSDFCU slab [zl][y][x] is one
// contiguous range of the global array; FUSIONPY slab [x][y][zl] is a strided window of
// the global [x][y][z] array (2-D copy with pitch Dz).
static int copy_slab(tsdf_vol* v, float* host, float* dev, int global_layout, bool to_host) {
  if (!host || !dev) return TSDF_OK;
  const tsdf_config& c = v->cfg;
  cudaMemcpyKind kind = to_host ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  if (!global_layout || (v->dzl == c.dim_z)) {
    if (to_host) CU_CHECK(cudaMemcpyAsync(host, dev, (size_t)v->n_local * 4, kind, v->s_compute));
    else         CU_CHECK(cudaMemcpyAsync(dev, host, (size_t)v->n_local * 4, kind, v->s_compute));
    return TSDF_OK;
  }
  if (!v->zfast) {
    float* h = host + (size_t)c.z_begin * c.dim_y * c.dim_x;
    if (to_host) CU_CHECK(cudaMemcpyAsync(h, dev, (size_t)v->n_local * 4, kind, v->s_compute));
    else         CU_CHECK(cudaMemcpyAsync(dev, h, (size_t)v->n_local * 4, kind, v->s_compute));
    return TSDF_OK;
  }
  float* h = host + c.z_begin;
  const size_t rows = (size_t)c.dim_x * c.dim_y;
  if (to_host)
    CU_CHECK(cudaMemcpy2DAsync(h, (size_t)c.dim_z * 4, dev, (size_t)v->dzl * 4, (size_t)v->dzl * 4, rows, kind, v->s_compute));
  else
    CU_CHECK(cudaMemcpy2DAsync(dev, (size_t)v->dzl * 4, h, (size_t)c.dim_z * 4, (size_t)v->dzl * 4, rows, kind, v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_download(tsdf_vol* v, float* tsdf, float* weight, float* color, int global_layout) {
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  REQUIRE(!color || v->d_color, "volume has no colour array");
  if ((rc = copy_slab(v, tsdf, v->d_tsdf, global_layout, true))) return rc;
  if ((rc = copy_slab(v, weight, v->d_weight, global_layout, true))) return rc;
  if ((rc = copy_slab(v, color, v->d_color, global_layout, true))) return rc;
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_download_range(tsdf_vol* v, int64_t first_voxel, int64_t n_voxels, float* tsdf,
                                   float* weight, float* color) {
  REQUIRE(v, "null handle");
  REQUIRE(first_voxel >= 0 && n_voxels >= 0 && first_voxel <= v->n_local && n_voxels <= v->n_local - first_voxel,
          "voxel range outside the slab");
  REQUIRE(!color || v->d_color, "volume has no colour array");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  const size_t bytes = (size_t)n_voxels * sizeof(float);
  float* host[3] = {tsdf, weight, color};
  const float* dev[3] = {v->d_tsdf, v->d_weight, v->d_color};
  for (int i = 0; i < 3; ++i)
    if (host[i] && bytes) CU_CHECK(cudaMemcpyAsync(host[i], dev[i] + first_voxel, bytes, cudaMemcpyDeviceToHost, v->s_compute));
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_upload(tsdf_vol* v, const float* tsdf, const float* weight, const float* color,
                           int global_layout) {
  if (v) ++v->epoch;  // voxel arrays change: invalidates the cached extraction mark pass
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  REQUIRE(!color || v->d_color, "volume has no colour array");
  if (weight) v->intw = false;  // arbitrary weights: the integer-weight fast path is off until reset
  if ((rc = copy_slab(v, const_cast<float*>(tsdf), v->d_tsdf, global_layout, false))) return rc;
  if ((rc = copy_slab(v, const_cast<float*>(weight), v->d_weight, global_layout, false))) return rc;
  if ((rc = copy_slab(v, const_cast<float*>(color), v->d_color, global_layout, false))) return rc;
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_extract_points_device(tsdf_vol* v, int predicate, float tsdf_thresh,
                                          float weight_thresh, float* d_xyz, uint8_t* d_rgb,
                                          int64_t* d_idx, int64_t cap, int64_t* count) {
  REQUIRE(v && count, "null argument");
  REQUIRE(predicate == TSDF_PRED_NONZERO || predicate == TSDF_PRED_ABS_LT, "bad predicate");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, d_xyz, d_rgb, d_idx, cap, count);
  if (rc) return rc;
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_extract_points(tsdf_vol* v, int predicate, float tsdf_thresh, float weight_thresh,
                                   float* xyz, uint8_t* rgb, int64_t* idx, int64_t cap, int64_t* count) {
  REQUIRE(v && count, "null argument");
  REQUIRE(predicate == TSDF_PRED_NONZERO || predicate == TSDF_PRED_ABS_LT, "bad predicate");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  int64_t n = 0;
  rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &n);
  if (rc) return rc;
  *count = n;
  if (!xyz && !rgb && !idx) return TSDF_OK;
  if (cap < n) {
    tsdf_set_error("tsdf_extract_points: capacity %lld < %lld points", (long long)cap, (long long)n);
    return TSDF_E_CAPACITY;
  }
  if (n == 0) return TSDF_OK;
  // outputs live in the handle's grow-only device buffers; each array goes to the caller with its own
  // asynchronous copy (straight into pinned memory, or through the pinned staging area)
  OutBuffers& ob = v->out;
  if (xyz && (rc = ob.need_dev(0, (size_t)n * 12))) return rc;
  if (rgb && (rc = ob.need_dev(1, (size_t)n * 3))) return rc;
  if (idx && (rc = ob.need_dev(2, (size_t)n * 8))) return rc;
  float* d_xyz = xyz ? static_cast<float*>(ob.d[0]) : nullptr;
  uint8_t* d_rgb = rgb ? static_cast<uint8_t*>(ob.d[1]) : nullptr;
  int64_t* d_idx = idx ? static_cast<int64_t*>(ob.d[2]) : nullptr;
  rc = extract_device(v, predicate, tsdf_thresh, weight_thresh, d_xyz, d_rgb, d_idx, n, &n);
  if (rc) return rc;
  if (xyz && (rc = copy_to_host(&ob, xyz, d_xyz, (size_t)n * 12, v->s_compute))) return rc;
  if (rgb && (rc = copy_to_host(&ob, rgb, d_rgb, (size_t)n * 3, v->s_compute))) return rc;
  if (idx && (rc = copy_to_host(&ob, idx, d_idx, (size_t)n * 8, v->s_compute))) return rc;
  CU_CHECK(cudaStreamSynchronize(v->s_compute));
  return TSDF_OK;
}

// ---- mesh readback (plugin SaveMesh, src/TSDFfusion.py.in:48-53) -------------------------------
extern "C" int tsdf_extract_mesh_device(tsdf_vol* v, float* d_verts, float* d_norms, uint8_t* d_colors,
                                        int64_t vcap, int32_t* d_faces, int64_t fcap, int64_t* n_verts,
                                        int64_t* n_faces) {
  REQUIRE(v && n_verts && n_faces, "null argument");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  return mesh_device(v, d_verts, d_norms, d_colors, vcap, d_faces, fcap, n_verts, n_faces);
}

extern "C" int tsdf_extract_mesh(tsdf_vol* v, float* verts, float* norms, uint8_t* colors, int64_t vcap,
                                 int32_t* faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces) {
  REQUIRE(v && n_verts && n_faces, "null argument");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  int64_t nv = 0, nf = 0;
  rc = mesh_device(v, nullptr, nullptr, nullptr, 0, nullptr, 0, &nv, &nf);
  if (rc) return rc;
  *n_verts = nv;
  *n_faces = nf;
  if (!verts && !norms && !colors && !faces) return TSDF_OK;
  if ((verts || norms || colors) && vcap < nv) {
    tsdf_set_error("tsdf_extract_mesh: vertex capacity %lld < %lld", (long long)vcap, (long long)nv);
    return TSDF_E_CAPACITY;
  }
  if (faces && fcap < nf) {
    tsdf_set_error("tsdf_extract_mesh: face capacity %lld < %lld", (long long)fcap, (long long)nf);
    return TSDF_E_CAPACITY;
  }
  OutBuffers& ob = v->out;
  if (verts && nv && (rc = ob.need_dev(0, (size_t)nv * 12))) return rc;
  if (norms && nv && (rc = ob.need_dev(1, (size_t)nv * 12))) return rc;
  if (colors && nv && (rc = ob.need_dev(2, (size_t)nv * 3))) return rc;
  if (faces && nf && (rc = ob.need_dev(3, (size_t)nf * 12))) return rc;
  float* d_v = (verts && nv) ? static_cast<float*>(ob.d[0]) : nullptr;
  float* d_n = (norms && nv) ? static_cast<float*>(ob.d[1]) : nullptr;
  uint8_t* d_c = (colors && nv) ? static_cast<uint8_t*>(ob.d[2]) : nullptr;
  int32_t* d_f = (faces && nf) ? static_cast<int32_t*>(ob.d[3]) : nullptr;
  if (d_v || d_n || d_c || d_f) {
    rc = mesh_device(v, d_v, d_n, d_c, nv, d_f, nf, &nv, &nf);
    if (rc) return rc;
    if (d_v && (rc = copy_to_host(&ob, verts, d_v, (size_t)nv * 12, v->s_compute))) return rc;
    if (d_n && (rc = copy_to_host(&ob, norms, d_n, (size_t)nv * 12, v->s_compute))) return rc;
    if (d_c && (rc = copy_to_host(&ob, colors, d_c, (size_t)nv * 3, v->s_compute))) return rc;
    if (d_f && (rc = copy_to_host(&ob, faces, d_f, (size_t)nf * 12, v->s_compute))) return rc;
    CU_CHECK(cudaStreamSynchronize(v->s_compute));
  }
  return TSDF_OK;
}

// fusion.meshwrite: ascii PLY, "%f %f %f %f %f %f %d %d %d" per vertex, "3 %d %d %d" per face.
extern "C" int tsdf_save_mesh_ply(tsdf_vol* v, const char* path) {
  REQUIRE(v && path, "null argument");
  REQUIRE(v->dzl == v->cfg.dim_z, "tsdf_save_mesh_ply needs a whole-grid handle");
  int64_t nv = 0, nf = 0;
  int rc = tsdf_extract_mesh(v, nullptr, nullptr, nullptr, 0, nullptr, 0, &nv, &nf);
  if (rc) return rc;
  std::vector<float> verts((size_t)(nv ? nv : 1) * 3), norms((size_t)(nv ? nv : 1) * 3);
  std::vector<uint8_t> cols((size_t)(nv ? nv : 1) * 3);
  std::vector<int32_t> faces((size_t)(nf ? nf : 1) * 3);
  if (nv) {
    rc = tsdf_extract_mesh(v, verts.data(), norms.data(), cols.data(), nv, faces.data(), nf, &nv, &nf);
    if (rc) return rc;
  }
  return write_mesh_ply(path, verts.data(), norms.data(), cols.data(), nv, faces.data(), nf);
}

int write_mesh_ply(const char* path, const float* verts, const float* norms, const uint8_t* cols, int64_t nv,
                   const int32_t* faces, int64_t nf) {
  FILE* fp = fopen(path, "w");
  if (!fp) { tsdf_set_error("tsdf_save_mesh_ply: cannot open %s", path); return TSDF_E_IO; }
  fprintf(fp, "ply\nformat ascii 1.0\nelement vertex %d\n", (int)nv);
  fprintf(fp, "property float x\nproperty float y\nproperty float z\n");
  fprintf(fp, "property float nx\nproperty float ny\nproperty float nz\n");
  fprintf(fp, "property uchar red\nproperty uchar green\nproperty uchar blue\n");
  fprintf(fp, "element face %d\nproperty list uchar int vertex_index\nend_header\n", (int)nf);
  for (int64_t i = 0; i < nv; ++i)
    fprintf(fp, "%f %f %f %f %f %f %d %d %d\n", verts[3 * i], verts[3 * i + 1], verts[3 * i + 2], norms[3 * i],
            norms[3 * i + 1], norms[3 * i + 2], cols[3 * i], cols[3 * i + 1], cols[3 * i + 2]);
  for (int64_t i = 0; i < nf; ++i) fprintf(fp, "3 %d %d %d\n", faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
  const bool bad = ferror(fp) != 0;
  if (fclose(fp) != 0 || bad) { tsdf_set_error("tsdf_save_mesh_ply: write failed"); return TSDF_E_IO; }
  return TSDF_OK;
}

// PLY: header text of src/tsdf.cu:186-192, then float x,y,z per kept voxel.
extern "C" int tsdf_save_ply(tsdf_vol* v, const char* path, int predicate, float tsdf_thresh,
                             float weight_thresh) {
  REQUIRE(v && path, "null argument");
  REQUIRE(v->dzl == v->cfg.dim_z, "tsdf_save_ply needs a whole-grid handle");
  int64_t n = 0;
  int rc = tsdf_extract_points(v, predicate, tsdf_thresh, weight_thresh, nullptr, nullptr, nullptr, 0, &n);
  if (rc) return rc;
  std::vector<float> xyz((size_t)(n ? n : 1) * 3);
  if (n) {
    rc = tsdf_extract_points(v, predicate, tsdf_thresh, weight_thresh, xyz.data(), nullptr, nullptr, n, &n);
    if (rc) return rc;
  }
  return write_points_ply(path, xyz.data(), n);
}

int write_points_ply(const char* path, const float* xyz, int64_t n) {
  FILE* fp = fopen(path, "w");
  if (!fp) { tsdf_set_error("tsdf_save_ply: cannot open %s", path); return TSDF_E_IO; }
  fprintf(fp, "ply\n");
  fprintf(fp, "format binary_little_endian 1.0\n");
  fprintf(fp, "element vertex %d\n", (int)n);
  fprintf(fp, "property float x\n");
  fprintf(fp, "property float y\n");
  fprintf(fp, "property float z\n");
  fprintf(fp, "end_header\n");
  size_t wr = fwrite(xyz, sizeof(float), (size_t)n * 3, fp);
  fclose(fp);
  if (wr != (size_t)n * 3) { tsdf_set_error("tsdf_save_ply: short write"); return TSDF_E_IO; }
  return TSDF_OK;
}

// BIN: 8-float header (dims as floats, origin, voxel_size, trunc) + tsdf (src/tsdf.cu:116-132).
extern "C" int tsdf_save_bin(tsdf_vol* v, const char* path) {
  REQUIRE(v && path, "null argument");
  REQUIRE(v->dzl == v->cfg.dim_z, "tsdf_save_bin needs a whole-grid handle");
  std::vector<float> t((size_t)v->n_local);
  int rc = tsdf_download(v, t.data(), nullptr, nullptr, 0);
  if (rc) return rc;
  return write_bin(path, v->cfg, t.data(), t.size());
}

int write_bin(const char* path, const tsdf_config& c, const float* tsdf, size_t n) {
  FILE* fp = fopen(path, "wb");
  if (!fp) { tsdf_set_error("tsdf_save_bin: cannot open %s", path); return TSDF_E_IO; }
  float h[8] = {(float)c.dim_x, (float)c.dim_y, (float)c.dim_z, c.origin[0], c.origin[1], c.origin[2],
                c.voxel_size, c.trunc_margin};
  size_t wr = fwrite(h, sizeof(float), 8, fp);
  wr += fwrite(tsdf, sizeof(float), n, fp);
  fclose(fp);
  if (wr != 8 + n) { tsdf_set_error("tsdf_save_bin: short write"); return TSDF_E_IO; }
  return TSDF_OK;
}

extern "C" int tsdf_load_bin(tsdf_vol* v, const char* path) {
  if (v) ++v->epoch;  // voxel arrays change: invalidates the cached extraction mark pass
  REQUIRE(v && path, "null argument");
  REQUIRE(v->dzl == v->cfg.dim_z, "tsdf_load_bin needs a whole-grid handle");
  FILE* fp = fopen(path, "rb");
  if (!fp) { tsdf_set_error("tsdf_load_bin: cannot open %s", path); return TSDF_E_IO; }
  float h[8];
  const tsdf_config& c = v->cfg;
  if (fread(h, sizeof(float), 8, fp) != 8 || (int)h[0] != c.dim_x || (int)h[1] != c.dim_y ||
      (int)h[2] != c.dim_z) {
    fclose(fp);
    tsdf_set_error("tsdf_load_bin: header of %s does not match the volume dims", path);
    return TSDF_E_IO;
  }
  std::vector<float> t((size_t)v->n_local);
  size_t rd = fread(t.data(), sizeof(float), t.size(), fp);
  fclose(fp);
  if (rd != t.size()) { tsdf_set_error("tsdf_load_bin: short read"); return TSDF_E_IO; }
  return tsdf_upload(v, t.data(), nullptr, nullptr, 0);
}

// ---- per-instance origin on the device (K9) -----------------------------------------------------
extern "C" int tsdf_instance_origin(int device, const float* depth, const uint8_t* mask, int im_h,
                                    int im_w, const float K[9], float origin_out[3]) {
  REQUIRE(depth && K && origin_out && im_h > 0 && im_w > 0, "bad argument");
  int ndev = 0;
  int rc = tsdf_device_count(&ndev);
  if (rc) return rc;
  DeviceGuard g(device);
  const size_t px = (size_t)im_h * im_w;
  float* d_depth = nullptr; uint8_t* d_mask = nullptr; float* d_out = nullptr;
  const float* dd = depth; const uint8_t* dm = mask;
  auto cleanup = [&]() { cudaFree(d_depth); cudaFree(d_mask); cudaFree(d_out); };
  cudaError_t e = cudaMalloc(&d_out, 12);
  if (e == cudaSuccess && ptr_kind(depth) != PTR_DEVICE) {
    e = cudaMalloc(&d_depth, px * 4);
    if (e == cudaSuccess) e = cudaMemcpy(d_depth, depth, px * 4, cudaMemcpyHostToDevice);
    dd = d_depth;
  }
  if (e == cudaSuccess && mask && ptr_kind(mask) != PTR_DEVICE) {
    e = cudaMalloc(&d_mask, px);
    if (e == cudaSuccess) e = cudaMemcpy(d_mask, mask, px, cudaMemcpyHostToDevice);
    dm = d_mask;
  }
  if (e != cudaSuccess) { cleanup(); tsdf_set_error("tsdf_instance_origin: %s", cudaGetErrorString(e)); return TSDF_E_CUDA; }
  rc = launch_instance_origin(dd, dm, im_h, im_w, K[0], K[2], K[4], K[5], d_out, 0);
  if (!rc) {
    e = cudaMemcpy(origin_out, d_out, 12, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { tsdf_set_error("tsdf_instance_origin: %s", cudaGetErrorString(e)); rc = TSDF_E_CUDA; }
  }
  cleanup();
  return rc;
}

// ---- memory helpers -------------------------------------------------------------------------------
extern "C" int tsdf_host_alloc(void** p, size_t bytes) {
  REQUIRE(p, "null argument");
  CU_CHECK(cudaHostAlloc(p, bytes, cudaHostAllocDefault));
  return TSDF_OK;
}
extern "C" int tsdf_host_free(void* p) {
  if (p) CU_CHECK(cudaFreeHost(p));
  return TSDF_OK;
}
extern "C" int tsdf_device_alloc(int device, void** p, size_t bytes) {
  REQUIRE(p, "null argument");
  DeviceGuard g(device);
  CU_CHECK(cudaMalloc(p, bytes));
  return TSDF_OK;
}
extern "C" int tsdf_device_free(int device, void* p) {
  DeviceGuard g(device);
  if (p) CU_CHECK(cudaFree(p));
  return TSDF_OK;
}
extern "C" int tsdf_memcpy_h2d(int device, void* dst, const void* src, size_t bytes) {
  DeviceGuard g(device);
  CU_CHECK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return TSDF_OK;
}
extern "C" int tsdf_memcpy_d2h(int device, void* dst, const void* src, size_t bytes) {
  DeviceGuard g(device);
  CU_CHECK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return TSDF_OK;
}

// ---- pose ingest ------------------------------------------------------------
extern "C" int tsdf_read_bundle_poses(const char* path, float* Twc, int cap, int* n_cams) {
  REQUIRE(path && n_cams, "null argument");
  FILE* f = fopen(path, "r");
  if (!f) { tsdf_set_error("tsdf_read_bundle_poses: cannot open %s", path); return TSDF_E_IO; }
  auto fail = [&](const char* what) {
    fclose(f);
    tsdf_set_error("tsdf_read_bundle_poses: %s in %s", what, path);
    return TSDF_E_IO;
  };
  char line[1024];
  long m = 0, n = 0;
  if (!fgets(line, sizeof line, f) || sscanf(line, "%ld %ld", &m, &n) < 1 || m < 0) return fail("bad header");
  *n_cams = (int)m;
  if (!Twc || cap == 0) { fclose(f); return TSDF_OK; }
  if (cap < m) {
    fclose(f);
    tsdf_set_error("tsdf_read_bundle_poses: capacity %d < %ld cameras", cap, m);
    return TSDF_E_CAPACITY;
  }
  for (long i = 0; i < m; ++i) {
    float R[9], t[3];
    if (!fgets(line, sizeof line, f)) return fail("truncated file");          // intrinsics line
    for (int j = 0; j < 3; ++j)                                                 // rows of Rcw
      if (!fgets(line, sizeof line, f) || sscanf(line, "%f %f %f", &R[3 * j], &R[3 * j + 1], &R[3 * j + 2]) != 3)
        return fail("bad rotation row");
    if (!fgets(line, sizeof line, f) || sscanf(line, "%f %f %f", &t[0], &t[1], &t[2]) != 3)
      return fail("bad translation row");
    float* T = Twc + 16 * i;
    // Rwc = Rcw^T, Ow = -Rwc*tcw (ORB-SLAM2 KeyFrame::SetPose), fp32, left to right
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) T[4 * r + c] = R[3 * c + r];
      T[4 * r + 3] = -((R[0 + r] * t[0] + R[3 + r] * t[1]) + R[6 + r] * t[2]);
    }
    T[12] = T[13] = T[14] = 0.0f;
    T[15] = 1.0f;
  }
  fclose(f);
  return TSDF_OK;
}

// CameraTrajectory.txt of the stereo / lidar run (src/ORB_SLAM2/System.cc:455-508, written by
// examples/stereo_kitti.cpp:113): one line per tracked frame, the 3x4 row-major [Rwc | twc] relative
// to the first keyframe — already camera-to-world, what Integrate takes; the fourth row is added.
extern "C" int tsdf_read_trajectory_kitti(const char* path, float* Twc, int cap, int* n_poses) {
  REQUIRE(path && n_poses, "null argument");
  FILE* f = fopen(path, "r");
  if (!f) { tsdf_set_error("tsdf_read_trajectory_kitti: cannot open %s", path); return TSDF_E_IO; }
  char line[2048];
  int n = 0, rc = TSDF_OK;
  while (fgets(line, sizeof line, f)) {
    float v[12];
    const int got = sscanf(line, "%f %f %f %f %f %f %f %f %f %f %f %f", &v[0], &v[1], &v[2], &v[3], &v[4], &v[5],
                           &v[6], &v[7], &v[8], &v[9], &v[10], &v[11]);
    if (got < 0) continue;  // empty line (sscanf: EOF before the first conversion)
    if (got != 12) {
      tsdf_set_error("tsdf_read_trajectory_kitti: line %d of %s has %d values, not 12", n + 1, path, got);
      rc = TSDF_E_IO;
      break;
    }
    if (Twc && n < cap) {
      float* T = Twc + 16 * (size_t)n;
      for (int k = 0; k < 12; ++k) T[k] = v[k];
      T[12] = T[13] = T[14] = 0.0f;
      T[15] = 1.0f;
    }
    ++n;
  }
  fclose(f);
  if (rc) return rc;
  *n_poses = n;
  if (Twc && cap > 0 && cap < n) {
    tsdf_set_error("tsdf_read_trajectory_kitti: capacity %d < %d poses", cap, n);
    return TSDF_E_CAPACITY;
  }
  return TSDF_OK;
}

// frame_id.txt (src/ORB_SLAM2/System.cc:1104-1121, examples/stereo_kitti.cpp:117): the frame number
// of every good keyframe, one per line, in the keyframe order of bundle.txt — which image of the
// sequence belongs to camera i of the bundle.
extern "C" int tsdf_read_frame_ids(const char* path, int64_t* ids, int cap, int* n_ids) {
  REQUIRE(path && n_ids, "null argument");
  FILE* f = fopen(path, "r");
  if (!f) { tsdf_set_error("tsdf_read_frame_ids: cannot open %s", path); return TSDF_E_IO; }
  char line[256];
  int n = 0, rc = TSDF_OK;
  while (fgets(line, sizeof line, f)) {
    long long v = 0;
    char extra = 0;
    const int got = sscanf(line, "%lld %c", &v, &extra);
    if (got < 0) continue;  // empty line
    if (got != 1 || v < 0) {
      tsdf_set_error("tsdf_read_frame_ids: bad line %d in %s", n + 1, path);
      rc = TSDF_E_IO;
      break;
    }
    if (ids && n < cap) ids[n] = (int64_t)v;
    ++n;
  }
  fclose(f);
  if (rc) return rc;
  *n_ids = n;
  if (ids && cap > 0 && cap < n) {
    tsdf_set_error("tsdf_read_frame_ids: capacity %d < %d ids", cap, n);
    return TSDF_E_CAPACITY;
  }
  return TSDF_OK;
}

extern "C" int tsdf_read_associations(const char* path, double* t_rgb, double* t_depth, char* rgb_names,
                                      char* depth_names, int name_stride, int cap, int* n_lines) {
  REQUIRE(path && n_lines, "null argument");
  REQUIRE(cap >= 0 && ((!rgb_names && !depth_names) || name_stride > 1), "bad capacity / name_stride");
  FILE* f = fopen(path, "r");
  if (!f) { tsdf_set_error("tsdf_read_associations: cannot open %s", path); return TSDF_E_IO; }
  const bool fill = cap > 0 && (t_rgb || t_depth || rgb_names || depth_names);
  char line[2048], a[1024], b[1024];
  int n = 0, rc = TSDF_OK;
  while (fgets(line, sizeof line, f)) {
    size_t len = strlen(line);
    while (len && (line[len - 1] == '\n' || line[len - 1] == '\r' || line[len - 1] == ' ' || line[len - 1] == '\t')) line[--len] = 0;
    if (!len) continue;
    double ta = 0, tb = 0;
    if (sscanf(line, "%lf %1023s %lf %1023s", &ta, a, &tb, b) != 4) {
      tsdf_set_error("tsdf_read_associations: line %d of %s is not '<t> <rgb> <t> <depth>'", n + 1, path);
      rc = TSDF_E_IO;
      break;
    }
    if (fill) {
      if (n >= cap) {
        tsdf_set_error("tsdf_read_associations: capacity %d is too small for %s", cap, path);
        rc = TSDF_E_CAPACITY;
        break;
      }
      if ((rgb_names && (int)strlen(a) >= name_stride) || (depth_names && (int)strlen(b) >= name_stride)) {
        tsdf_set_error("tsdf_read_associations: a file name on line %d does not fit name_stride %d", n + 1, name_stride);
        rc = TSDF_E_CAPACITY;
        break;
      }
      if (t_rgb) t_rgb[n] = ta;
      if (t_depth) t_depth[n] = tb;
      if (rgb_names) strcpy(rgb_names + (size_t)n * name_stride, a);
      if (depth_names) strcpy(depth_names + (size_t)n * name_stride, b);
    }
    ++n;
  }
  fclose(f);
  if (rc == TSDF_OK) *n_lines = n;
  return rc;
}

// ---- per-frame timeline (evidence for the multi-GPU step chain) --------------------
int trace_mark(tsdf_vol* v, int which, cudaStream_t st) {
  if (!v->tracing) return TSDF_OK;
  if (which == 0) {   // a new frame: four fresh slots
    v->tev_used = (v->tev_used + 3) / 4 * 4;
    while (v->tev.size() < v->tev_used + 4) {
      cudaEvent_t e;
      CU_CHECK(cudaEventCreate(&e));
      v->tev.push_back(e);
    }
    v->tev_used += 4;
  }
  if (v->tev_used < 4) return TSDF_OK;
  CU_CHECK(cudaEventRecord(v->tev[v->tev_used - 4 + which], st));
  return TSDF_OK;
}

extern "C" int tsdf_trace_enable(tsdf_vol* v, int enable) {
  REQUIRE(v, "null handle");
  int rc = tsdf_sync(v);
  if (rc) return rc;
  v->tracing = enable != 0;
  v->tev_used = 0;
  return TSDF_OK;
}

extern "C" int tsdf_trace_read(tsdf_vol* v, float* ms4, int cap_frames, int* n_frames) {
  REQUIRE(v && n_frames, "null argument");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  const int n = (int)(v->tev_used / 4);
  *n_frames = n;
  if (!ms4 || n == 0) return TSDF_OK;
  for (int f = 0; f < n && f < cap_frames; ++f)
    for (int k = 0; k < 4; ++k) {
      float ms = 0;
      if (cudaEventElapsedTime(&ms, v->tev[0], v->tev[4 * f + k]) != cudaSuccess) { cudaGetLastError(); ms = -1.0f; }
      ms4[4 * f + k] = ms;
    }
  return TSDF_OK;
}

// ---- measurement hooks ---------------------------------------------------------------------------
extern "C" int tsdf_timer_start(tsdf_vol* v) {
  REQUIRE(v, "null handle");
  DeviceGuard g(v->cfg.device);
  CU_CHECK(cudaEventRecord(v->ev_t0, v->s_compute));
  return TSDF_OK;
}

extern "C" int tsdf_timer_stop(tsdf_vol* v, float* ms) {
  REQUIRE(v && ms, "null argument");
  DeviceGuard g(v->cfg.device);
  CU_CHECK(cudaStreamSynchronize(v->s_copy));   // everything staged so far is enqueued/finished
  if (v->lseq > 0)   // ... and the last step's result has reached host memory
    CU_CHECK(cudaStreamWaitEvent(v->s_compute, v->ev_pub_done[(v->lseq - 1) % kWrittenRing], 0));
  CU_CHECK(cudaEventRecord(v->ev_t1, v->s_compute));
  CU_CHECK(cudaEventSynchronize(v->ev_t1));
  CU_CHECK(cudaEventElapsedTime(ms, v->ev_t0, v->ev_t1));
  return tsdf_sync(v);
}

extern "C" int tsdf_kernel_timing(tsdf_vol* v, int enable) {
  REQUIRE(v, "null handle");
  int rc = tsdf_sync(v);
  if (rc) return rc;
  v->ktiming = enable != 0;
  v->kev_used = 0;
  return TSDF_OK;
}

extern "C" int tsdf_kernel_timing_read(tsdf_vol* v, double* sum_ms, int64_t* launches, int reset) {
  REQUIRE(v && sum_ms && launches, "null argument");
  DeviceGuard g(v->cfg.device);
  int rc = tsdf_sync(v);
  if (rc) return rc;
  double s = 0;
  for (size_t i = 0; i + 1 < v->kev_used; i += 2) {
    float ms = 0;
    CU_CHECK(cudaEventElapsedTime(&ms, v->kev[i], v->kev[i + 1]));
    s += ms;
  }
  *sum_ms = s;
  *launches = (int64_t)(v->kev_used / 2);
  if (reset) v->kev_used = 0;
  return TSDF_OK;
}
