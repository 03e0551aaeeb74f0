// extract.cu — surface / point-cloud extraction (K5) as warp-ballot compaction.
//
// Replaces the two serial host scans of TSDF::SaveVoxelGrid2SurfacePointCloud (reference
// src/tsdf.cu:170-218).  Output order is ascending reference linear index, i.e. exactly the
// order the reference's loop emits (deterministic, no atomics on the output cursor):
//   pass 1  k_mark : 128-bit coalesced read of tsdf+weight (8 B/voxel, the only full sweep),
//                    predicate -> four warp ballots per 128-voxel chunk (16 B) + per-tile count
//   scan    k_scan_local + k_scan_carry : exclusive scan of the tile counts (4096 per block)
//   pass 2  k_emit : reads only the ballots; rank = popc of lower bits; writes xyz
//                    (origin + (float)i*vs, src/tsdf.cu:206-208), GLOBAL reference index and
//                    rgb (colour gathered for kept voxels only)
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kChunksPerWarp = 8;                       // 128-voxel chunks per warp
constexpr int kChunksPerTile = kWarps * kChunksPerWarp;  // 64
constexpr int64_t kTileVox = (int64_t)kChunksPerTile * 128;

__device__ __forceinline__ bool keep(int pred, float t, float w, float tt, float wt) {
  if (pred == TSDF_PRED_NONZERO) return fabsf(t) != 0.0f && w > wt;  // src/tsdf.cu:179,198
  return fabsf(t) < tt && w > wt;                                    // src/obselete/tsdf.cpp:16,33
}

// ballots[(chunk*4 + j)] bit t  <=>  voxel chunk*128 + 4*t + j is kept
__global__ void __launch_bounds__(kThreads)
k_mark(const float* __restrict__ tsdf, const float* __restrict__ weight, int64_t n, int pred,
       float tt, float wt, uint32_t* __restrict__ ballots, uint32_t* __restrict__ tile_count) {
  __shared__ uint32_t s_cnt[kWarps];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.x;
  uint32_t cnt = 0;
  const bool aligned = (n & 3) == 0;
#pragma unroll 2
  for (int c = 0; c < kChunksPerWarp; ++c) {
    const int64_t chunk = tile * kChunksPerTile + warp * kChunksPerWarp + c;
    const int64_t e0 = chunk * 128 + 4 * lane;
    float t[4] = {0, 0, 0, 0}, w[4] = {0, 0, 0, 0};
    if (aligned && e0 + 3 < n) {
      float4 a = __ldcs(reinterpret_cast<const float4*>(tsdf + e0));
      float4 b = __ldcs(reinterpret_cast<const float4*>(weight + e0));
      t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w;
      w[0] = b.x; w[1] = b.y; w[2] = b.z; w[3] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (e0 + j < n) { t[j] = tsdf[e0 + j]; w[j] = weight[e0 + j]; }
        else            { t[j] = 0.0f; w[j] = -1e30f; }  // never kept (needs w > wt; see below)
    }
    uint32_t b[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      bool k = (e0 + j < n) && keep(pred, t[j], w[j], tt, wt);
      b[j] = __ballot_sync(0xffffffffu, k);
      cnt += __popc(b[j]);
    }
    if (chunk * 128 < n && lane < 4)
      ballots[chunk * 4 + lane] = lane == 0 ? b[0] : lane == 1 ? b[1] : lane == 2 ? b[2] : b[3];
  }
  // cnt is warp-uniform (sum of ballots)
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kWarps; ++i) s += s_cnt[i];
    tile_count[tile] = s;
  }
}

// Exclusive scan of m uint32 counts into int64 offsets; offsets[m] = total.  Two launches:
//   k_scan_local  block b scans its 4096 counts (4 per thread, warp shuffles + one barrier), writes the
//                 block-local exclusive offsets and the block's total
//   k_scan_carry  block b sums the totals of the blocks before it (a few hundred at most) and adds
//                 them to its 4096 offsets; the last block also writes offsets[m]
// (the first form was ONE block walking the whole array: 0.30 ms for the 524288 mesh tiles of a
// 1024^3 volume, 2.4 us per 4096 entries, a fifth of the mesh sweep).  blockIdx.y selects one of up
// to two (count, offset, totals) triples of the same length (mesh.cu scans vertices and triangles).
constexpr int kScanBlock = 1024, kScanPer = 4, kScanTile = kScanBlock * kScanPer;
struct ScanP {
  const uint32_t* cnt[2];
  int64_t* off[2];
  int64_t* tot[2];  // ceil(m / 4096) block totals
  uint32_t* list[2];   // optional: indices of the non-zero counts, in no particular order ...
  uint32_t* nlist[2];  // ... and how many there are (zeroed by the caller)
  int64_t m;
};

__global__ void __launch_bounds__(kScanBlock)
k_scan_local(ScanP p) {
  __shared__ uint32_t s_warp[32];
  const uint32_t* __restrict__ cnt = p.cnt[blockIdx.y];
  int64_t* __restrict__ off = p.off[blockIdx.y];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t i0 = (int64_t)blockIdx.x * kScanTile + kScanPer * tid;
  uint32_t c[kScanPer];
#pragma unroll
  for (int j = 0; j < kScanPer; ++j) c[j] = (i0 + j < p.m) ? cnt[i0 + j] : 0u;
  // a block's 4096 counts are tile counts (<= 3 * 2048 vertices, <= 5 * 2048 triangles, <= 8192 points):
  // the block-local sums fit 32 bits
  const uint32_t mine = c[0] + c[1] + c[2] + c[3];
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  const uint32_t w = s_warp[lane];
  uint32_t wi = w;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
    if (lane >= o) wi += v;
  }
  uint32_t run = __shfl_sync(0xffffffffu, wi - w, warp) + incl - mine;
#pragma unroll
  for (int j = 0; j < kScanPer; ++j) {
    if (i0 + j < p.m) off[i0 + j] = run;
    run += c[j];
  }
  if (tid == kScanBlock - 1) p.tot[blockIdx.y][blockIdx.x] = run;
}

__global__ void __launch_bounds__(kScanBlock)
k_scan_carry(ScanP p) {
  __shared__ int64_t s_part[32];
  int64_t* __restrict__ off = p.off[blockIdx.y];
  const int64_t* __restrict__ tot = p.tot[blockIdx.y];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int64_t sum = 0;
  for (int j = tid; j < (int)blockIdx.x; j += kScanBlock) sum += tot[j];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) s_part[warp] = sum;
  __syncthreads();
  int64_t carry = s_part[lane];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) carry += __shfl_xor_sync(0xffffffffu, carry, o);
  const int64_t i0 = (int64_t)blockIdx.x * kScanTile + kScanPer * tid;
  if (blockIdx.x > 0) {
#pragma unroll
    for (int j = 0; j < kScanPer; ++j)
      if (i0 + j < p.m) off[i0 + j] += carry;
  }
  if (blockIdx.x == gridDim.x - 1 && tid == 0) off[p.m] = carry + tot[blockIdx.x];
  // list of the entries with a non-zero count (the emit kernels of mesh.cu launch one block per
  // listed tile instead of one per tile): one atomic per warp
  uint32_t* __restrict__ list = p.list[blockIdx.y];
  if (list) {
    const uint32_t* __restrict__ cnt = p.cnt[blockIdx.y];
    uint32_t flags = 0;
#pragma unroll
    for (int j = 0; j < kScanPer; ++j)
      if (i0 + j < p.m && cnt[i0 + j] != 0u) flags |= 1u << j;
    const uint32_t mine = __popc(flags);
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    uint32_t base = 0;
    if (lane == 31 && incl) base = atomicAdd(p.nlist[blockIdx.y], incl);
    uint32_t pos = __shfl_sync(0xffffffffu, base, 31) + incl - mine;
#pragma unroll
    for (int j = 0; j < kScanPer; ++j)
      if (flags & (1u << j)) list[pos++] = (uint32_t)(i0 + j);
  }
}

struct EmitP {
  int64_t n;
  int dx, dy, dz, dzl, z0;
  int zfast;
  float ox, oy, oz, vs;
  const float* color;
};

__global__ void __launch_bounds__(kThreads)
k_emit(const uint32_t* __restrict__ ballots, const int64_t* __restrict__ tile_off, EmitP p,
       float* __restrict__ xyz, uint8_t* __restrict__ rgb, int64_t* __restrict__ idx) {
  __shared__ uint32_t s_b[kChunksPerTile * 4];
  __shared__ uint32_t s_choff[kChunksPerTile + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t base = tile_off[tile];
  if (tile_off[tile + 1] == base) return;  // empty tile (block-uniform)
  {
    const int64_t chunk = tile * kChunksPerTile + (threadIdx.x >> 2);
    s_b[threadIdx.x] = (chunk * 128 < p.n) ? ballots[chunk * 4 + (threadIdx.x & 3)] : 0u;
  }
  __syncthreads();
  if (threadIdx.x < kChunksPerTile) {
    s_choff[threadIdx.x + 1] = __popc(s_b[threadIdx.x * 4]) + __popc(s_b[threadIdx.x * 4 + 1]) +
                               __popc(s_b[threadIdx.x * 4 + 2]) + __popc(s_b[threadIdx.x * 4 + 3]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    s_choff[0] = 0;
    for (int i = 1; i <= kChunksPerTile; ++i) s_choff[i] += s_choff[i - 1];
  }
  __syncthreads();
  const uint32_t lt = (1u << lane) - 1u;
  for (int c = 0; c < kChunksPerWarp; ++c) {
    const int q = warp * kChunksPerWarp + c;
    if (s_choff[q + 1] == s_choff[q]) continue;
    const uint32_t b0 = s_b[q * 4], b1 = s_b[q * 4 + 1], b2 = s_b[q * 4 + 2], b3 = s_b[q * 4 + 3];
    int64_t pos = base + s_choff[q] + __popc(b0 & lt) + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt);
    const uint32_t bb[4] = {b0, b1, b2, b3};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (!((bb[j] >> lane) & 1u)) continue;
      const int64_t i = (tile * kChunksPerTile + q) * 128 + 4 * lane + j;  // local linear index
      int x, y, zl;
      if (p.zfast) {
        const int64_t row = i / p.dzl;
        zl = (int)(i - row * p.dzl);
        x = (int)(row / p.dy);
        y = (int)(row - (int64_t)x * p.dy);
      } else {  // src/tsdf.cu:201-203
        const int64_t row = i / p.dx;
        x = (int)(i - row * p.dx);
        zl = (int)(row / p.dy);
        y = (int)(row - (int64_t)zl * p.dy);
      }
      const int z = zl + p.z0;
      if (xyz) {
        xyz[3 * pos + 0] = p.ox + (float)x * p.vs;  // :206
        xyz[3 * pos + 1] = p.oy + (float)y * p.vs;  // :207
        xyz[3 * pos + 2] = p.oz + (float)z * p.vs;  // :208
      }
      if (idx) idx[pos] = p.zfast ? ((int64_t)x * p.dy + y) * p.dz + z : ((int64_t)z * p.dy + y) * p.dx + x;
      if (rgb) {
        uint32_t cc = p.color ? (uint32_t)p.color[i] : 0u;  // packed value is an exact integer
        rgb[3 * pos + 0] = (uint8_t)(cc & 255u);
        rgb[3 * pos + 1] = (uint8_t)((cc >> 8) & 255u);
        rgb[3 * pos + 2] = (uint8_t)((cc >> 16) & 255u);
      }
      ++pos;
    }
  }
}

}  // namespace

size_t scan_scratch_bytes(int64_t m) {  // block totals of one array, 256-byte padded
  return ((size_t)((m + kScanTile - 1) / kScanTile + 1) * 8 + 255) & ~(size_t)255;
}

int launch_scan(const uint32_t* d_cnt, int64_t m, int64_t* d_off, int64_t* d_tot, cudaStream_t st,
                const uint32_t* d_cnt_b, int64_t* d_off_b, int64_t* d_tot_b, uint32_t* d_list,
                uint32_t* d_list_b, uint32_t* d_nlist2) {
  if (m <= 0) {
    CU_CHECK(cudaMemsetAsync(d_off, 0, 8, st));
    if (d_off_b) CU_CHECK(cudaMemsetAsync(d_off_b, 0, 8, st));
    return TSDF_OK;
  }
  ScanP p;
  p.cnt[0] = d_cnt; p.off[0] = d_off; p.tot[0] = d_tot;
  p.cnt[1] = d_cnt_b; p.off[1] = d_off_b; p.tot[1] = d_tot_b;
  p.list[0] = d_list; p.list[1] = d_list_b;
  p.nlist[0] = d_nlist2; p.nlist[1] = d_nlist2 ? d_nlist2 + 1 : nullptr;
  p.m = m;
  const dim3 grid((unsigned)((m + kScanTile - 1) / kScanTile), d_cnt_b ? 2 : 1);
  k_scan_local<<<grid, kScanBlock, 0, st>>>(p);
  CU_LAUNCHED();
  k_scan_carry<<<grid, kScanBlock, 0, st>>>(p);
  CU_LAUNCHED();
  return TSDF_OK;
}

int ensure_scratch(tsdf_vol* v, size_t bytes) {
  if (v->scratch_bytes >= bytes) return TSDF_OK;
  if (v->d_scratch) cudaFree(v->d_scratch);
  v->d_scratch = nullptr;
  v->mark_valid = false;
  v->scratch_bytes = 0;
  CU_CHECK(cudaMalloc(&v->d_scratch, bytes));
  v->scratch_bytes = bytes;
  return TSDF_OK;
}

// Outputs are device pointers (may be null).  *count always receives the match count.
int extract_device(tsdf_vol* v, int predicate, float tsdf_thresh, float weight_thresh, float* d_xyz,
                   uint8_t* d_rgb, int64_t* d_idx, int64_t cap, int64_t* count) {
  const int64_t n = v->n_local;
  const int64_t n_tiles = (n + kTileVox - 1) / kTileVox;
  const int64_t n_chunks = (n + 127) / 128;
  // scratch layout: [tile_off int64 (n_tiles+1)] [tile_count u32 n_tiles] [ballots u32 4*n_chunks] [scan totals]
  size_t off_bytes = ((size_t)(n_tiles + 1) * 8 + 255) & ~(size_t)255;
  size_t cnt_bytes = ((size_t)n_tiles * 4 + 255) & ~(size_t)255;
  size_t bal_bytes = ((size_t)n_chunks * 16 + 255) & ~(size_t)255;
  int rc = ensure_scratch(v, off_bytes + cnt_bytes + bal_bytes + scan_scratch_bytes(n_tiles));
  if (rc) return rc;
  int64_t* d_tot = reinterpret_cast<int64_t*>(reinterpret_cast<char*>(v->d_scratch) + off_bytes + cnt_bytes + bal_bytes);
  int64_t* d_off = reinterpret_cast<int64_t*>(v->d_scratch);
  uint32_t* d_cnt = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(v->d_scratch) + off_bytes);
  uint32_t* d_bal = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(v->d_scratch) + off_bytes + cnt_bytes);
  cudaStream_t st = v->s_compute;
  // The usual pattern is a sizing call followed by the real one: the ballots and offsets of the
  // first are still valid for the second when nothing has touched the volume in between.
  const bool cached = v->mark_valid && v->mark_epoch == v->epoch && v->mark_pred == predicate &&
                      v->mark_tt == tsdf_thresh && v->mark_wt == weight_thresh;
  int64_t total = v->mark_total;
  if (!cached) {
    v->mark_valid = false;
    k_mark<<<(unsigned)n_tiles, kThreads, 0, st>>>(v->d_tsdf, v->d_weight, n, predicate, tsdf_thresh,
                                                   weight_thresh, d_bal, d_cnt);
    CU_LAUNCHED();
    if ((rc = launch_scan(d_cnt, n_tiles, d_off, d_tot, st))) return rc;
    CU_CHECK(cudaMemcpyAsync(&total, d_off + n_tiles, 8, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaStreamSynchronize(st));
    v->mark_valid = true; v->mark_epoch = v->epoch; v->mark_pred = predicate;
    v->mark_tt = tsdf_thresh; v->mark_wt = weight_thresh; v->mark_total = total;
  }
  if (count) *count = total;
  if (!d_xyz && !d_rgb && !d_idx) return TSDF_OK;  // sizing call
  if (cap < total) {
    tsdf_set_error("extract: capacity %lld < %lld points", (long long)cap, (long long)total);
    return TSDF_E_CAPACITY;
  }
  if (total == 0) return TSDF_OK;
  EmitP p;
  p.n = n;
  p.dx = v->cfg.dim_x; p.dy = v->cfg.dim_y; p.dz = v->cfg.dim_z; p.dzl = v->dzl; p.z0 = v->cfg.z_begin;
  p.zfast = v->zfast ? 1 : 0;
  p.ox = v->cfg.origin[0]; p.oy = v->cfg.origin[1]; p.oz = v->cfg.origin[2]; p.vs = v->cfg.voxel_size;
  p.color = v->d_color;
  k_emit<<<(unsigned)n_tiles, kThreads, 0, st>>>(d_bal, d_off, p, d_xyz, d_rgb, d_idx);
  CU_LAUNCHED();
  return TSDF_OK;   // NOT synchronised: the caller's copies / sends follow on the same stream
}

// ---- merge of per-slab point runs (multi-GPU readback: comm.cu, group.cu) ------------------------
namespace {
// Records of all ranks sit rank after rank in G_* (rank r: [offs[r], offs[r+1])), each rank's run
// ascending by GLOBAL reference index, all indices distinct.  The merged position of a record is its
// position in its own run plus, for every other run, the number of records with a smaller index
// (a binary search): no atomics, deterministic, ascending global order.
__global__ void k_merge_scatter(int64_t total, int nranks, const int64_t* __restrict__ offs,
                                const float* __restrict__ g_xyz, const int64_t* __restrict__ g_idx,
                                const uint8_t* __restrict__ g_rgb, float* __restrict__ o_xyz,
                                int64_t* __restrict__ o_idx, uint8_t* __restrict__ o_rgb) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < total; j += (int64_t)gridDim.x * blockDim.x) {
    int r = 0;
    while (r + 1 < nranks && j >= offs[r + 1]) ++r;
    const int64_t key = g_idx[j];
    int64_t pos = j - offs[r];
    for (int q = 0; q < nranks; ++q) {
      if (q == r) continue;
      int64_t lo = offs[q], hi = offs[q + 1];
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (g_idx[mid] < key) lo = mid + 1; else hi = mid;
      }
      pos += lo - offs[q];
    }
    o_idx[pos] = key;
    o_xyz[3 * pos] = g_xyz[3 * j]; o_xyz[3 * pos + 1] = g_xyz[3 * j + 1]; o_xyz[3 * pos + 2] = g_xyz[3 * j + 2];
    if (g_rgb) { o_rgb[3 * pos] = g_rgb[3 * j]; o_rgb[3 * pos + 1] = g_rgb[3 * j + 1]; o_rgb[3 * pos + 2] = g_rgb[3 * j + 2]; }
  }
}

}  // namespace

int launch_merge_scatter(int64_t total, int nranks, const int64_t* d_offs, const float* g_xyz, const int64_t* g_idx,
                         const uint8_t* g_rgb, float* o_xyz, int64_t* o_idx, uint8_t* o_rgb, cudaStream_t st) {
  if (total <= 0) return TSDF_OK;
  const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
  k_merge_scatter<<<blocks, 256, 0, st>>>(total, nranks, d_offs, g_xyz, g_idx, g_rgb, o_xyz, o_idx, o_rgb);
  CU_LAUNCHED();
  return TSDF_OK;
}

// ---- point extraction for callers that may not allocate (tsdf_extract_points_static) ---------------
// No counterpart in the reference.  Same three passes as extract_device (k_mark -> scan -> k_emit), but
// the ballots / tile counts / offsets live in ONE file-scope device array instead of a cudaMalloc'ed,
// grow-on-demand buffer owned by a handle, and the voxel arrays and outputs are the caller's.  The host
// path below calls no cudaMalloc / cudaFree / cudaHostAlloc / stream or event creation.
namespace {

constexpr int64_t kStaticMaxVoxels = (int64_t)1 << 27;  // 512^3; the reference's own grids are 200^3 / 500^3
constexpr size_t pad256(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr int64_t kStaticMaxTiles = (kStaticMaxVoxels + kTileVox - 1) / kTileVox;
constexpr size_t kStaticOffBytes = pad256((size_t)(kStaticMaxTiles + 1) * 8);
constexpr size_t kStaticCntBytes = pad256((size_t)kStaticMaxTiles * 4);
constexpr size_t kStaticBalBytes = pad256((size_t)((kStaticMaxVoxels + 127) / 128) * 16);
constexpr size_t kStaticTotBytes = pad256((size_t)((kStaticMaxTiles + kScanTile - 1) / kScanTile + 1) * 8);
constexpr size_t kStaticScratchBytes = kStaticOffBytes + kStaticCntBytes + kStaticBalBytes + kStaticTotBytes;  // ~16.2 MiB

// Lives in every context that loads this module, from load until process exit.  Host code must never
// hand `g_static_scratch` itself to a kernel: on the host the name is only a handle for the runtime's
// symbol table; the device address comes from cudaGetSymbolAddress (static_scratch_for below).
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
void* g_static_ptr[kStaticMaxDevices];          // device address of g_static_scratch, per device ordinal
std::atomic<bool> g_static_known[kStaticMaxDevices];

// Looks the array's device address up once per device.  `device` must be the current device (each
// device has its own copy of the array at its own address).
int static_scratch_for(int device, unsigned char** out) {
  if (!g_static_known[device].load(std::memory_order_acquire)) {
    void* p = nullptr;
    CU_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) {
      tsdf_set_error("tsdf_extract_points_static: no device address for the static scratch on device %d", device);
      return TSDF_E_CUDA;
    }
    g_static_ptr[device] = p;   // two racing first calls store the same value
    g_static_known[device].store(true, std::memory_order_release);
  }
  *out = static_cast<unsigned char*>(g_static_ptr[device]);
  return TSDF_OK;
}

}  // namespace

extern "C" int tsdf_extract_points_static(const tsdf_config* cfg, const float* d_tsdf, const float* d_weight,
                                          const float* d_color, int predicate, float tsdf_thresh,
                                          float weight_thresh, float* d_xyz, uint8_t* d_rgb, int64_t* d_idx,
                                          int64_t cap, int64_t* count, void* stream) {
  REQUIRE(cfg && d_tsdf && d_weight && count, "null argument");
  REQUIRE(cfg->dim_x > 0 && cfg->dim_y > 0 && cfg->dim_z > 0, "grid dims must be positive");
  REQUIRE(predicate == TSDF_PRED_NONZERO || predicate == TSDF_PRED_ABS_LT, "unknown predicate");
  REQUIRE(cfg->semantics == TSDF_SEM_TSDFCU || cfg->semantics == TSDF_SEM_FUSIONPY, "unknown semantics");
  int z0 = cfg->z_begin, z1 = cfg->z_end;
  if (z0 == 0 && z1 == 0) z1 = cfg->dim_z;
  REQUIRE(z0 >= 0 && z0 < z1 && z1 <= cfg->dim_z, "bad z slab");
  const int dzl = z1 - z0;
  const int64_t n = (int64_t)cfg->dim_x * cfg->dim_y * dzl;
  if (n > kStaticMaxVoxels) {
    tsdf_set_error("tsdf_extract_points_static: %lld voxels exceed the static scratch limit of %lld "
                   "(2^27 = 512^3); larger grids need tsdf_extract_points_device on a handle",
                   (long long)n, (long long)kStaticMaxVoxels);
    return TSDF_E_INVALID;
  }
  int ndev = 0;
  int rc = tsdf_device_count(&ndev);
  if (rc) return rc;
  REQUIRE(cfg->device >= 0 && cfg->device < ndev && cfg->device < kStaticMaxDevices, "device ordinal out of range");
  DeviceGuard g(cfg->device);
  unsigned char* base = nullptr;
  if ((rc = static_scratch_for(cfg->device, &base))) return rc;
  const int64_t n_tiles = (n + kTileVox - 1) / kTileVox;
  const int64_t n_chunks = (n + 127) / 128;
  // same carve-up as extract_device, at the offsets of the LARGEST grid (so the layout is call-independent)
  int64_t* d_off = reinterpret_cast<int64_t*>(base);
  uint32_t* d_cnt = reinterpret_cast<uint32_t*>(base + kStaticOffBytes);
  uint32_t* d_bal = reinterpret_cast<uint32_t*>(base + kStaticOffBytes + kStaticCntBytes);
  int64_t* d_tot = reinterpret_cast<int64_t*>(base + kStaticOffBytes + kStaticCntBytes + kStaticBalBytes);
  (void)n_chunks;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Nothing in the array has to start at zero: k_mark writes every ballot word and tile count of this
  // grid, the scan writes every offset (and off[n_tiles]) before k_emit reads them, so whatever the
  // previous call left behind is overwritten before it is read.
  k_mark<<<(unsigned)n_tiles, kThreads, 0, st>>>(d_tsdf, d_weight, n, predicate, tsdf_thresh, weight_thresh,
                                                 d_bal, d_cnt);
  CU_LAUNCHED();
  if ((rc = launch_scan(d_cnt, n_tiles, d_off, d_tot, st))) return rc;
  int64_t total = 0;
  CU_CHECK(cudaMemcpyAsync(&total, d_off + n_tiles, 8, cudaMemcpyDeviceToHost, st));
  CU_CHECK(cudaStreamSynchronize(st));
  *count = total;
  if (!d_xyz && !d_rgb && !d_idx) return TSDF_OK;  // sizing call
  if (cap < total) {
    tsdf_set_error("tsdf_extract_points_static: capacity %lld < %lld points", (long long)cap, (long long)total);
    return TSDF_E_CAPACITY;
  }
  if (total == 0) return TSDF_OK;
  EmitP p;
  p.n = n;
  p.dx = cfg->dim_x; p.dy = cfg->dim_y; p.dz = cfg->dim_z; p.dzl = dzl; p.z0 = z0;
  p.zfast = cfg->semantics == TSDF_SEM_FUSIONPY ? 1 : 0;
  p.ox = cfg->origin[0]; p.oy = cfg->origin[1]; p.oz = cfg->origin[2]; p.vs = cfg->voxel_size;
  p.color = d_color;
  k_emit<<<(unsigned)n_tiles, kThreads, 0, st>>>(d_bal, d_off, p, d_xyz, d_rgb, d_idx);
  CU_LAUNCHED();
  CU_CHECK(cudaStreamSynchronize(st));
  return TSDF_OK;
}
