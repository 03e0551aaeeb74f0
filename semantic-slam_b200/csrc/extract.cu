// extract.cu — surface / point-cloud extraction (K5) as warp-ballot compaction.
//
// Replaces the two serial host scans of TSDF::SaveVoxelGrid2SurfacePointCloud (reference
// src/tsdf.cu:170-218).  Output order is ascending reference linear index, i.e. exactly the
// order the reference's loop emits (deterministic, no atomics on the output cursor):
//   pass 1  k_mark : 128-bit coalesced read of tsdf+weight (8 B/voxel, the only full sweep),
//                    predicate -> four warp ballots per 128-voxel chunk (16 B) + per-tile count
//   scan    k_scan : exclusive scan of the tile counts (one block; <= a few 100k entries)
//   pass 2  k_emit : reads only the ballots; rank = popc of lower bits; writes xyz
//                    (origin + (float)i*vs, src/tsdf.cu:206-208), GLOBAL reference index and
//                    rgb (colour gathered for kept voxels only)
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

// Exclusive scan of m uint32 counts into int64 offsets; offsets[m] = total.  One block of
// 1024 threads walks the array in coalesced 4096-entry tiles (uint4 per thread), scans each
// tile with warp shuffles and carries the running total (<= 131072 entries at 1024^3).
// A second (count, offset) pair of the same length rides in block 1 of the same launch (mesh.cu).
__global__ void __launch_bounds__(1024)
k_scan(const uint32_t* __restrict__ cnt, int64_t m, int64_t* __restrict__ off,
       const uint32_t* __restrict__ cnt_b, int64_t* __restrict__ off_b) {
  if (blockIdx.x == 1) { cnt = cnt_b; off = off_b; }
  __shared__ int64_t s_warp[32];
  __shared__ int64_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < m; base += 4096) {
    const int64_t i0 = base + 4 * tid;
    uint32_t c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] = (i0 + j < m) ? cnt[i0 + j] : 0u;
    const int64_t mine = (int64_t)c[0] + c[1] + c[2] + c[3];
    int64_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int64_t w = s_warp[lane], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t v = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += v;
      }
      s_warp[lane] = wi - w;  // exclusive prefix of the warp totals
    }
    __syncthreads();
    int64_t run = s_carry + s_warp[warp] + incl - mine;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (i0 + j < m) off[i0 + j] = run;
      run += c[j];
    }
    __syncthreads();                      // everyone has read s_carry / s_warp
    if (tid == 1023) s_carry = run;       // inclusive total so far
    __syncthreads();
  }
  if (tid == 0) off[m] = s_carry;
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

int launch_scan(const uint32_t* d_cnt, int64_t m, int64_t* d_off, cudaStream_t st,
                const uint32_t* d_cnt_b, int64_t* d_off_b) {
  k_scan<<<d_cnt_b ? 2 : 1, 1024, 0, st>>>(d_cnt, m, d_off, d_cnt_b, d_off_b);
  CU_CHECK(cudaGetLastError());
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
  // scratch layout: [tile_off int64 (n_tiles+1)] [tile_count u32 n_tiles] [ballots u32 4*n_chunks]
  size_t off_bytes = ((size_t)(n_tiles + 1) * 8 + 255) & ~(size_t)255;
  size_t cnt_bytes = ((size_t)n_tiles * 4 + 255) & ~(size_t)255;
  size_t bal_bytes = (size_t)n_chunks * 16;
  int rc = ensure_scratch(v, off_bytes + cnt_bytes + bal_bytes);
  if (rc) return rc;
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
    CU_CHECK(cudaGetLastError());
    k_scan<<<1, 1024, 0, st>>>(d_cnt, n_tiles, d_off, nullptr, nullptr);
    CU_CHECK(cudaGetLastError());
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
  CU_CHECK(cudaGetLastError());
  CU_CHECK(cudaStreamSynchronize(st));
  return TSDF_OK;
}
