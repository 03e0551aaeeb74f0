// mesh.cu — triangle-mesh readback: level-0 marching cubes of the tsdf array as warp-ballot
// compaction.  Replaces fusion.TSDFVolume.get_mesh(), which the reference's plugin calls from
// SaveMesh (reference src/TSDFfusion.py.in:48-53: `verts,faces,norms,colors = tsdf_vol.get_mesh()`
// then `fusion.meshwrite`); upstream runs skimage's marching cubes on the host over the whole
// volume.  skimage is un-vendored and absent, so the mesh is DEFINED by the restatement in
// oracle/tsdf_oracle.c (oracle_mesh) and checked against it bit for bit; parity with skimage's
// vertex/face ORDER is unpinned.
//
// Indexed mesh, every vertex emitted once:
//   vertex   one per grid edge (voxel i, axis a) whose two voxels differ in (tsdf < 0);
//            id = rank of (i, a) in ascending (storage index, axis) order
//   k_mc_mark   one thread per voxel, 8 corner loads (coalesced along the fast axis, the seven
//               neighbours come from L1/L2): three edge ballots per 32-voxel group, the cell's
//               8-bit case, per-tile vertex and triangle counts; one 16-byte record per group
//               {ballot_x, ballot_y, ballot_z, vertices before this group in its tile}
//   k_scan      (extract.cu) exclusive scans of the two tile-count arrays
//   k_mc_verts  rank = tile offset + record prefix + popc of lower ballot bits; writes position
//               (t = a/(a-b) along the edge), gradient normal, colour of the nearer voxel
//   k_mc_faces  per cell: triangles of mc_table.inc; a vertex id is ONE 16-byte record load + popc
// Algorithmic bytes: 4 B/voxel read + 1.5 B/voxel written by the mark pass (the only sweep), then
// 1 B/voxel re-read of the case bytes and O(surface) output.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kIters = 8;
constexpr int kGroupsPerTile = kWarps * kIters;             // 64 groups of 32 voxels
constexpr int64_t kTileVox = (int64_t)kGroupsPerTile * 32;  // 2048

__device__ __align__(16) const uint8_t d_mc_table[256][16] = {
#include "mc_table.inc"
};

struct MeshP {
  int64_t n;          // voxels of the slab
  int n_fast, n_mid;  // storage: i = (slow*n_mid + mid)*n_fast + fast
  int zfast;          // FUSIONPY [x][y][z] (fast = z) or TSDFCU [z][y][x] (fast = x)
  int dim[3];         // local dims along x, y, z
  int goff[3];        // global index of local 0 along x, y, z (z: the slab's z_begin)
  int64_t st[3];      // storage strides along x, y, z
  float o[3], vs;
  const float* t;
  const float* color;
};

__device__ __forceinline__ void decode(const MeshP& p, int64_t i, int co[3]) {
  int f, m, s;
  if (p.n <= 0x7fffffffLL) {
    const uint32_t u = (uint32_t)i, row = u / (uint32_t)p.n_fast;
    f = (int)(u - row * (uint32_t)p.n_fast);
    s = (int)(row / (uint32_t)p.n_mid);
    m = (int)(row - (uint32_t)s * (uint32_t)p.n_mid);
  } else {
    const int64_t row = i / p.n_fast;
    f = (int)(i - row * p.n_fast);
    s = (int)(row / p.n_mid);
    m = (int)(row - (int64_t)s * p.n_mid);
  }
  co[1] = m;
  if (p.zfast) { co[0] = s; co[2] = f; } else { co[0] = f; co[2] = s; }
}

__global__ void __launch_bounds__(kThreads)
k_mc_mark(MeshP p, uint4* __restrict__ recs, uint8_t* __restrict__ cases,
          uint32_t* __restrict__ tile_verts, uint32_t* __restrict__ tile_tris) {
  __shared__ uint32_t s_b[3][kGroupsPerTile];
  __shared__ uint32_t s_vc[kGroupsPerTile], s_tc[kGroupsPerTile];
  __shared__ uint8_t s_ntri[256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.x;
  s_ntri[threadIdx.x] = d_mc_table[threadIdx.x][0];
  __syncthreads();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
    const int g = it * kWarps + warp;
    const int64_t i = tile * kTileVox + (int64_t)g * 32 + lane;
    bool ex = false, ey = false, ez = false;
    uint32_t cs = 0;
    if (i < p.n) {
      int co[3];
      decode(p, i, co);
      const bool hx = co[0] + 1 < p.dim[0], hy = co[1] + 1 < p.dim[1], hz = co[2] + 1 < p.dim[2];
      const float* q = p.t + i;
      const bool in0 = __ldg(q) < 0.0f;
      if (hx) ex = in0 != (__ldg(q + p.st[0]) < 0.0f);
      if (hy) ey = in0 != (__ldg(q + p.st[1]) < 0.0f);
      if (hz) ez = in0 != (__ldg(q + p.st[2]) < 0.0f);
      if (hx && hy && hz) {
        // corner c = dx + 2*dy + 4*dz
        cs = (uint32_t)in0;
        cs |= (uint32_t)(in0 != ex) << 1;
        cs |= (uint32_t)(in0 != ey) << 2;
        cs |= (uint32_t)(__ldg(q + p.st[0] + p.st[1]) < 0.0f) << 3;
        cs |= (uint32_t)(in0 != ez) << 4;
        cs |= (uint32_t)(__ldg(q + p.st[0] + p.st[2]) < 0.0f) << 5;
        cs |= (uint32_t)(__ldg(q + p.st[1] + p.st[2]) < 0.0f) << 6;
        cs |= (uint32_t)(__ldg(q + p.st[0] + p.st[1] + p.st[2]) < 0.0f) << 7;
      }
      cases[i] = (uint8_t)cs;
    }
    const uint32_t bx = __ballot_sync(0xffffffffu, ex);
    const uint32_t by = __ballot_sync(0xffffffffu, ey);
    const uint32_t bz = __ballot_sync(0xffffffffu, ez);
    const uint32_t nt = __reduce_add_sync(0xffffffffu, (uint32_t)s_ntri[cs]);
    if (lane == 0) {
      s_b[0][g] = bx; s_b[1][g] = by; s_b[2][g] = bz;
      s_vc[g] = __popc(bx) + __popc(by) + __popc(bz);
      s_tc[g] = nt;
    }
  }
  __syncthreads();
  if (warp == 0) {
    // exclusive scan of the 64 group vertex counts: lane owns groups 2*lane, 2*lane + 1
    const uint32_t c0 = s_vc[2 * lane], c1 = s_vc[2 * lane + 1];
    uint32_t incl = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const uint32_t excl = incl - (c0 + c1);
    const int64_t g0 = tile * kGroupsPerTile + 2 * lane;
    recs[g0] = make_uint4(s_b[0][2 * lane], s_b[1][2 * lane], s_b[2][2 * lane], excl);
    recs[g0 + 1] = make_uint4(s_b[0][2 * lane + 1], s_b[1][2 * lane + 1], s_b[2][2 * lane + 1], excl + c0);
    const uint32_t tt = __reduce_add_sync(0xffffffffu, s_tc[2 * lane] + s_tc[2 * lane + 1]);
    if (lane == 31) tile_verts[tile] = incl;
    if (lane == 0) tile_tris[tile] = tt;
  }
}

// tsdf gradient at voxel u (central differences, one-sided at the slab faces)
__device__ __forceinline__ void grad_at(const MeshP& p, int64_t u, const int co[3], float g[3]) {
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const bool lo = co[a] > 0, hi = co[a] + 1 < p.dim[a];
    const float* q = p.t + u;
    if (lo && hi) g[a] = __fmul_rn(__fsub_rn(__ldg(q + p.st[a]), __ldg(q - p.st[a])), 0.5f);
    else if (hi)  g[a] = __fsub_rn(__ldg(q + p.st[a]), __ldg(q));
    else if (lo)  g[a] = __fsub_rn(__ldg(q), __ldg(q - p.st[a]));
    else          g[a] = 0.0f;
  }
}

__global__ void __launch_bounds__(kThreads)
k_mc_verts(MeshP p, const uint4* __restrict__ recs, const int64_t* __restrict__ tile_off,
           float* __restrict__ verts, float* __restrict__ norms, uint8_t* __restrict__ cols) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t base = tile_off[tile];
  if (tile_off[tile + 1] == base) return;  // no vertex in this tile (block-uniform)
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
    const int g = it * kWarps + warp;
    const uint4 r = recs[tile * kGroupsPerTile + g];
    if (!(r.x | r.y | r.z)) continue;
    const uint32_t bits[3] = {(r.x >> lane) & 1u, (r.y >> lane) & 1u, (r.z >> lane) & 1u};
    if (!(bits[0] | bits[1] | bits[2])) continue;
    const int64_t i = tile * kTileVox + (int64_t)g * 32 + lane;
    int64_t pos = base + r.w + __popc(r.x & lt) + __popc(r.y & lt) + __popc(r.z & lt);
    int co[3];
    decode(p, i, co);
    const float va = __ldg(p.t + i);
    float g0[3];
    if (norms) grad_at(p, i, co, g0);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      if (!bits[a]) continue;
      const float vb = __ldg(p.t + i + p.st[a]);
      const float t = __fdiv_rn(va, __fsub_rn(va, vb));
      if (verts) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const float gk = (float)(co[k] + p.goff[k]);
          verts[3 * pos + k] = __fadd_rn(p.o[k], __fmul_rn(k == a ? __fadd_rn(gk, t) : gk, p.vs));
        }
      }
      if (cols) {
        const int gi = co[a] + p.goff[a];
        int k = (int)rintf(__fadd_rn((float)gi, t));  // np.round: half to even
        k = k < gi ? gi : (k > gi + 1 ? gi + 1 : k);
        const uint32_t cc = p.color ? (uint32_t)__ldg(p.color + i + (int64_t)(k - gi) * p.st[a]) : 0u;
        cols[3 * pos + 0] = (uint8_t)(cc & 255u);
        cols[3 * pos + 1] = (uint8_t)((cc >> 8) & 255u);
        cols[3 * pos + 2] = (uint8_t)((cc >> 16) & 255u);
      }
      if (norms) {
        int c1[3] = {co[0], co[1], co[2]};
        c1[a] += 1;
        float g1[3], nn[3];
        grad_at(p, i + p.st[a], c1, g1);
#pragma unroll
        for (int k = 0; k < 3; ++k) nn[k] = __fadd_rn(g0[k], __fmul_rn(t, __fsub_rn(g1[k], g0[k])));
        const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nn[0], nn[0]), __fmul_rn(nn[1], nn[1])),
                                               __fmul_rn(nn[2], nn[2])));
#pragma unroll
        for (int k = 0; k < 3; ++k) norms[3 * pos + k] = len > 0.0f ? __fdiv_rn(nn[k], len) : 0.0f;
      }
      ++pos;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
k_mc_faces(MeshP p, const uint4* __restrict__ recs, const uint8_t* __restrict__ cases,
           const int64_t* __restrict__ vert_off, const int64_t* __restrict__ tri_off,
           int32_t* __restrict__ faces) {
  __shared__ uint32_t s_tab[256 * 4];     // the table, 16 bytes per case
  __shared__ uint32_t s_goff[kGroupsPerTile + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t base = tri_off[tile];
  if (tri_off[tile + 1] == base) return;  // no triangle in this tile (block-uniform)
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&d_mc_table[0][0]);
#pragma unroll
    for (int k = 0; k < 4; ++k) s_tab[threadIdx.x + k * kThreads] = src[threadIdx.x + k * kThreads];
  }
  __syncthreads();
  const uint8_t* tab = reinterpret_cast<const uint8_t*>(s_tab);
  uint32_t my_case[kIters];
#pragma unroll
  for (int it = 0; it < kIters; ++it) {
    const int g = it * kWarps + warp;
    const int64_t i = tile * kTileVox + (int64_t)g * 32 + lane;
    my_case[it] = i < p.n ? (uint32_t)cases[i] : 0u;
    const uint32_t nt = __reduce_add_sync(0xffffffffu, (uint32_t)tab[my_case[it] * 16]);
    if (lane == 0) s_goff[g + 1] = nt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    s_goff[0] = 0;
    for (int g = 1; g <= kGroupsPerTile; ++g) s_goff[g] += s_goff[g - 1];
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < kIters; ++it) {
    const int g = it * kWarps + warp;
    if (s_goff[g + 1] == s_goff[g]) continue;  // warp-uniform
    const uint8_t* row = tab + my_case[it] * 16;
    const uint32_t nt = row[0];
    uint32_t incl = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const int64_t i = tile * kTileVox + (int64_t)g * 32 + lane;
    int64_t f = base + s_goff[g] + (incl - nt);
    for (uint32_t k = 0; k < nt; ++k, ++f) {
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const uint32_t e = row[1 + 3 * k + j], a = e >> 2, q = e & 3u;
        // the two fixed coordinates of edge e, in axis order (tools/gen_mc_table.py)
        const uint32_t d0 = a == 0 ? 0u : (q & 1u);
        const uint32_t d1 = a == 1 ? 0u : (a == 0 ? (q & 1u) : (q >> 1));
        const uint32_t d2 = a == 2 ? 0u : (q >> 1);
        const int64_t u = i + d0 * p.st[0] + d1 * p.st[1] + d2 * p.st[2];
        const int64_t gu = u >> 5;
        const uint32_t lu = (uint32_t)(u & 31), ltu = (1u << lu) - 1u;
        const uint4 r = recs[gu];
        int64_t id = vert_off[gu / kGroupsPerTile] + r.w + __popc(r.x & ltu) + __popc(r.y & ltu) + __popc(r.z & ltu);
        if (a > 0) id += (r.x >> lu) & 1u;
        if (a > 1) id += (r.y >> lu) & 1u;
        faces[3 * f + j] = (int32_t)id;
      }
    }
  }
}

void fill_mesh_params(const tsdf_vol* v, MeshP* p) {
  const tsdf_config& c = v->cfg;
  p->n = v->n_local;
  p->zfast = v->zfast ? 1 : 0;
  p->dim[0] = c.dim_x; p->dim[1] = c.dim_y; p->dim[2] = v->dzl;
  p->goff[0] = 0; p->goff[1] = 0; p->goff[2] = c.z_begin;
  if (v->zfast) {
    p->n_fast = v->dzl; p->n_mid = c.dim_y;
    p->st[2] = 1; p->st[1] = v->dzl; p->st[0] = (int64_t)c.dim_y * v->dzl;
  } else {
    p->n_fast = c.dim_x; p->n_mid = c.dim_y;
    p->st[0] = 1; p->st[1] = c.dim_x; p->st[2] = (int64_t)c.dim_x * c.dim_y;
  }
  p->o[0] = c.origin[0]; p->o[1] = c.origin[1]; p->o[2] = c.origin[2];
  p->vs = c.voxel_size;
  p->t = v->d_tsdf;
  p->color = v->d_color;
}

}  // namespace

// Outputs are device pointers (any may be null).  *n_verts / *n_faces always receive the counts;
// with all outputs null this is the sizing call (its mark pass is reused by the next call while
// the voxel arrays are unchanged).
int mesh_device(tsdf_vol* v, float* d_verts, float* d_norms, uint8_t* d_cols, int64_t vcap,
                int32_t* d_faces, int64_t fcap, int64_t* n_verts, int64_t* n_faces) {
  const int64_t n = v->n_local;
  const int64_t n_tiles = (n + kTileVox - 1) / kTileVox;
  // scratch: [vert_off i64 (T+1)] [tri_off i64 (T+1)] [tile_verts u32 T] [tile_tris u32 T]
  //          [records uint4 64*T] [cases u8 2048*T]
  const size_t off_b = ((size_t)(n_tiles + 1) * 8 + 255) & ~(size_t)255;
  const size_t cnt_b = ((size_t)n_tiles * 4 + 255) & ~(size_t)255;
  const size_t rec_b = (size_t)n_tiles * kGroupsPerTile * 16;
  const size_t cas_b = (size_t)n_tiles * kTileVox;
  const size_t need = 2 * off_b + 2 * cnt_b + rec_b + cas_b;
  if (v->mesh_bytes < need) {
    if (v->d_mesh) cudaFree(v->d_mesh);
    v->d_mesh = nullptr; v->mesh_bytes = 0; v->mesh_valid = false;
    CU_CHECK(cudaMalloc(&v->d_mesh, need));
    v->mesh_bytes = need;
  }
  char* s = reinterpret_cast<char*>(v->d_mesh);
  int64_t* d_voff = reinterpret_cast<int64_t*>(s);
  int64_t* d_toff = reinterpret_cast<int64_t*>(s + off_b);
  uint32_t* d_vcnt = reinterpret_cast<uint32_t*>(s + 2 * off_b);
  uint32_t* d_tcnt = reinterpret_cast<uint32_t*>(s + 2 * off_b + cnt_b);
  uint4* d_recs = reinterpret_cast<uint4*>(s + 2 * off_b + 2 * cnt_b);
  uint8_t* d_cases = reinterpret_cast<uint8_t*>(s + 2 * off_b + 2 * cnt_b + rec_b);
  cudaStream_t st = v->s_compute;
  MeshP p;
  fill_mesh_params(v, &p);
  if (!(v->mesh_valid && v->mesh_epoch == v->epoch)) {
    v->mesh_valid = false;
    k_mc_mark<<<(unsigned)n_tiles, kThreads, 0, st>>>(p, d_recs, d_cases, d_vcnt, d_tcnt);
    CU_CHECK(cudaGetLastError());
    int rc = launch_scan(d_vcnt, n_tiles, d_voff, st);
    if (rc) return rc;
    if ((rc = launch_scan(d_tcnt, n_tiles, d_toff, st))) return rc;
    int64_t tot[2] = {0, 0};
    CU_CHECK(cudaMemcpyAsync(&tot[0], d_voff + n_tiles, 8, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaMemcpyAsync(&tot[1], d_toff + n_tiles, 8, cudaMemcpyDeviceToHost, st));
    CU_CHECK(cudaStreamSynchronize(st));
    v->mesh_valid = true; v->mesh_epoch = v->epoch;
    v->mesh_nv = tot[0]; v->mesh_nf = tot[1];
  }
  const int64_t nv = v->mesh_nv, nf = v->mesh_nf;
  if (n_verts) *n_verts = nv;
  if (n_faces) *n_faces = nf;
  if (!d_verts && !d_norms && !d_cols && !d_faces) return TSDF_OK;  // sizing call
  if ((d_verts || d_norms || d_cols) && vcap < nv) {
    tsdf_set_error("mesh: vertex capacity %lld < %lld", (long long)vcap, (long long)nv);
    return TSDF_E_CAPACITY;
  }
  if (d_faces && fcap < nf) {
    tsdf_set_error("mesh: face capacity %lld < %lld", (long long)fcap, (long long)nf);
    return TSDF_E_CAPACITY;
  }
  if (d_faces && nv > 0x7fffffffLL) {
    tsdf_set_error("mesh: %lld vertices do not fit the int32 face indices of the PLY layout", (long long)nv);
    return TSDF_E_CAPACITY;
  }
  if (nv && (d_verts || d_norms || d_cols)) {
    k_mc_verts<<<(unsigned)n_tiles, kThreads, 0, st>>>(p, d_recs, d_voff, d_verts, d_norms, d_cols);
    CU_CHECK(cudaGetLastError());
  }
  if (nf && d_faces) {
    k_mc_faces<<<(unsigned)n_tiles, kThreads, 0, st>>>(p, d_recs, d_cases, d_voff, d_toff, d_faces);
    CU_CHECK(cudaGetLastError());
  }
  CU_CHECK(cudaStreamSynchronize(st));
  return TSDF_OK;
}
