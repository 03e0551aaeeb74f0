// integrate.cu — the voxel-integration kernel (K1) for sm_100a.
//
// Replaces __global__ GpuIntegrate (reference src/tsdf.cu:15-60) and the third-party
// integrate kernel behind src/TSDFfusion.py.in:43.  NOT a port: the reference maps one
// thread to a y row and walks x serially (uncoalesced, K/pose re-read from global memory
// per voxel).  Here
//   * the slab is stored with the reference-order fast axis contiguous ([x][y][z] for the
//     plugin surface, [z][y][x] for class TSDF); a thread owns a run of 4 voxels on that
//     axis and moves tsdf / weight / packed-RGB as 128-bit accesses, a warp covers 512 B
//     per array per row;
//   * pose, intrinsics and grid constants are kernel parameters (constant bank);
//   * the camera transform is split into a per-row part (staged once per block in shared
//     memory) and a per-run part (registers), so the per-voxel cost is three adds;
//   * volume loads are issued only when at least one voxel of the run is updated;
//   * k frames can be fused into one pass (state read once, updated k times, written once).
//
// Arithmetic: this file is compiled with -fmad=false; every fp32 operation below is a
// single IEEE round-to-nearest op in the reference's source order,
// so results are bit-identical to oracle/tsdf_oracle.c.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxRowsPerBlock = 128;

template <int NF>
struct LaunchP {
  GridP g;
  FrameP f[NF];
};

template <int VEC> struct VecT;
template <> struct VecT<4> { using type = float4; };
template <> struct VecT<1> { using type = float; };

template <int VEC>
__device__ __forceinline__ void load_vec(const float* p, float (&v)[VEC]) {
  if constexpr (VEC == 4) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  } else {
    v[0] = *p;
  }
}
template <int VEC>
__device__ __forceinline__ void store_vec(float* p, const float (&v)[VEC]) {
  if constexpr (VEC == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    *p = v[0];
  }
}

// One voxel, one frame: decides validity and produces the truncated distance and the
// pixel's packed colour.  Mirrors src/tsdf.cu:39-53.
template <bool COLOR, bool LE_TRUNC, bool MASK>
__device__ __forceinline__ bool probe(const FrameP& f, const GridP& g, const uint8_t* mask, float camx,
                                      float camy, float camz, float& dist, float& ncol) {
  bool ok = camz > 0.0f;                                  // :39
  float u = f.fx * (camx / camz) + f.cx;                  // :41
  float v = f.fy * (camy / camz) + f.cy;                  // :42
  float ru = roundf(u), rv = roundf(v);
  ok = ok && ru >= 0.0f && ru < (float)g.W && rv >= 0.0f && rv < (float)g.H;  // :43
  int pix = ok ? (int)rv * g.W + (int)ru : 0;
  float d;
  if constexpr (COLOR) {
    uint2 p = __ldg(reinterpret_cast<const uint2*>(f.img) + pix);  // :45 (+ colour)
    d = __uint_as_float(p.x);
    ncol = __uint_as_float(p.y);
  } else {
    d = __ldg(reinterpret_cast<const float*>(f.img) + pix);
    ncol = 0.0f;
  }
  // per-instance mask: masked depth = depth * (mask/255) (reference src/Engine.cpp:192-193),
  // i.e. the sample is invalid (0) where the 8-bit mask rounds to 0
  if constexpr (MASK) d = __ldg(mask + pix) >= 128 ? d : 0.0f;
  ok = ok && !(d <= 0.0f);                                // :46 (d > max_depth folded into prep)
  float diff = d - camz;                                  // :48
  if constexpr (LE_TRUNC) ok = ok && !(diff <= -g.trunc); // :49
  else                    ok = ok && !(diff < -g.trunc);
  dist = fminf(1.0f, diff / g.trunc);                     // :53
  return ok;
}

// Colour running average on the float-packed value (upstream algorithm).
__device__ __forceinline__ float blend_color(float oc, float nc, float w0, float w1) {
  float ob = floorf(oc / 65536.0f);
  float og = floorf((oc - ob * 65536.0f) / 256.0f);
  float orr = oc - ob * 65536.0f - og * 256.0f;
  float nb = floorf(nc / 65536.0f);
  float ng = floorf((nc - nb * 65536.0f) / 256.0f);
  float nr = nc - nb * 65536.0f - ng * 256.0f;
  nb = fminf(roundf((ob * w0 + nb) / w1), 255.0f);
  ng = fminf(roundf((og * w0 + ng) / w1), 255.0f);
  nr = fminf(roundf((orr * w0 + nr) / w1), 255.0f);
  return nb * 65536.0f + ng * 256.0f + nr;
}

// ZFAST : slab layout [x][y][zl]  (row = x*Dy + y, fast = zl)   — plugin surface
// !ZFAST: slab layout [zl][y][x]  (row = zl*Dy + y, fast = x)   — class TSDF, src/tsdf.cu:52
// Row table entry per frame: ZFAST  {A0,A1,A2}            A_k = c[k]*t0 + c[4+k]*t1
//                            !ZFAST {Y0,Y1,Y2,Z0,Z1,Z2}   Y_k = c[4+k]*t1, Z_k = c[8+k]*t2
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC, bool MASK>
__device__ __forceinline__ void integrate_body(const LaunchP<NF>& P, const uint8_t* mask, float* tab,
                                               int tpr_log2, int rows_per_block) {
  constexpr int NT = ZFAST ? 3 : 6;
  const GridP& g = P.g;
  const int n_rows = g.n_mid * g.n_slow;
  const int row0 = blockIdx.x * rows_per_block;
  if (row0 >= n_rows) return;  // batched launches size the grid for the largest instance

  // ---- prologue: per-row terms of the camera transform -> shared memory ----------------
  for (int lr = threadIdx.x; lr < rows_per_block; lr += kThreads) {
    int r = row0 + lr;
    if (r < n_rows) {
      int slow = r / g.n_mid, mid = r - slow * g.n_mid;
      float by = g.oy + (float)mid * g.vs;                         // :28
#pragma unroll
      for (int f = 0; f < NF; ++f) {
        const float* c = P.f[f].c;
        float t1 = by - c[7];                                      // :34
        float* e = tab + (lr * NF + f) * NT;
        if constexpr (ZFAST) {
          float bx = g.ox + (float)slow * g.vs;                    // :27
          float t0 = bx - c[3];                                    // :33
          e[0] = c[0] * t0 + c[4] * t1;
          e[1] = c[1] * t0 + c[5] * t1;
          e[2] = c[2] * t0 + c[6] * t1;
        } else {
          float bz = g.oz + (float)(g.z0 + slow) * g.vs;           // :29 (global z)
          float t2 = bz - c[11];                                   // :35
          e[0] = c[4] * t1; e[1] = c[5] * t1; e[2] = c[6] * t1;
          e[3] = c[8] * t2; e[4] = c[9] * t2; e[5] = c[10] * t2;
        }
      }
    }
  }
  __syncthreads();

  // ---- this thread's run on the fast axis -----------------------------------------------
  const int tpr = 1 << tpr_log2;
  const int tf = threadIdx.x & (tpr - 1);
  const int tr = threadIdx.x >> tpr_log2;
  const int rp = kThreads >> tpr_log2;  // rows processed in parallel by the block
  const int f0 = (blockIdx.y * tpr + tf) * VEC;
  if (f0 >= g.n_fast) return;

  float bf[VEC];  // base-frame coordinate of each voxel of the run along the fast axis
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    if constexpr (ZFAST) bf[j] = g.oz + (float)(g.z0 + f0 + j) * g.vs;   // :29
    else                 bf[j] = g.ox + (float)(f0 + j) * g.vs;           // :27
  }

  unsigned int n_written = 0;
  for (int lr = tr; lr < rows_per_block; lr += rp) {
    const int r = row0 + lr;
    if (r >= n_rows) break;
    bool okf[NF][VEC];
    float distf[NF][VEC];
    float ncolf[NF][VEC];
    bool any = false;
#pragma unroll
    for (int f = 0; f < NF; ++f) {
      const float* c = P.f[f].c;
      const float* e = tab + (lr * NF + f) * NT;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        float camx, camy, camz;
        if constexpr (ZFAST) {
          float t2 = bf[j] - c[11];                 // :35
          camx = e[0] + c[8] * t2;                  // :36
          camy = e[1] + c[9] * t2;                  // :37
          camz = e[2] + c[10] * t2;                 // :38
        } else {
          float t0 = bf[j] - c[3];                  // :33
          camx = (c[0] * t0 + e[0]) + e[3];
          camy = (c[1] * t0 + e[1]) + e[4];
          camz = (c[2] * t0 + e[2]) + e[5];
        }
        okf[f][j] = probe<COLOR, LE_TRUNC, MASK>(P.f[f], g, mask, camx, camy, camz, distf[f][j], ncolf[f][j]);
        any = any || okf[f][j];
      }
    }
    if (!any) continue;  // no volume traffic for a run with no update

    const int64_t base = (int64_t)r * g.n_fast + f0;
    float t[VEC], w[VEC], col[VEC];
    load_vec<VEC>(g.tsdf + base, t);
    load_vec<VEC>(g.weight + base, w);
    if constexpr (COLOR) load_vec<VEC>(g.color + base, col);
#pragma unroll
    for (int f = 0; f < NF; ++f) {
      const float obs = P.f[f].obs_weight;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        if (okf[f][j]) {
          float w0 = w[j];                          // :54
          float w1 = w0 + obs;                      // :55
          w[j] = w1;                                // :56
          t[j] = (t[j] * w0 + distf[f][j]) / w1;    // :57
          if constexpr (COLOR) col[j] = blend_color(col[j], ncolf[f][j], w0, w1);
          ++n_written;
        }
      }
    }
    store_vec<VEC>(g.tsdf + base, t);
    store_vec<VEC>(g.weight + base, w);
    if constexpr (COLOR) store_vec<VEC>(g.color + base, col);
  }

  // exact N_written: warp-reduce, one atomic per warp
  for (int o = 16; o > 0; o >>= 1) n_written += __shfl_down_sync(0xffffffffu, n_written, o);
  if ((threadIdx.x & 31) == 0 && n_written) atomicAdd(g.written, (unsigned long long)n_written);
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC>
__global__ void __launch_bounds__(kThreads)
k_integrate(const __grid_constant__ LaunchP<NF> P, int tpr_log2, int rows_per_block) {
  __shared__ float tab[kMaxRowsPerBlock * NF * (ZFAST ? 3 : 6)];
  integrate_body<ZFAST, COLOR, LE_TRUNC, NF, VEC, false>(P, nullptr, tab, tpr_log2, rows_per_block);
}

// Batched per-instance volumes (config 5): blockIdx.z selects a descriptor {grid, pose, mask}
// from global memory; one launch covers every active instance.
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int VEC>
__global__ void __launch_bounds__(kThreads)
k_integrate_batch(const BatchDesc* __restrict__ descs, int tpr_log2, int rows_per_block) {
  __shared__ float tab[kMaxRowsPerBlock * (ZFAST ? 3 : 6)];
  __shared__ LaunchP<1> P;
  __shared__ const uint8_t* mask;
  const BatchDesc& d = descs[blockIdx.z];
  if (threadIdx.x == 0) { P.g = d.g; P.f[0] = d.f; mask = d.mask; }
  __syncthreads();
  if (mask) integrate_body<ZFAST, COLOR, LE_TRUNC, 1, VEC, true>(P, mask, tab, tpr_log2, rows_per_block);
  else      integrate_body<ZFAST, COLOR, LE_TRUNC, 1, VEC, false>(P, nullptr, tab, tpr_log2, rows_per_block);
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC>
int launch_t(const tsdf_vol* v, const GridP& g, const FrameP* frames, cudaStream_t st) {
  LaunchP<NF> P;
  P.g = g;
  for (int i = 0; i < NF; ++i) P.f[i] = frames[i];
  int vec_per_row = (g.n_fast + VEC - 1) / VEC;
  int tpr_log2 = 0;
  while ((1 << tpr_log2) < vec_per_row && (1 << tpr_log2) < kThreads) ++tpr_log2;
  int tpr = 1 << tpr_log2;
  int rp = kThreads / tpr;
  int rows_per_block = rp * 16;
  if (rows_per_block > kMaxRowsPerBlock) rows_per_block = kMaxRowsPerBlock;
  if (rows_per_block < rp) rows_per_block = rp;
  int n_rows = g.n_mid * g.n_slow;
  dim3 grid((n_rows + rows_per_block - 1) / rows_per_block, (vec_per_row + tpr - 1) / tpr, 1);
  k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC><<<grid, kThreads, 0, st>>>(P, tpr_log2, rows_per_block);
  CU_CHECK(cudaGetLastError());
  return TSDF_OK;
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int VEC>
int launch_nf(const tsdf_vol* v, const GridP& g, const FrameP* frames, int nf, cudaStream_t st) {
  switch (nf) {
    case 1: return launch_t<ZFAST, COLOR, LE_TRUNC, 1, VEC>(v, g, frames, st);
    case 2: return launch_t<ZFAST, COLOR, LE_TRUNC, 2, VEC>(v, g, frames, st);
    case 4: return launch_t<ZFAST, COLOR, LE_TRUNC, 4, VEC>(v, g, frames, st);
    case 8: return launch_t<ZFAST, COLOR, LE_TRUNC, 8, VEC>(v, g, frames, st);
  }
  tsdf_set_error("launch_integrate: unsupported fused frame count %d", nf);
  return TSDF_E_INVALID;
}

}  // namespace

// nf in {1,2,4,8}.  The semantics of the handle pick layout and truncation compare.
int launch_integrate(tsdf_vol* v, const FrameP* frames, int nf, bool color, cudaStream_t st) {
  GridP g;
  const tsdf_config& c = v->cfg;
  g.n_fast = v->zfast ? v->dzl : c.dim_x;
  g.n_mid = c.dim_y;
  g.n_slow = v->zfast ? c.dim_x : v->dzl;
  g.z0 = c.z_begin;
  g.ox = c.origin[0]; g.oy = c.origin[1]; g.oz = c.origin[2];
  g.vs = c.voxel_size; g.trunc = c.trunc_margin;
  g.W = c.im_w; g.H = c.im_h;
  g.tsdf = v->d_tsdf; g.weight = v->d_weight; g.color = color ? v->d_color : nullptr;
  g.written = v->d_written;
  const bool vec4 = (g.n_fast % 4) == 0;
  const bool le = c.semantics == TSDF_SEM_TSDFCU;
#define DISPATCH(ZF, CO, LE)                                                     \
  (vec4 ? launch_nf<ZF, CO, LE, 4>(v, g, frames, nf, st) : launch_nf<ZF, CO, LE, 1>(v, g, frames, nf, st))
  // layout and truncation compare are both fixed by the semantics (tsdf_create)
  (void)le;
  if (v->zfast) return color ? DISPATCH(true, true, false) : DISPATCH(true, false, false);
  return color ? DISPATCH(false, true, true) : DISPATCH(false, false, true);
#undef DISPATCH
}

// All instances of a batch share semantics, colour mode and grid dims class (VEC); the grid
// is sized for the largest instance.  descs: n_active device descriptors.
int launch_integrate_batch(const BatchDesc* d_descs, int n_active, int max_fast, int max_rows,
                           bool zfast, bool color, bool vec4, cudaStream_t st) {
  const int VECv = vec4 ? 4 : 1;
  int vec_per_row = (max_fast + VECv - 1) / VECv;
  int tpr_log2 = 0;
  while ((1 << tpr_log2) < vec_per_row && (1 << tpr_log2) < kThreads) ++tpr_log2;
  int tpr = 1 << tpr_log2;
  int rp = kThreads / tpr;
  int rows_per_block = rp * 16;
  if (rows_per_block > kMaxRowsPerBlock) rows_per_block = kMaxRowsPerBlock;
  if (rows_per_block < rp) rows_per_block = rp;
  dim3 grid((max_rows + rows_per_block - 1) / rows_per_block, (vec_per_row + tpr - 1) / tpr, n_active);
#define LB(ZF, CO, LE, V) k_integrate_batch<ZF, CO, LE, V><<<grid, kThreads, 0, st>>>(d_descs, tpr_log2, rows_per_block)
  if (zfast) { if (color) { if (vec4) LB(true, true, false, 4); else LB(true, true, false, 1); }
               else       { if (vec4) LB(true, false, false, 4); else LB(true, false, false, 1); } }
  else       { if (color) { if (vec4) LB(false, true, true, 4); else LB(false, true, true, 1); }
               else       { if (vec4) LB(false, false, true, 4); else LB(false, false, true, 1); } }
#undef LB
  CU_CHECK(cudaGetLastError());
  return TSDF_OK;
}
