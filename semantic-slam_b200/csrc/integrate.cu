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
//   * two instantiations, picked per launch by a host-side frustum test: a dense-frame kernel
//     whose run is branch-free (batched reciprocals and gathers, select update) and a
//     sparse-frame kernel with a conservative, bit-exact segment cull (frustum, far depth,
//     max-depth tiles, mask box) decided by warp ballots;
//   * the kernel is written for NF frames per pass; only NF = 1 is instantiated (a fused pass
//     measured slower than NF passes, see launch_integrate).
//
// Arithmetic: this file is compiled with -fmad=false; every fp32 operation below is a
// single IEEE round-to-nearest op in the reference's source order,
// so results are bit-identical to oracle/tsdf_oracle.c.
#include <algorithm>
#include <climits>
#include <cmath>
#include "common.cuh"
#include "exactmath.cuh"

namespace {

// Dense kernels test the exponent window of camz once per run of 4 instead of per voxel: every fp32
// operation of the camera transform is monotone in the fast index, so the run's camz values lie
// between its two end voxels'; a run with an end outside the window (behind the camera, or absurdly
// far) takes the literal path as a whole.  -DKI_NO_RUNSAFE: A/B builds.
#ifdef KI_NO_RUNSAFE
constexpr bool kRunSafe = false;
#else
constexpr bool kRunSafe = true;
#endif
#ifndef KI_MINBLOCKS
#define KI_MINBLOCKS 4   // (A/B builds: -DKI_MINBLOCKS=3 lets the dense kernel take 80 registers)
#endif
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

template <bool L2_ONLY>
__device__ __forceinline__ void prefetch_line(const float* p, int mode) {
  if (L2_ONLY || mode == 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  else                      asm volatile("prefetch.global.L1 [%0];" ::"l"(p));  // A/B only (legacy kernel)
}

template <int VEC>
__device__ __forceinline__ void load_vec(const float* p, float (&v)[VEC]) {
  if constexpr (VEC == 4) {
    // (plain loads/stores: the streaming .cs hints measured 0.6 % slower on the dense frame)
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

// ---------------------------------------------------------------------------------------
// Per-voxel arithmetic.  Two spellings with identical results:
//   MATH_EXACT   (default) the reference's source-order IEEE operations evaluated with the
//                shared-reciprocal / Markstein / round-down-magic forms of exactmath.cuh
//   MATH_LITERAL the literal C spelling (`/`, roundf, floorf) — kept as the A/B baseline
//                (tsdf_set_variant(v, 1)) and as the fallback for out-of-range operands
// ---------------------------------------------------------------------------------------
constexpr int MATH_EXACT = 0, MATH_LITERAL = 1;

// ---------------------------------------------------------------------------------------
// Staged image tile (north_star: "depth and color tiles are staged through shared memory or TMA").
// When the pixel rectangle the block's voxel tile can project to (the block-level hull test computes
// it) fits one TMA box of kStageW x kStageH pixels, one elected thread fetches that box of the
// prepared frame into shared memory with cp.async.bulk.tensor.2d behind an mbarrier, and the block's
// gathers read shared memory instead of L1/L2: same pixel index -> same bits (src/tsdf.cu:45).
// Blocks whose rectangle does not fit keep the __ldg gather.  A/B variant (tsdf_set_variant bit 20, libtsdf_b200_dev.so only):
// on the C2 / C3 geometries 2 % / 1 % of the in-view blocks fit (a 128-voxel run along z spreads over
// ~100-200 pixels of parallax), see DESIGN.md §4.1.
// ---------------------------------------------------------------------------------------
constexpr int kStageW = 64, kStageH = 32;                 // pixels; 64 * 32 * 16 B = 32 KB
struct Stage {
  const float4* tile;   // shared memory, [kStageH][kStageW]
  int on, x0, y0;
};
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

struct Hit {          // result of probing one voxel against one frame
  bool ok;
  float dist;         // fmin(1, diff/trunc)                                   :53
  float c0, c1, c2;   // pixel colour channels as exact floats (colour mode)
};

template <bool COLOR, bool MASK>
__device__ __forceinline__ float fetch(const FrameP& f, const uint8_t* mask, int pix, Hit& h, const Stage* st = nullptr,
                                       int sidx = 0) {  // h: caller's registers
  float d;
  if constexpr (COLOR) {
    float4 p;
    if (st != nullptr && st->on) p = st->tile[sidx];   // staged box (block-uniform branch)
    else p = __ldg(reinterpret_cast<const float4*>(f.img) + pix);  // :45 + colour, one gather
    d = p.x; h.c0 = p.y; h.c1 = p.z; h.c2 = p.w;
  } else {
    d = __ldg(reinterpret_cast<const float*>(f.img) + pix);
    h.c0 = h.c1 = h.c2 = 0.0f;
  }
  // per-instance mask: masked depth = depth * (mask/255) (reference src/Engine.cpp:192-193)
  if constexpr (MASK) d = __ldg(mask + pix) >= 128 ? d : 0.0f;
  return d;
}

// Literal spelling of src/tsdf.cu:39-53.
template <bool COLOR, bool LE_TRUNC, bool MASK>
__device__ __forceinline__ Hit probe_literal_body(const FrameP& f, const GridP& g, const uint8_t* mask,
                                                  float camx, float camy, float camz) {
  Hit h;
  bool ok = camz > 0.0f;                                  // :39
  float u = f.fx * (camx / camz) + f.cx;                  // :41
  float v = f.fy * (camy / camz) + f.cy;                  // :42
  float ru = roundf(u), rv = roundf(v);
  ok = ok && ru >= 0.0f && ru < (float)g.W && rv >= 0.0f && rv < (float)g.H;  // :43
  int pix = ok ? (int)rv * g.W + (int)ru : 0;
  float d = fetch<COLOR, MASK>(f, mask, pix, h);
  ok = ok && !(d <= 0.0f);                                // :46 (d > max_depth folded into prep)
  float diff = d - camz;                                  // :48
  if constexpr (LE_TRUNC) ok = ok && !(diff <= -g.trunc); // :49
  else                    ok = ok && !(diff < -g.trunc);
  h.dist = fminf(1.0f, diff / g.trunc);                   // :53
  h.ok = ok;
  return h;
}
// out-of-line copy used as the rare fallback of the exact path (keeps the hot loop small)
template <bool COLOR, bool LE_TRUNC, bool MASK>
__device__ __noinline__ Hit probe_literal_cold(const void* img, float fx, float cx, float fy, float cy,
                                               float trunc, int W, int H, const uint8_t* mask,
                                               float camx, float camy, float camz) {
  FrameP f; f.img = img; f.fx = fx; f.cx = cx; f.fy = fy; f.cy = cy;
  GridP g; g.trunc = trunc; g.W = W; g.H = H;
  return probe_literal_body<COLOR, LE_TRUNC, MASK>(f, g, mask, camx, camy, camz);
}

// Same decisions and the same bits with the exactmath forms.  Operands outside the safe
// exponent window (never seen on real data) take the literal path.  One-piece form used by
// the dense (CULL = false) instantiation: the compiler is free to overlap the four gathers of
// a run with the projections.
template <bool COLOR, bool LE_TRUNC, bool MASK>
__device__ __forceinline__ Hit probe_exact(const FrameP& f, const GridP& g, const uint8_t* mask,
                                           float camx, float camy, float camz) {
  const bool inr = em::in_safe_range(camz);               // implies camz > 0            :39
  if (__builtin_expect(!inr && camz > 0.0f, 0))
    return probe_literal_cold<COLOR, LE_TRUNC, MASK>(f.img, f.fx, f.cx, f.fy, f.cy, g.trunc, g.W, g.H, mask,
                                                     camx, camy, camz);
  Hit h;
  const float yz = em::rcp_rn(camz);
  const float u = f.fx * em::div_rn(camx, camz, yz) + f.cx;   // :41
  const float v = f.fy * em::div_rn(camy, camz, yz) + f.cy;   // :42
  // (int)roundf(u) in [0,W)  <=>  -0.5 < u < W-0.5   (roundf is half-away-from-zero)   :43
  bool ok = inr && u > -0.5f && u < g.w_lim && v > -0.5f && v < g.h_lim;
  const int px = em::round_half_away_nonneg_int(u);
  const int py = em::round_half_away_nonneg_int(v);
  const int pix = ok ? py * g.W + px : 0;
  const float d = fetch<COLOR, MASK>(f, mask, pix, h);
  ok = ok && !(d <= 0.0f);                                // :46
  const float diff = d - camz;                            // :48
  if constexpr (LE_TRUNC) ok = ok && !(diff <= -g.trunc); // :49
  else                    ok = ok && !(diff < -g.trunc);
  h.dist = fminf(1.0f, em::div_rn(diff, g.trunc, g.inv_trunc));  // :53
  h.ok = ok;
  return h;
}

// The same in two stages (CULL = true instantiation), so that a run
// whose voxels all project outside the image stops before the gather.  Operands outside the
// safe exponent window (never seen on real data) take the literal path (caller's job).
struct Proj { bool ok; int pix; int px, py; };   // (px, py: only the staged-tile variant reads them)


// KNOWN_SAFE: the caller has shown camz inside the exponent window (dense kernels: once per run).
// INSIDE: the HOST has proved (slab_inside_with_margin: double-precision projection of the slab's 8 corners,
// one pixel of margin, an fp32 error bound of a quarter pixel) that every voxel of the launch has camz in
// the window and projects strictly inside the image, so the tests of :39 and :43 are true for every voxel
// and are not evaluated: 4 compares, their combination and the pixel-index select leave the hot loop.
template <bool SELECT, bool KNOWN_SAFE = false, bool INSIDE = false>
__device__ __forceinline__ Proj project_exact(const FrameP& f, const GridP& g, float camx, float camy, float camz) {
  const bool inr = KNOWN_SAFE || INSIDE || em::in_safe_range(camz); // implies camz > 0            :39
  const float yz = em::rcp_rn(camz);
  const float u = f.fx * em::div_rn(camx, camz, yz) + f.cx;   // :41
  const float v = f.fy * em::div_rn(camy, camz, yz) + f.cy;   // :42
  // (int)roundf(u) in [0,W)  <=>  -0.5 < u < W-0.5   (roundf is half-away-from-zero)   :43
  Proj p;
  const int px = em::round_half_away_nonneg_int(u);
  const int py = em::round_half_away_nonneg_int(v);
  p.px = px; p.py = py;
  if constexpr (INSIDE) {
    p.ok = true;
    p.pix = py * g.W + px;
    return p;
  }
  p.ok = inr && u > -0.5f && u < g.w_lim && v > -0.5f && v < g.h_lim;
  // select, not branch: the rounding of an out-of-image u/v is harmless, and a branch here
  // would sit between the four gathers of a run
  if constexpr (SELECT) {
    int pix = py * g.W + px, okm = p.ok ? -1 : 0;
    asm("and.b32 %0, %1, %2;" : "=r"(p.pix) : "r"(pix), "r"(okm));
  } else {
    p.pix = p.ok ? py * g.W + px : 0;
  }
  return p;
}

template <bool COLOR, bool LE_TRUNC, bool MASK>
__device__ __forceinline__ Hit sample_exact(const FrameP& f, const GridP& g, const uint8_t* mask, const Proj& p,
                                            float camz, const Stage* st = nullptr) {
  Hit h;
  int sidx = 0;
  if (st != nullptr && st->on && p.ok) sidx = (p.py - st->y0) * kStageW + (p.px - st->x0);
  const float d = fetch<COLOR, MASK>(f, mask, p.pix, h, st, sidx);
  bool ok = p.ok && !(d <= 0.0f);                         // :46
  const float diff = d - camz;                            // :48
  if constexpr (LE_TRUNC) ok = ok && !(diff <= -g.trunc); // :49
  else                    ok = ok && !(diff < -g.trunc);
  h.dist = fminf(1.0f, em::div_rn(diff, g.trunc, g.inv_trunc));  // :53
  h.ok = ok;
  return h;
}

// ---------------------------------------------------------------------------------------
// Conservative segment cull.  Along the fast axis every fp32 operation of the camera
// transform (:27-38) is a correctly rounded, hence weakly monotone, function of the voxel
// index, so the COMPUTED camera coordinates of every voxel of a segment lie in the box
// spanned by the computed coordinates of its two end voxels.  Correctly rounded division,
// multiplication by fx > 0 and addition are monotone too, so evaluating :41-42 at the right
// box corner bounds every voxel's computed u / v from above or below — in fp32, with the
// voxels' own roundings, no error analysis needed.  Returns true only if EVERY voxel of the
// box is skipped by one of :39 (behind the camera), :43 (outside the image) or :49 (farther
// than the frame's largest depth plus the truncation).  Never true on NaN/Inf input.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ bool box_is_skipped(const FrameP& f, const GridP& g, float xa, float ya, float za,
                                               float xb, float yb, float zb, const int* __restrict__ tail,
                                               int* rect = nullptr /* out: pxl, pxh, pyl, pyh when known */) {
  const float big = 3.0e38f;
  const bool fin = fabsf(xa) <= big && fabsf(ya) <= big && fabsf(za) <= big &&
                   fabsf(xb) <= big && fabsf(yb) <= big && fabsf(zb) <= big;
  if (!fin || !(f.fx > 0.0f) || !(f.fy > 0.0f)) return false;
  const float xlo = fminf(xa, xb), xhi = fmaxf(xa, xb);
  const float ylo = fminf(ya, yb), yhi = fmaxf(ya, yb);
  const float zlo = fminf(za, zb), zhi = fmaxf(za, zb);
  if (zhi <= 0.0f) return true;                                   // :39 for every voxel
  if (!(zlo > 0.0f)) return false;                                // straddles the camera plane
  const float dmax = __int_as_float(__ldg(tail));                 // largest depth of the frame
  if (dmax - zlo < -g.trunc) return true;                         // :49: d - z <= dmax - zlo
  // the four bounds below need RN quotients; inside the exponent window the shared-reciprocal
  // form is the IEEE quotient (exactmath.cuh) at a fraction of the cost of eight `/`
  if (!em::in_safe_range(zlo) || !em::in_safe_range(zhi)) return false;
  auto num_ok = [](float x) { return x == 0.0f || em::in_safe_range(fabsf(x)); };  // no under/overflow in div_rn
  if (!(num_ok(xlo) && num_ok(xhi) && num_ok(ylo) && num_ok(yhi))) return false;
  const float rlo = em::rcp_rn(zlo), rhi = em::rcp_rn(zhi);
  const bool a0 = xhi >= 0.0f, a1 = xlo >= 0.0f, a2 = yhi >= 0.0f, a3 = ylo >= 0.0f;
  const float umax = f.fx * em::div_rn(xhi, a0 ? zlo : zhi, a0 ? rlo : rhi) + f.cx;
  const float umin = f.fx * em::div_rn(xlo, a1 ? zhi : zlo, a1 ? rhi : rlo) + f.cx;
  const float vmax = f.fy * em::div_rn(yhi, a2 ? zlo : zhi, a2 ? rlo : rhi) + f.cy;
  const float vmin = f.fy * em::div_rn(ylo, a3 ? zhi : zlo, a3 ? rhi : rlo) + f.cy;
  // :43 left of / right of / above / below the image
  if (umax <= -0.5f || umin >= g.w_lim || vmax <= -0.5f || vmin >= g.h_lim) return true;
  // Occlusion / invalid-depth test on the max-depth tiles of k_frame_prep.  Every voxel of the
  // box that passes :43 reads a pixel (px, py) = (floor(u + 1/2), floor(v + 1/2)) with u in
  // [umin, umax], v in [vmin, vmax]; the pixel rectangle below contains them all (directed
  // roundings), so the largest depth over the tiles it touches bounds every depth the box can
  // see.  0 there = only invalid pixels (:46); otherwise :49 as for the whole-frame maximum.
  const int pxl = max(0, (int)floorf(__fadd_rd(fmaxf(umin, -0.5f), 0.5f)));
  const int pxh = min(g.W - 1, (int)floorf(__fadd_ru(fminf(umax, g.w_lim), 0.5f)));
  const int pyl = max(0, (int)floorf(__fadd_rd(fmaxf(vmin, -0.5f), 0.5f)));
  const int pyh = min(g.H - 1, (int)floorf(__fadd_ru(fminf(vmax, g.h_lim), 0.5f)));
  if (rect) { rect[0] = pxl; rect[1] = pxh; rect[2] = pyl; rect[3] = pyh; }
  // (batched instances: the mask is 0 outside the instance's box, i.e. depth * 0 = invalid, :46)
  if (pxh < f.mb_x0 || pxl > f.mb_x1 || pyh < f.mb_y0 || pyl > f.mb_y1) return true;
  const int tw0 = tiles_across(g.W, kTile0Log2), th0 = tiles_across(g.H, kTile0Log2);
  const int* tiles = tail + kTailHeadWords;
  int sh = kTile0Log2, tw = tw0;
  if ((pxh >> sh) - (pxl >> sh) > 2 || (pyh >> sh) - (pyl >> sh) > 2) {
    tiles += tw0 * th0;
    sh = kTile1Log2;
    tw = tiles_across(g.W, kTile1Log2);
    if ((pxh >> sh) - (pxl >> sh) > 2 || (pyh >> sh) - (pyl >> sh) > 2) return false;  // too wide: no local bound
  }
  int m = 0;
  for (int ty = pyl >> sh; ty <= (pyh >> sh); ++ty)
    for (int tx = pxl >> sh; tx <= (pxh >> sh); ++tx) m = max(m, __ldg(tiles + ty * tw + tx));
  if (m == 0) return true;                                        // :46 nothing valid to hit
  return __int_as_float(m) - zlo < -g.trunc;                      // :49 behind everything it can hit
}

// fast-axis terms of the camera transform of voxel index k (global on z), and their
// combination with the row terms — the ONE spelling used by the hot loop and by the cull
template <bool ZFAST>
__device__ __forceinline__ float fast_coord(const GridP& g, int k) {
  if constexpr (ZFAST) return g.oz + (float)(g.z0 + k) * g.vs;    // :29
  else                 return g.ox + (float)k * g.vs;              // :27
}
template <bool ZFAST>
__device__ __forceinline__ void fast_terms(const float* c, float bf, float& p0, float& p1, float& p2) {
  const float tt = bf - (ZFAST ? c[11] : c[3]);                    // :35 / :33
  p0 = (ZFAST ? c[8] : c[0]) * tt;
  p1 = (ZFAST ? c[9] : c[1]) * tt;
  p2 = (ZFAST ? c[10] : c[2]) * tt;
}
template <bool ZFAST>
__device__ __forceinline__ void cam_coords(const float4& ea, const float4& eb, float p0, float p1, float p2,
                                           float& camx, float& camy, float& camz) {
  if constexpr (ZFAST) {
    camx = ea.x + p0;                         // :36
    camy = ea.y + p1;                         // :37
    camz = ea.z + p2;                         // :38
  } else {
    camx = (p0 + ea.x) + eb.x;
    camy = (p1 + ea.y) + eb.y;
    camz = (p2 + ea.z) + eb.z;
  }
}

// Row terms of the camera transform of row (slow, mid) — the ONE spelling used by the row table
// and by the block-level hull test.  ZFAST: e0 = {A0,A1,A2}; !ZFAST: e0 = {Y0,Y1,Y2}, e1 = {Z0,Z1,Z2}.
template <bool ZFAST>
__device__ __forceinline__ void row_terms(const GridP& g, const float* c, int slow, int mid, float4& e0, float4& e1) {
  const float by = g.oy + (float)mid * g.vs;                       // :28
  const float t1 = by - c[7];                                      // :34
  if constexpr (ZFAST) {
    const float bx = g.ox + (float)slow * g.vs;                    // :27
    const float t0 = bx - c[3];                                    // :33
    e0 = make_float4(c[0] * t0 + c[4] * t1, c[1] * t0 + c[5] * t1, c[2] * t0 + c[6] * t1, 0.0f);
  } else {
    const float bz = g.oz + (float)(g.z0 + slow) * g.vs;           // :29 (global z)
    const float t2 = bz - c[11];                                   // :35
    e0 = make_float4(c[4] * t1, c[5] * t1, c[6] * t1, 0.0f);
    e1 = make_float4(c[8] * t2, c[9] * t2, c[10] * t2, 0.0f);
  }
}

// One voxel update, src/tsdf.cu:54-57 (+ colour).  WMODE: 0 general, 1 integer weights.
template <bool COLOR, int MATH, int WMODE>
__device__ __forceinline__ void update_voxel(float& t, float& w, float& col, const Hit& h, float obs) {
  const float w0 = w;                                   // :54
  const float w1 = w0 + obs;                            // :55
  w = w1;                                               // :56
  if constexpr (MATH == MATH_LITERAL) {
    t = (t * w0 + h.dist) / w1;                         // :57
    if constexpr (COLOR) col = em::blend_literal(col, h.c0, h.c1, h.c2, w0, w1);
  } else {
    const float yw = em::rcp_rn(w1);
    t = em::div_rn(t * w0 + h.dist, w1, yw);            // :57
    if constexpr (COLOR) {
      if constexpr (WMODE == 1) col = em::blend_intw(col, h.c0, h.c1, h.c2, w0, yw);
      else                      col = em::blend_exact(col, h.c0, h.c1, h.c2, w0, w1, yw);
    }
  }
}

// ZFAST : slab layout [x][y][zl]  (row = x*Dy + y, fast = zl)   — plugin surface
// !ZFAST: slab layout [zl][y][x]  (row = zl*Dy + y, fast = x)   — class TSDF, src/tsdf.cu:52
// Row table entry per frame (padded to 128-bit loads):
//   ZFAST  {A0,A1,A2,-}                A_k = c[k]*t0 + c[4+k]*t1
//   !ZFAST {Y0,Y1,Y2,-, Z0,Z1,Z2,-}    Y_k = c[4+k]*t1, Z_k = c[8+k]*t2
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC, bool MASK, int MATH, int WMODE, int CULL, bool STAGE = false>
__device__ __forceinline__ void integrate_body(const LaunchP<NF>& P, const uint8_t* mask, float4* tab,
                                               int tpr_log2, int rows_per_block, int flags,
                                               const void* tmap = nullptr, float4* stage_tile = nullptr) {
  constexpr int NT = ZFAST ? 1 : 2;  // float4 per row per frame
  // CULL: 0 dense (no segment cull, branch-free run), 3 dense + host-proved inside the image (project_exact
  // INSIDE), 2 sparse (segment cull), 1 legacy one-piece probe (A/B)
  constexpr bool kDense = CULL == 0 || CULL == 3;
  constexpr bool kInside = CULL == 3;
  const GridP& g = P.g;
  const int pf = flags & 3;            // software prefetch: 0 off, 1 L2, 2 L1
  const int n_rows = g.n_mid * g.n_slow;
  const int row0 = blockIdx.x * rows_per_block;
  if (row0 >= n_rows) return;  // batched launches size the grid for the largest instance
  // a peer-pull wait of this frame timed out (comm.cu): the prepared frame is stale, leave the
  // volume untouched.  Grid-uniform; one L2-resident word.
  if (P.f[0].skip != nullptr && *reinterpret_cast<const volatile int*>(P.f[0].skip) != 0) return;

  // ---- prologue 0 (sparse kernel): can ANY voxel of this block be touched? ----------------
  // The computed camera coordinates of voxel k of row (slow, mid) are e(slow, mid) + p_k (ZFAST)
  // or (p_k + y(mid)) + z(slow): every fp32 operation on the way is correctly rounded, hence
  // weakly monotone in each operand, and the direction in which a coordinate moves with slow,
  // mid or k is fixed by the sign of one pose coefficient.  So over the index box
  // [s0, s1] x [mlo, mhi] x [a, b] that contains the block's voxels, each computed coordinate lies
  // between the values the SAME operation sequence gives at the box's corners — in fp32, no error
  // analysis.  One box_is_skipped on that hull, by warp 0 while the other warps build the row
  // table (the kernel is issue-bound on sparse frames: 8 warps repeating the test cost 17 % of all
  // warp instructions on S-kitti), lets a block outside the frustum, behind the camera or behind
  // everything the frame sees leave at the row-table barrier.
  __shared__ int blk_skip;
  __shared__ int stage_xy[3];                      // on, x0, y0
  __shared__ __align__(8) unsigned long long stage_bar;
  const bool blk_test = CULL == 2 && NF == 1 && (flags & 4);
  if constexpr (STAGE) { if (threadIdx.x == 0) stage_xy[0] = 0; }
  if constexpr (CULL == 2 && NF == 1) {
    if (blk_test && threadIdx.x < 32) {
      const int r_last = min(row0 + rows_per_block, n_rows) - 1;
      const int s0 = row0 / g.n_mid, s1 = r_last / g.n_mid;
      const int mlo = s0 == s1 ? row0 - s0 * g.n_mid : 0;
      const int mhi = s0 == s1 ? r_last - s0 * g.n_mid : g.n_mid - 1;
      const float* c = P.f[0].c;
      float elo[3], ehi[3], zlo[3] = {0.f, 0.f, 0.f}, zhi[3] = {0.f, 0.f, 0.f};
      float4 e[2], z4[2];
      if constexpr (ZFAST) {
        float4 q[4];
        row_terms<ZFAST>(g, c, s0, mlo, q[0], z4[0]);
        row_terms<ZFAST>(g, c, s0, mhi, q[1], z4[0]);
        row_terms<ZFAST>(g, c, s1, mlo, q[2], z4[0]);
        row_terms<ZFAST>(g, c, s1, mhi, q[3], z4[0]);
        e[0] = make_float4(fminf(fminf(q[0].x, q[1].x), fminf(q[2].x, q[3].x)), fminf(fminf(q[0].y, q[1].y), fminf(q[2].y, q[3].y)),
                           fminf(fminf(q[0].z, q[1].z), fminf(q[2].z, q[3].z)), 0.f);
        e[1] = make_float4(fmaxf(fmaxf(q[0].x, q[1].x), fmaxf(q[2].x, q[3].x)), fmaxf(fmaxf(q[0].y, q[1].y), fmaxf(q[2].y, q[3].y)),
                           fmaxf(fmaxf(q[0].z, q[1].z), fmaxf(q[2].z, q[3].z)), 0.f);
        // a non-finite corner must disable the test (fminf / fmaxf drop NaN operands)
        const float chk = ((q[0].x + q[1].x) + (q[2].x + q[3].x)) + ((q[0].y + q[1].y) + (q[2].y + q[3].y)) +
                          ((q[0].z + q[1].z) + (q[2].z + q[3].z));
        if (!(fabsf(chk) <= 3.0e38f)) e[0].x = __int_as_float(0x7fc00000);
      } else {
        float4 ya, yb, za, zb;
        row_terms<ZFAST>(g, c, s0, mlo, ya, za);
        row_terms<ZFAST>(g, c, s1, mhi, yb, zb);
        e[0] = make_float4(fminf(ya.x, yb.x), fminf(ya.y, yb.y), fminf(ya.z, yb.z), 0.f);
        e[1] = make_float4(fmaxf(ya.x, yb.x), fmaxf(ya.y, yb.y), fmaxf(ya.z, yb.z), 0.f);
        z4[0] = make_float4(fminf(za.x, zb.x), fminf(za.y, zb.y), fminf(za.z, zb.z), 0.f);
        z4[1] = make_float4(fmaxf(za.x, zb.x), fmaxf(za.y, zb.y), fmaxf(za.z, zb.z), 0.f);
        const float chk = ((ya.x + yb.x) + (ya.y + yb.y)) + ((ya.z + yb.z) + (za.x + zb.x)) + ((za.y + zb.y) + (za.z + zb.z));
        if (!(fabsf(chk) <= 3.0e38f)) e[0].x = __int_as_float(0x7fc00000);
        zlo[0] = z4[0].x; zlo[1] = z4[0].y; zlo[2] = z4[0].z;
        zhi[0] = z4[1].x; zhi[1] = z4[1].y; zhi[2] = z4[1].z;
      }
      elo[0] = e[0].x; elo[1] = e[0].y; elo[2] = e[0].z;
      ehi[0] = e[1].x; ehi[1] = e[1].y; ehi[2] = e[1].z;
      const int tprb = 1 << tpr_log2;
      const int a = blockIdx.y * tprb * VEC;
      int b = a + tprb * VEC - 1;
      if (b > g.n_fast - 1) b = g.n_fast - 1;
      float pa[3], pb[3], lo[3], hi[3];
      fast_terms<ZFAST>(c, fast_coord<ZFAST>(g, a), pa[0], pa[1], pa[2]);
      fast_terms<ZFAST>(c, fast_coord<ZFAST>(g, b), pb[0], pb[1], pb[2]);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float plo = fminf(pa[k], pb[k]), phi = fmaxf(pa[k], pb[k]);
        if constexpr (ZFAST) {
          lo[k] = elo[k] + plo;
          hi[k] = ehi[k] + phi;
        } else {
          lo[k] = (plo + elo[k]) + zlo[k];
          hi[k] = (phi + ehi[k]) + zhi[k];
        }
      }
      // (a NaN planted above, or produced by the additions, fails box_is_skipped's finiteness test)
      int rect[4] = {0, -1, 0, -1};
      const bool skip = box_is_skipped(P.f[0], g, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                                       reinterpret_cast<const int*>(P.f[0].dmax), STAGE ? rect : nullptr);
      if (threadIdx.x == 0) blk_skip = skip;
      if constexpr (STAGE) {
        // every pixel a voxel of this block can read lies in rect: if it fits one box, fetch the box
        if (threadIdx.x == 0 && !skip && rect[1] >= rect[0] && rect[3] >= rect[2] &&
            rect[1] - rect[0] < kStageW && rect[3] - rect[2] < kStageH) {
          const unsigned bar = smem_u32(&stage_bar), dst = smem_u32(stage_tile);
          asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
          asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kStageW * kStageH * 16) : "memory");
          asm volatile(
              "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
              ::"r"(dst), "l"(tmap), "r"(4 * rect[0]), "r"(rect[2]), "r"(bar) : "memory");
          stage_xy[0] = 1; stage_xy[1] = rect[0]; stage_xy[2] = rect[2];
        }
      }
    }
  }

  // ---- prologue 1: per-row terms of the camera transform -> shared memory ---------------
  // (warps 1-7 when warp 0 runs the block test: 224 threads cover the <= 128 rows in one pass)
  const int t_first = blk_test ? 32 : 0;
  for (int lr = (int)threadIdx.x - t_first; lr < rows_per_block; lr += kThreads - t_first) {
    int r = row0 + lr;
    if (lr >= 0 && r < n_rows) {
      int slow = r / g.n_mid, mid = r - slow * g.n_mid;
#pragma unroll
      for (int f = 0; f < NF; ++f) {
        float4* e = tab + (lr * NF + f) * NT;
        float4 e0, e1;
        row_terms<ZFAST>(g, P.f[f].c, slow, mid, e0, e1);
        e[0] = e0;
        if constexpr (!ZFAST) e[1] = e1;
      }
    }
  }
  __syncthreads();
#ifdef TSDF_B200_DEV
#define DBG_ADD(k, n) do { if (g.dbg) atomicAdd(g.dbg + (k), (unsigned long long)(n)); } while (0)
#else
#define DBG_ADD(k, n) do { } while (0)
#endif
  if constexpr (CULL == 2 && NF == 1) {
    if (threadIdx.x == 0) { DBG_ADD(0, 1); if (blk_test && blk_skip) DBG_ADD(1, 1); }
    if (blk_test && blk_skip) return;   // block-uniform
  }
  Stage stg;
  stg.tile = stage_tile; stg.on = 0; stg.x0 = 0; stg.y0 = 0;
  if constexpr (STAGE) {
    if (stage_xy[0]) {   // block-uniform: wait for the box (phase 0 of the barrier)
      const unsigned bar = smem_u32(&stage_bar);
      unsigned done = 0;
      while (!done) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar) : "memory");
      }
      stg.on = 1; stg.x0 = stage_xy[1]; stg.y0 = stage_xy[2];
    }
  }

  // ---- this thread's run on the fast axis -----------------------------------------------
  const int tpr = 1 << tpr_log2;
  const int tf = threadIdx.x & (tpr - 1);
  const int tr = threadIdx.x >> tpr_log2;
  const int rp = kThreads >> tpr_log2;  // rows processed in parallel by the block
  const int f0 = (blockIdx.y * tpr + tf) * VEC;

  // ---- prologue 2: which of this warp's row segments no frame can touch -------------------
  // A warp works on one segment of the fast axis (32 runs = 128 voxels; the whole row window
  // when a row has fewer than 32 runs, several rows per warp then) for every row of its row
  // sequence lr = tr, tr+rp, ...  Lane k of a row group tests iteration k (k + group size, ...)
  // with box_is_skipped; one ballot per pass hands every lane the bit mask of its own
  // sequence.  No shared memory, no block barrier: warps start their rows independently.
  unsigned long long rowmask = 0ull;  // bit i: iteration i of this thread's row loop is culled
  if constexpr (!kDense) {
    const int tw_log2 = tpr_log2 < 5 ? tpr_log2 : 5;
    const int tw = 1 << tw_log2;                       // lanes per row group inside the warp
    const int lane = threadIdx.x & 31;
    const int kk = lane & (tw - 1);
    const int gshift = lane & ~(tw - 1);
    const unsigned gmask = tw == 32 ? 0xffffffffu : ((1u << tw) - 1u);
    const int n_iter = (rows_per_block - tr + rp - 1) / rp;   // same for every tr (powers of two)
    const int ka = (blockIdx.y * tpr + (tf & ~31)) * VEC;
    int kb = ka + tw * VEC - 1;
    if (kb > g.n_fast - 1) kb = g.n_fast - 1;
    // box test of fast-axis voxels [a, b] of local row lr against every frame
    auto seg_skipped = [&](int lr, int a, int b) {
      const float bfa = fast_coord<ZFAST>(g, a), bfb = fast_coord<ZFAST>(g, b);
      bool skip = true;
#pragma unroll
      for (int f = 0; f < NF; ++f) {
        const float4* e = tab + (lr * NF + f) * NT;
        const float4 ea = e[0];
        float4 eb = ea;
        if constexpr (!ZFAST) eb = e[1];
        float p0, p1, p2, xa, ya, za, xb, yb, zb;
        fast_terms<ZFAST>(P.f[f].c, bfa, p0, p1, p2);
        cam_coords<ZFAST>(ea, eb, p0, p1, p2, xa, ya, za);
        fast_terms<ZFAST>(P.f[f].c, bfb, p0, p1, p2);
        cam_coords<ZFAST>(ea, eb, p0, p1, p2, xb, yb, zb);
        skip = skip && box_is_skipped(P.f[f], g, xa, ya, za, xb, yb, zb, reinterpret_cast<const int*>(P.f[f].dmax));
      }
      return skip;
    };
    if (tw == 32 && n_iter <= 64) {
      // Usual case (one segment per warp): all 32 lanes work — in each round of 16 row iterations
      // lane L tests two quarter-segments (8 runs each) of row (L & 15), half (L >> 4); a row
      // segment is culled iff its four quarter boxes are, which is tighter than one box around it.
      const int half = lane >> 4;
      for (int base = 0; base < n_iter; base += 16) {
        const int it = base + (lane & 15);
        const int lr = tr + it * rp;
        bool skip = false;
        if (it < n_iter && row0 + lr < n_rows && ka <= kb) {
          skip = true;
          const int qlen = 8 * VEC;
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int qa = ka + (2 * half + q) * qlen;
            const int qb = qa + qlen - 1 < kb ? qa + qlen - 1 : kb;
            if (qa <= kb) skip = skip && seg_skipped(lr, qa, qb);
          }
        }
        const unsigned b = __ballot_sync(0xffffffffu, skip);
        rowmask |= (unsigned long long)(b & (b >> 16) & 0xffffu) << base;
      }
    } else {
      for (int pass = 0; (pass << tw_log2) < n_iter; ++pass) {
        const int it = (pass << tw_log2) + kk;
        const int lr = tr + it * rp;
        const bool skip = it < n_iter && row0 + lr < n_rows && ka <= kb && seg_skipped(lr, ka, kb);
        const unsigned b = __ballot_sync(0xffffffffu, skip);
        rowmask |= (unsigned long long)((b >> gshift) & gmask) << (pass << tw_log2);
      }
    }
  }
  // A thread whose run lies beyond the row (ragged fast axis) does no row work but stays: the
  // closing warp reduction of n_written names every lane in its mask.
  const bool live = f0 < g.n_fast;

  // per-run terms: the products of the fast-axis coordinate (registers, frame 0 only when
  // several frames are fused: the others are recomputed per row to bound register use)
  float bf[VEC];
  float q0[VEC], q1[VEC], q2[VEC];
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    bf[j] = fast_coord<ZFAST>(g, f0 + j);
    fast_terms<ZFAST>(P.f[0].c, bf[j], q0[j], q1[j], q2[j]);
  }

  unsigned int n_written = 0;
  // Software prefetch of the NEXT row's three 128-bit runs: the volume loads can only be
  // issued after the depth gather has decided validity, so without it every row exposes a
  // full DRAM latency (profiles/r1_v2: 45 % of stall samples on the first use of the loaded
  // weight).  A row is prefetched when the row before it was updated (coherent frames), and
  // the first row of a block unless it is culled.
  bool any_prev = true;
  // running pointers to this thread's run in the current row (the row stride is block-uniform): the
  // round-1 loop rebuilt three 64-bit addresses per row for the loads and three more per prefetch,
  // ~45 of the ~470 instructions of a row iteration
  const int64_t step = (int64_t)rp * g.n_fast;
  const int64_t off0 = (int64_t)(row0 + tr) * g.n_fast + f0;
  float* pt = g.tsdf + off0;
  float* pw = g.weight + off0;
  float* pc = COLOR ? g.color + off0 : nullptr;
  if (pf && live && row0 + tr < n_rows && !(!kDense && (rowmask & 1ull))) {   // first row of the block
    prefetch_line<CULL != 1>(pt, pf);
    prefetch_line<CULL != 1>(pw, pf);
    if constexpr (COLOR) prefetch_line<CULL != 1>(pc, pf);
  }
  for (int lr = tr; live && lr < rows_per_block; lr += rp, rowmask >>= 1, pt += step, pw += step, pc += COLOR ? step : 0) {
    const int r = row0 + lr;
    if (r >= n_rows) break;
    if constexpr (!kDense) {
      if ((threadIdx.x & 31) == 0) { DBG_ADD(2, 1); if (rowmask & 1ull) DBG_ADD(3, 1); }
      if (rowmask & 1ull) { any_prev = false; continue; }
    }
    if (pf && any_prev && lr + rp < rows_per_block && r + rp < n_rows) {
      prefetch_line<CULL != 1>(pt + step, pf);
      prefetch_line<CULL != 1>(pw + step, pf);
      if constexpr (COLOR) prefetch_line<CULL != 1>(pc + step, pf);
    }
    Hit hit[NF][VEC];
    bool any = false;
#pragma unroll
    for (int f = 0; f < NF; ++f) {
      const float4* e = tab + (lr * NF + f) * NT;
      const float4 ea = e[0];
      float4 eb = ea;
      if constexpr (!ZFAST) eb = e[1];
      float camx[VEC], camy[VEC], camz[VEC];
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        float p0, p1, p2;
        if (f == 0) { p0 = q0[j]; p1 = q1[j]; p2 = q2[j]; }
        else fast_terms<ZFAST>(P.f[f].c, bf[j], p0, p1, p2);
        cam_coords<ZFAST>(ea, eb, p0, p1, p2, camx[j], camy[j], camz[j]);
      }
      if constexpr (MATH == MATH_LITERAL) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          hit[f][j] = probe_literal_body<COLOR, LE_TRUNC, MASK>(P.f[f], g, mask, camx[j], camy[j], camz[j]);
          any = any || hit[f][j].ok;
        }
      } else if constexpr (CULL == 1) {  // legacy one-piece probe (A/B baseline, variant bit 10)
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          hit[f][j] = probe_exact<COLOR, LE_TRUNC, MASK>(P.f[f], g, mask, camx[j], camy[j], camz[j]);
          any = any || hit[f][j].ok;
        }
      } else {
        constexpr bool kEnds = kRunSafe && kDense && VEC == 4;
        bool odd = false;  // a positive camz outside the exponent window of the exact forms
        if constexpr (kEnds) {
          odd = !(em::in_safe_range(camz[0]) && em::in_safe_range(camz[VEC - 1]));   // (dense frames: never)
        } else {
#pragma unroll
          for (int j = 0; j < VEC; ++j) odd = odd || (!em::in_safe_range(camz[j]) && camz[j] > 0.0f);
        }
        if (__builtin_expect(odd, 0)) {
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            hit[f][j] = probe_literal_cold<COLOR, LE_TRUNC, MASK>(P.f[f].img, P.f[f].fx, P.f[f].cx, P.f[f].fy, P.f[f].cy,
                                                                  g.trunc, g.W, g.H, mask, camx[j], camy[j], camz[j]);
            any = any || hit[f][j].ok;
          }
        } else {
          Proj pr[VEC];
          bool in_image = false;
#pragma unroll
          for (int j = 0; j < VEC; ++j) {
            pr[j] = project_exact<true, kEnds, kInside>(P.f[f], g, camx[j], camy[j], camz[j]);
            in_image = in_image || pr[j].ok;
          }
          // CULL == 2 (sparse frames): a run that projects outside the image stops before the
          // gather.  CULL == 0 (dense frames): no early-out, the run is branch-free from here, so
          // the four gathers issue back to back.
          if constexpr (CULL == 2) { DBG_ADD(4, 1); if (in_image) DBG_ADD(5, 1); }
          if (kDense || in_image) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
              hit[f][j] = sample_exact<COLOR, LE_TRUNC, MASK>(P.f[f], g, mask, pr[j], camz[j], STAGE ? &stg : nullptr);
              any = any || hit[f][j].ok;
            }
          } else {
#pragma unroll
            for (int j = 0; j < VEC; ++j) hit[f][j].ok = false;
          }
        }
      }
    }
    any_prev = any;
    if constexpr (CULL == 2) { if (any) DBG_ADD(6, 1); }
    if (!any) continue;  // no volume traffic for a run with no update

    float t[VEC], w[VEC], col[VEC];
    load_vec<VEC>(pt, t);
    load_vec<VEC>(pw, w);
    if constexpr (COLOR) load_vec<VEC>(pc, col);
    else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) col[j] = 0.0f;
    }
#pragma unroll
    for (int f = 0; f < NF; ++f) {
      const float obs = P.f[f].obs_weight;
      // Dense frames (CULL == 0): nearly every voxel of a written run is updated, so the update
      // is computed unconditionally and selected — no divergence bookkeeping around 60
      // instructions.  Sparse frames keep the branch: measured (S-real / S-kitti, round-robin)
      // the select form costs 7-16 % there, and a per-run hybrid of the two 6 %.
      // (A per-run restructuring of this update with exact free-space short cuts — skip :53's quotient
      // when diff >= trunc for the whole run, skip :57's when the stored tsdf is 1 as well — was built
      // and measured in round 2: 8 % SLOWER at full clocks, profiles/r2_v3_dense_update_ab.jsonl; the
      // kernel is HBM-bound and the extra control flow costs more than the arithmetic it saves.)
      if constexpr (kDense) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          float nt = t[j], nw = w[j], nc = col[j];
          update_voxel<COLOR, MATH, WMODE>(nt, nw, nc, hit[f][j], obs);
          const bool ok = hit[f][j].ok;
          t[j] = ok ? nt : t[j];
          w[j] = ok ? nw : w[j];
          col[j] = ok ? nc : col[j];
          n_written += ok ? 1u : 0u;
        }
      } else {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
          if (hit[f][j].ok) {
            update_voxel<COLOR, MATH, WMODE>(t[j], w[j], col[j], hit[f][j], obs);
            ++n_written;
          }
        }
      }
    }
    store_vec<VEC>(pt, t);
    store_vec<VEC>(pw, w);
    if constexpr (COLOR) store_vec<VEC>(pc, col);
  }

  // exact N_written: warp-reduce, one atomic per warp
  for (int o = 16; o > 0; o >>= 1) n_written += __shfl_down_sync(0xffffffffu, n_written, o);
  if ((threadIdx.x & 31) == 0 && n_written) atomicAdd(g.written, (unsigned long long)n_written);
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC, int MATH, int WMODE, int CULL>
__global__ void __launch_bounds__(kThreads, KI_MINBLOCKS)   // 64 registers: four blocks per SM keep the HBM stream fed
k_integrate(const __grid_constant__ LaunchP<NF> P, int tpr_log2, int rows_per_block, int flags) {
  __shared__ float4 tab[kMaxRowsPerBlock * NF * (ZFAST ? 1 : 2)];
  integrate_body<ZFAST, COLOR, LE_TRUNC, NF, VEC, false, MATH, WMODE, CULL>(P, nullptr, tab, tpr_log2, rows_per_block, flags);
}

// ---------------------------------------------------------------------------------------
// Sparse-frame kernel with QUARTER-SEGMENT work lists (default for sparse frames on compact tiles).
// The plain sparse kernel culls a 128-voxel row segment only when all four of its 32-voxel quarter
// boxes are culled, and then probes the whole segment: measured on the C2 / C3 shapes
// (tools/sparse_counters.py, profiles/r2_v3_sparse_counters.jsonl) only 37 % / 16 % of the runs it
// projects reach the gather, because segments straddling the frustum, an occluder or the depth
// range drag their dead quarters along — and the kernel is issue-bound, so dead lanes cost what
// live ones do.  Here the quarter tests' ballots become a per-warp WORK LIST of live (row, quarter)
// items (64 bits: 16 rows x 4 quarters), and every warp iteration takes FOUR items, one per group of
// 8 lanes — 8 lanes x 4 voxels = one quarter, still 128-bit accesses and whole 128-byte lines.  A
// lane's fast-axis position now changes per item, so the fast-axis terms of the block's 128-voxel
// window live in a shared-memory table next to the row table (one LDS.128 per voxel).
// Same tests, same per-voxel arithmetic, same bits as the plain kernels (tests/test_cull_gpu.py).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int nth_set_bit64(unsigned long long m, int n) {   // n-th (0-based) set bit, n < popc(m)
  const unsigned lo = (unsigned)m, hi = (unsigned)(m >> 32);
  const int pl = __popc(lo);
  return n < pl ? (int)__fns(lo, 0, n + 1) : 32 + (int)__fns(hi, 0, n - pl + 1);
}

constexpr int kSegQ = 128;                           // voxels of a block's fast-axis window (32 runs of 4)

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE, bool MASK>
__device__ __forceinline__ void integrate_q_body(const LaunchP<1>& P, const uint8_t* mask, float4* tab, float4* tabf,
                                                 int rows_per_block, int flags) {
  constexpr int NT = ZFAST ? 1 : 2;
  constexpr int VEC = 4;
  constexpr int kSeg = kSegQ;
  __shared__ int blk_skip;
  const GridP& g = P.g;
  const FrameP& F = P.f[0];
  const int pf = flags & 3;
  const int n_rows = g.n_mid * g.n_slow;
  const int row0 = blockIdx.x * rows_per_block;
  if (row0 >= n_rows) return;
  if (F.skip != nullptr && *reinterpret_cast<const volatile int*>(F.skip) != 0) return;
  const int ka = blockIdx.y * kSeg;                  // first voxel of the window
  int kb = ka + kSeg - 1;
  if (kb > g.n_fast - 1) kb = g.n_fast - 1;
  const float* c = F.c;

  // ---- prologue 0: block-level hull test (warp 0), as in integrate_body ------------------------
  const bool blk_test = (flags & 4) != 0;
  if (blk_test && threadIdx.x < 32) {
    const int r_last = min(row0 + rows_per_block, n_rows) - 1;
    const int s0 = row0 / g.n_mid, s1 = r_last / g.n_mid;
    const int mlo = s0 == s1 ? row0 - s0 * g.n_mid : 0;
    const int mhi = s0 == s1 ? r_last - s0 * g.n_mid : g.n_mid - 1;
    float lo[3], hi[3];
    float4 q[4], z4[2];
    float elo[3], ehi[3], zlo[3] = {0.f, 0.f, 0.f}, zhi[3] = {0.f, 0.f, 0.f};
    bool finite = true;
    if constexpr (ZFAST) {
      row_terms<ZFAST>(g, c, s0, mlo, q[0], z4[0]);
      row_terms<ZFAST>(g, c, s0, mhi, q[1], z4[0]);
      row_terms<ZFAST>(g, c, s1, mlo, q[2], z4[0]);
      row_terms<ZFAST>(g, c, s1, mhi, q[3], z4[0]);
      elo[0] = fminf(fminf(q[0].x, q[1].x), fminf(q[2].x, q[3].x)); ehi[0] = fmaxf(fmaxf(q[0].x, q[1].x), fmaxf(q[2].x, q[3].x));
      elo[1] = fminf(fminf(q[0].y, q[1].y), fminf(q[2].y, q[3].y)); ehi[1] = fmaxf(fmaxf(q[0].y, q[1].y), fmaxf(q[2].y, q[3].y));
      elo[2] = fminf(fminf(q[0].z, q[1].z), fminf(q[2].z, q[3].z)); ehi[2] = fmaxf(fmaxf(q[0].z, q[1].z), fmaxf(q[2].z, q[3].z));
      const float chk = ((q[0].x + q[1].x) + (q[2].x + q[3].x)) + ((q[0].y + q[1].y) + (q[2].y + q[3].y)) +
                        ((q[0].z + q[1].z) + (q[2].z + q[3].z));
      finite = fabsf(chk) <= 3.0e38f;   // (fminf / fmaxf drop NaN operands: a non-finite corner disables the test)
    } else {
      float4 ya, yb, za, zb;
      row_terms<ZFAST>(g, c, s0, mlo, ya, za);
      row_terms<ZFAST>(g, c, s1, mhi, yb, zb);
      elo[0] = fminf(ya.x, yb.x); elo[1] = fminf(ya.y, yb.y); elo[2] = fminf(ya.z, yb.z);
      ehi[0] = fmaxf(ya.x, yb.x); ehi[1] = fmaxf(ya.y, yb.y); ehi[2] = fmaxf(ya.z, yb.z);
      zlo[0] = fminf(za.x, zb.x); zlo[1] = fminf(za.y, zb.y); zlo[2] = fminf(za.z, zb.z);
      zhi[0] = fmaxf(za.x, zb.x); zhi[1] = fmaxf(za.y, zb.y); zhi[2] = fmaxf(za.z, zb.z);
      const float chk = ((ya.x + yb.x) + (ya.y + yb.y)) + ((ya.z + yb.z) + (za.x + zb.x)) + ((za.y + zb.y) + (za.z + zb.z));
      finite = fabsf(chk) <= 3.0e38f;
    }
    float pa[3], pb[3];
    fast_terms<ZFAST>(c, fast_coord<ZFAST>(g, ka), pa[0], pa[1], pa[2]);
    fast_terms<ZFAST>(c, fast_coord<ZFAST>(g, kb), pb[0], pb[1], pb[2]);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float plo = fminf(pa[k], pb[k]), phi = fmaxf(pa[k], pb[k]);
      if constexpr (ZFAST) { lo[k] = elo[k] + plo; hi[k] = ehi[k] + phi; }
      else                 { lo[k] = (plo + elo[k]) + zlo[k]; hi[k] = (phi + ehi[k]) + zhi[k]; }
    }
    const bool skip = finite && box_is_skipped(F, g, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                                               reinterpret_cast<const int*>(F.dmax));
    if (threadIdx.x == 0) blk_skip = skip;
  }
  // ---- prologue 1: row table (warps 1-7 while warp 0 tests) and fast-axis table -----------------
  const int t_first = blk_test ? 32 : 0;
  for (int lr = (int)threadIdx.x - t_first; lr < rows_per_block; lr += kThreads - t_first) {
    const int r = row0 + lr;
    if (lr >= 0 && r < n_rows) {
      const int slow = r / g.n_mid, mid = r - slow * g.n_mid;
      float4 e0, e1;
      row_terms<ZFAST>(g, c, slow, mid, e0, e1);
      tab[lr * NT] = e0;
      if constexpr (!ZFAST) tab[lr * NT + 1] = e1;
    }
  }
  if (threadIdx.x >= kThreads - kSeg) {   // the last four warps (not warp 0)
    const int k = (int)threadIdx.x - (kThreads - kSeg);
    float p0, p1, p2;
    fast_terms<ZFAST>(c, fast_coord<ZFAST>(g, ka + k), p0, p1, p2);
    tabf[k] = make_float4(p0, p1, p2, 0.0f);
  }
  __syncthreads();
#ifdef TSDF_B200_DEV
  if (threadIdx.x == 0) { DBG_ADD(0, 1); if (blk_test && blk_skip) DBG_ADD(1, 1); }
#endif
  if (blk_test && blk_skip) return;   // block-uniform

  // ---- prologue 2: the warp's work list -------------------------------------------------------
  const int lane = threadIdx.x & 31;
  const int tr = threadIdx.x >> 5;            // warp = row phase: rows tr, tr + 8, ...
  constexpr int rp = kThreads / 32;
  const int n_iter = (rows_per_block - tr + rp - 1) / rp;   // <= 16
  unsigned long long work = 0ull;             // bit Q * 16 + it: quarter Q of row iteration `it` is live
  {
    const int it = lane & 15, half = lane >> 4;
    const int lr = tr + it * rp;
    const bool rowok = it < n_iter && row0 + lr < n_rows;
    bool live[2] = {false, false};
    if (rowok) {
      const float4 ea = tab[lr * NT];
      float4 eb = ea;
      if constexpr (!ZFAST) eb = tab[lr * NT + 1];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int qa = ka + (2 * half + q) * 32;
        const int qb = qa + 31 < kb ? qa + 31 : kb;
        if (qa <= kb) {
          const float4 fa = tabf[qa - ka], fb = tabf[qb - ka];
          float xa, ya, za, xb, yb, zb;
          cam_coords<ZFAST>(ea, eb, fa.x, fa.y, fa.z, xa, ya, za);
          cam_coords<ZFAST>(ea, eb, fb.x, fb.y, fb.z, xb, yb, zb);
          live[q] = !box_is_skipped(F, g, xa, ya, za, xb, yb, zb, reinterpret_cast<const int*>(F.dmax));
        }
      }
    }
    const unsigned b0 = __ballot_sync(0xffffffffu, live[0]);   // bit L: quarter 2 * (L >> 4)     of row L & 15
    const unsigned b1 = __ballot_sync(0xffffffffu, live[1]);   // bit L: quarter 2 * (L >> 4) + 1 of row L & 15
    work = (unsigned long long)(b0 & 0xffffu) | ((unsigned long long)(b1 & 0xffffu) << 16) |
           ((unsigned long long)(b0 >> 16) << 32) | ((unsigned long long)(b1 >> 16) << 48);
  }
  const int total = __popcll(work);
#ifdef TSDF_B200_DEV
  if (lane == 0) { DBG_ADD(2, 4 * n_iter); DBG_ADD(3, 4 * n_iter - total); }
#endif

  // ---- main loop: four live quarters per warp iteration ------------------------------------------
  // (Prefetching the NEXT item's volume lines one iteration ahead was tried: no gain — S-real 0.1559 vs
  // 0.1567 ms, S-kitti 0.0943 vs 0.0925 ms, profiles/r2_v4_sparse_ab.jsonl — and is not kept.)
  unsigned int n_written = 0;
  const int grp = lane >> 3, sub = lane & 7;
  for (int t0 = 0; t0 < total; t0 += 4) {
    const int item = t0 + grp;
    if (item >= total) continue;
    const int bit = nth_set_bit64(work, item);
    const int Q = bit >> 4, it = bit & 15;
    const int lr = tr + it * rp;
    const int r = row0 + lr;
    const int kf = Q * 32 + sub * VEC;              // offset in the window
    const int f0 = ka + kf;
    if (f0 >= g.n_fast) continue;                   // ragged fast axis (n_fast % 4 == 0: whole runs)
    const float4 ea = tab[lr * NT];
    float4 eb = ea;
    if constexpr (!ZFAST) eb = tab[lr * NT + 1];
    float camx[VEC], camy[VEC], camz[VEC];
    bool odd = false;
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const float4 fp = tabf[kf + j];
      cam_coords<ZFAST>(ea, eb, fp.x, fp.y, fp.z, camx[j], camy[j], camz[j]);
      odd = odd || (!em::in_safe_range(camz[j]) && camz[j] > 0.0f);
    }
    Hit hit[VEC];
    bool any = false;
    if (__builtin_expect(odd, 0)) {
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        hit[j] = probe_literal_cold<COLOR, LE_TRUNC, MASK>(F.img, F.fx, F.cx, F.fy, F.cy, g.trunc, g.W, g.H, mask,
                                                           camx[j], camy[j], camz[j]);
        any = any || hit[j].ok;
      }
    } else {
      Proj pr[VEC];
      bool in_image = false;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        pr[j] = project_exact<true>(F, g, camx[j], camy[j], camz[j]);
        in_image = in_image || pr[j].ok;
      }
      DBG_ADD(4, 1);
      if (!in_image) continue;                      // the run projects outside the image: no gather
      DBG_ADD(5, 1);
      const int64_t pb = (int64_t)r * g.n_fast + f0;
      if (pf) {   // the gather's latency hides part of the volume loads' (most gathered runs are updated)
        prefetch_line<true>(g.tsdf + pb, 1);
        prefetch_line<true>(g.weight + pb, 1);
        if constexpr (COLOR) prefetch_line<true>(g.color + pb, 1);
      }
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        hit[j] = sample_exact<COLOR, LE_TRUNC, MASK>(F, g, mask, pr[j], camz[j]);
        any = any || hit[j].ok;
      }
    }
    if (!any) continue;                             // no volume traffic for a run with no update
    DBG_ADD(6, 1);
    const int64_t base = (int64_t)r * g.n_fast + f0;
    float t[VEC], w[VEC], col[VEC];
    load_vec<VEC>(g.tsdf + base, t);
    load_vec<VEC>(g.weight + base, w);
    if constexpr (COLOR) load_vec<VEC>(g.color + base, col);
    else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) col[j] = 0.0f;
    }
    const float obs = F.obs_weight;
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      if (hit[j].ok) {
        update_voxel<COLOR, MATH_EXACT, WMODE>(t[j], w[j], col[j], hit[j], obs);
        ++n_written;
      }
    }
    store_vec<VEC>(g.tsdf + base, t);
    store_vec<VEC>(g.weight + base, w);
    if constexpr (COLOR) store_vec<VEC>(g.color + base, col);
  }
  for (int o = 16; o > 0; o >>= 1) n_written += __shfl_down_sync(0xffffffffu, n_written, o);
  if (lane == 0 && n_written) atomicAdd(g.written, (unsigned long long)n_written);
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE>
__global__ void __launch_bounds__(kThreads, KI_MINBLOCKS)
k_integrate_q(const __grid_constant__ LaunchP<1> P, int rows_per_block, int flags) {
  __shared__ float4 tab[kMaxRowsPerBlock * (ZFAST ? 1 : 2)];   // row terms
  __shared__ float4 tabf[kSegQ];                                // fast-axis terms {p0, p1, p2, -} of the window
  integrate_q_body<ZFAST, COLOR, LE_TRUNC, WMODE, false>(P, nullptr, tab, tabf, rows_per_block, flags);
}

// The same for batched per-instance volumes (config 5): blockIdx.z selects the descriptor.
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE>
__global__ void __launch_bounds__(kThreads, KI_MINBLOCKS)
k_integrate_q_batch(const BatchDesc* __restrict__ descs, int rows_per_block, int flags) {
  __shared__ float4 tab[kMaxRowsPerBlock * (ZFAST ? 1 : 2)];
  __shared__ float4 tabf[kSegQ];
  __shared__ LaunchP<1> P;
  __shared__ const uint8_t* mask;
  const BatchDesc& d = descs[blockIdx.z];
  if (threadIdx.x == 0) { P.g = d.g; P.f[0] = d.f; mask = d.mask; }
  __syncthreads();
  if (mask) integrate_q_body<ZFAST, COLOR, LE_TRUNC, WMODE, true>(P, mask, tab, tabf, rows_per_block, flags);
  else      integrate_q_body<ZFAST, COLOR, LE_TRUNC, WMODE, false>(P, nullptr, tab, tabf, rows_per_block, flags);
}

#ifdef TSDF_B200_DEV
struct TmapParam { alignas(64) unsigned char bytes[128]; };   // a CUtensorMap, by value (param space)

// Sparse-frame kernel with the staged image tile (A/B variant, see `Stage`): colour frames, 4-wide runs.
template <bool ZFAST, bool LE_TRUNC, int WMODE>
__global__ void __launch_bounds__(kThreads, KI_MINBLOCKS)
k_integrate_staged(const __grid_constant__ LaunchP<1> P, const __grid_constant__ TmapParam tm, int tpr_log2,
                   int rows_per_block, int flags) {
  __shared__ float4 tab[kMaxRowsPerBlock * (ZFAST ? 1 : 2)];
  extern __shared__ __align__(128) unsigned char stage_raw[];
  integrate_body<ZFAST, true, LE_TRUNC, 1, 4, false, MATH_EXACT, WMODE, 2, true>(
      P, nullptr, tab, tpr_log2, rows_per_block, flags, &tm, reinterpret_cast<float4*>(stage_raw));
}
#endif  // TSDF_B200_DEV

// ---------------------------------------------------------------------------------------
// Two frames per volume pass (the off-line keyframe replay of
// examples/label_instance_rgbd.cpp:79-110 hands the library many frames per volume).  Dense frames
// only (the host checks that the slab projects inside both images, as for CULL = 0).  The voxel
// state of a run is loaded ONCE at the top of the row — no validity to wait for on dense frames, and
// the whole probe of frame 0 stands between the loads and their first use — then updated by frame 0
// and by frame 1 in registers, in frame order (bit-identical to two passes), and stored once:
// 12 B per update instead of 24.  Unlike the generic NF-frame form of integrate_body (measured
// slower than NF passes: it keeps every frame's Hit live across the volume loads, 118-244
// registers), a frame's probe results die in its own update, so the live set is one frame's.
// ---------------------------------------------------------------------------------------
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE, bool INSIDE>
__global__ void __launch_bounds__(kThreads, 3)
k_integrate2(const __grid_constant__ LaunchP<2> P, int tpr_log2, int rows_per_block, int flags) {
  constexpr int NT = ZFAST ? 1 : 2;
  constexpr int VEC = 4;
  __shared__ float4 tab[kMaxRowsPerBlock * 2 * NT];
  const GridP& g = P.g;
  const int pf = flags & 3;
  const int n_rows = g.n_mid * g.n_slow;
  const int row0 = blockIdx.x * rows_per_block;
  if (row0 >= n_rows) return;
  if (P.f[0].skip != nullptr && *reinterpret_cast<const volatile int*>(P.f[0].skip) != 0) return;
  for (int lr = threadIdx.x; lr < rows_per_block; lr += kThreads) {
    const int r = row0 + lr;
    if (r < n_rows) {
      const int slow = r / g.n_mid, mid = r - slow * g.n_mid;
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        float4 e0, e1;
        row_terms<ZFAST>(g, P.f[f].c, slow, mid, e0, e1);
        tab[(lr * 2 + f) * NT] = e0;
        if constexpr (!ZFAST) tab[(lr * 2 + f) * NT + 1] = e1;
      }
    }
  }
  __syncthreads();
  const int tpr = 1 << tpr_log2;
  const int tf = threadIdx.x & (tpr - 1);
  const int tr = threadIdx.x >> tpr_log2;
  const int rp = kThreads >> tpr_log2;
  const int f0 = (blockIdx.y * tpr + tf) * VEC;
  const bool live = f0 < g.n_fast;
  float q[2][3][VEC];   // fast-axis products of both frames (registers)
#pragma unroll
  for (int j = 0; j < VEC; ++j) {
    const float bfj = fast_coord<ZFAST>(g, f0 + j);
#pragma unroll
    for (int f = 0; f < 2; ++f) fast_terms<ZFAST>(P.f[f].c, bfj, q[f][0][j], q[f][1][j], q[f][2][j]);
  }
  unsigned int n_written = 0;
  for (int lr = tr; live && lr < rows_per_block; lr += rp) {
    const int r = row0 + lr;
    if (r >= n_rows) break;
    const int64_t base = (int64_t)r * g.n_fast + f0;
    if (pf && lr + rp < rows_per_block && r + rp < n_rows) {
      const int64_t pb = (int64_t)(r + rp) * g.n_fast + f0;
      prefetch_line<true>(g.tsdf + pb, 1);
      prefetch_line<true>(g.weight + pb, 1);
      if constexpr (COLOR) prefetch_line<true>(g.color + pb, 1);
    }
    float t[VEC], w[VEC], col[VEC];
    load_vec<VEC>(g.tsdf + base, t);
    load_vec<VEC>(g.weight + base, w);
    if constexpr (COLOR) load_vec<VEC>(g.color + base, col);
    else {
#pragma unroll
      for (int j = 0; j < VEC; ++j) col[j] = 0.0f;
    }
    bool any = false;
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      const float4 ea = tab[(lr * 2 + f) * NT];
      float4 eb = ea;
      if constexpr (!ZFAST) eb = tab[(lr * 2 + f) * NT + 1];
      float camx[VEC], camy[VEC], camz[VEC];
      bool odd = false;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        cam_coords<ZFAST>(ea, eb, q[f][0][j], q[f][1][j], q[f][2][j], camx[j], camy[j], camz[j]);
        if constexpr (!kRunSafe) odd = odd || (!em::in_safe_range(camz[j]) && camz[j] > 0.0f);
      }
      if constexpr (kRunSafe) odd = !(em::in_safe_range(camz[0]) && em::in_safe_range(camz[VEC - 1]));
      Hit hit[VEC];
      if (__builtin_expect(odd, 0)) {
#pragma unroll
        for (int j = 0; j < VEC; ++j)
          hit[j] = probe_literal_cold<COLOR, LE_TRUNC, false>(P.f[f].img, P.f[f].fx, P.f[f].cx, P.f[f].fy, P.f[f].cy,
                                                              g.trunc, g.W, g.H, nullptr, camx[j], camy[j], camz[j]);
      } else {
        Proj pr[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) pr[j] = project_exact<true, kRunSafe, INSIDE>(P.f[f], g, camx[j], camy[j], camz[j]);
#pragma unroll
        for (int j = 0; j < VEC; ++j) hit[j] = sample_exact<COLOR, LE_TRUNC, false>(P.f[f], g, nullptr, pr[j], camz[j]);
      }
      const float obs = P.f[f].obs_weight;
#pragma unroll
      for (int j = 0; j < VEC; ++j) {
        float nt = t[j], nw = w[j], nc = col[j];
        update_voxel<COLOR, MATH_EXACT, WMODE>(nt, nw, nc, hit[j], obs);
        const bool ok = hit[j].ok;
        t[j] = ok ? nt : t[j];
        w[j] = ok ? nw : w[j];
        col[j] = ok ? nc : col[j];
        n_written += ok ? 1u : 0u;
        any = any || ok;
      }
    }
    if (any) {
      store_vec<VEC>(g.tsdf + base, t);
      store_vec<VEC>(g.weight + base, w);
      if constexpr (COLOR) store_vec<VEC>(g.color + base, col);
    }
  }
  for (int o = 16; o > 0; o >>= 1) n_written += __shfl_down_sync(0xffffffffu, n_written, o);
  if ((threadIdx.x & 31) == 0 && n_written) atomicAdd(g.written, (unsigned long long)n_written);
}

// Batched per-instance volumes (config 5): blockIdx.z selects a descriptor {grid, pose, mask}
// from global memory; one launch covers every active instance.
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int VEC, int WMODE>
__global__ void __launch_bounds__(kThreads)
k_integrate_batch(const BatchDesc* __restrict__ descs, int tpr_log2, int rows_per_block, int flags) {
  __shared__ float4 tab[kMaxRowsPerBlock * (ZFAST ? 1 : 2)];
  __shared__ LaunchP<1> P;
  __shared__ const uint8_t* mask;
  const BatchDesc& d = descs[blockIdx.z];
  if (threadIdx.x == 0) { P.g = d.g; P.f[0] = d.f; mask = d.mask; }
  __syncthreads();
  if (mask) integrate_body<ZFAST, COLOR, LE_TRUNC, 1, VEC, true, MATH_EXACT, WMODE, 2>(P, mask, tab, tpr_log2, rows_per_block, flags);
  else      integrate_body<ZFAST, COLOR, LE_TRUNC, 1, VEC, false, MATH_EXACT, WMODE, 2>(P, nullptr, tab, tpr_log2, rows_per_block, flags);
}

// Moves the exact N_written of the launch(es) before it to the pinned host mirror (mapped
// memory: this IS the step's 8-byte device->host read) and re-arms the counter, so the
// compute stream carries no memset / memcpy operations between integrate launches.
__global__ void k_publish(unsigned long long* __restrict__ written, unsigned long long* __restrict__ h_written,
                          int accumulate, int* __restrict__ tail, int tail_words) {
  if (threadIdx.x == 0) {
    const unsigned long long n = *written;
    *h_written = accumulate ? *h_written + n : n;
    *written = 0ull;
  }
  // the launch before this one consumed the frame slot `tail` belongs to: zero its max-depth tiles
  // now, so that the next prep into the slot needs no memset on the copy stream
  for (int i = threadIdx.x; i < tail_words; i += blockDim.x) tail[i] = 0;
}

// Per-launch choices (no global state: handles may be driven from different host threads).
struct LaunchOpts {
  int prefetch = 1;    // 0 off, 1 L2, 2 L1 (legacy kernel only)
  int kernel = 1;      // 0: dense kernel, 1: segment cull + staged probe, 2: legacy dense (A/B only)
  int rows_mult = 16;  // row iterations per thread
  bool balance = true; // dense kernel, few waves: size the blocks so that the grid fills whole waves
  bool compact = true;    // sparse kernel: compact block tiles (one warp per row segment)
  bool block_test = true; // sparse kernel: block-level hull test before the per-warp tests
  bool quarters = true;   // sparse kernel: quarter-segment work lists (k_integrate_q)
  const void* tmap = nullptr;  // sparse kernel: CUtensorMap of the prepared frame => staged image tile (A/B)
};

struct Geometry { int tpr_log2, rows_per_block; dim3 grid; };

Geometry plan(int n_fast, int n_rows, int vec, int nz, int rows_mult = 16, int max_tpr = kThreads) {
  Geometry q;
  int vec_per_row = (n_fast + vec - 1) / vec;
  q.tpr_log2 = 0;
  while ((kThreads >> q.tpr_log2) > kMaxRowsPerBlock) ++q.tpr_log2;   // one pass of rows must fit the row table
  while ((1 << q.tpr_log2) < vec_per_row && (1 << q.tpr_log2) < max_tpr) ++q.tpr_log2;
  int tpr = 1 << q.tpr_log2;
  int rp = kThreads / tpr;
  q.rows_per_block = rp * rows_mult;
  if (q.rows_per_block > kMaxRowsPerBlock) q.rows_per_block = kMaxRowsPerBlock;
  if (q.rows_per_block < rp) q.rows_per_block = rp;
  q.grid = dim3((n_rows + q.rows_per_block - 1) / q.rows_per_block, (vec_per_row + tpr - 1) / tpr, nz);
  return q;
}

// Small and mid-size dense volumes run only a few waves of blocks, so a last wave that is mostly
// empty costs a whole block time (200^3: 625 blocks on 592 resident slots = 2 block times for
// 1.06 waves of work).  Re-size the blocks so that the grid is just under a whole number of waves.
// Work per voxel is uniform in the dense kernel; the sparse kernel keeps small blocks (its work per
// block is not uniform and finer blocks balance better).
void balance_waves(Geometry* q, int n_rows, int slots) {
  const long blocks = (long)q->grid.x * q->grid.y * q->grid.z;
  if (slots <= 0 || blocks > 8L * slots) return;
  const int rp = kThreads >> q->tpr_log2;
  const long passes = (n_rows + rp - 1) / rp;
  const long waves = (blocks + slots - 1) / slots;
  const long gx_target = std::max(1L, waves * slots / ((long)q->grid.y * q->grid.z));
  long m = (passes + gx_target - 1) / gx_target;
  m = std::max<long>(std::min<long>(m, kMaxRowsPerBlock / rp), 1);
  q->rows_per_block = (int)m * rp;
  q->grid.x = (unsigned)((n_rows + q->rows_per_block - 1) / q->rows_per_block);
}

int device_sm_count() {
  // every GPU of a box is the same part; queried once (thread-safe static initialisation)
  static const int sms = [] {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      n = 0;
    return n;
  }();
  return sms;
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int NF, int VEC, int MATH, int WMODE>
int launch_t(const GridP& g, const FrameP* frames, const LaunchOpts& o, cudaStream_t st) {
  const int pf = o.prefetch;
  LaunchP<NF> P;
  P.g = g;
  for (int i = 0; i < NF; ++i) P.f[i] = frames[i];
  // sparse kernel: one warp per row segment (32 runs), 8 rows in parallel, so that a block is a
  // compact tile (128 rows x 128 voxels) the block-level hull test can reject as a whole
  Geometry q = plan(g.n_fast, g.n_mid * g.n_slow, VEC, 1, o.rows_mult, (o.kernel == 1 && o.compact) ? 32 : kThreads);
  if ((o.kernel == 0 || o.kernel == 3) && o.balance) {
    // resident blocks per SM of this instantiation (same on every B200); C++11 static
    // initialisation is thread-safe: handles may be driven from different host threads
    static const int occ = [] {
      int n = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
              &n, k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC, MATH, WMODE, 0>, kThreads, 0) != cudaSuccess)
        n = 0;
      return n;
    }();
    balance_waves(&q, g.n_mid * g.n_slow, occ * device_sm_count());
  }
#ifdef TSDF_B200_DEV
  if constexpr (COLOR && NF == 1 && VEC == 4 && MATH == MATH_EXACT) {
    if (o.kernel == 1 && o.tmap != nullptr && o.block_test) {
      TmapParam tm;
      memcpy(tm.bytes, o.tmap, 128);
      constexpr int kStageBytes = kStageW * kStageH * 16;
      static const bool attr_ok = cudaFuncSetAttribute(k_integrate_staged<ZFAST, LE_TRUNC, WMODE>,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kStageBytes) == cudaSuccess;
      if (attr_ok) {
        k_integrate_staged<ZFAST, LE_TRUNC, WMODE><<<q.grid, kThreads, kStageBytes, st>>>(P, tm, q.tpr_log2, q.rows_per_block, pf | 4);
        CU_LAUNCHED();
        return TSDF_OK;
      }
    }
  }
#endif
  if constexpr (NF == 1 && VEC == 4 && MATH == MATH_EXACT) {
    // sparse frames on compact tiles (one warp per 128-voxel segment, <= 16 row iterations per warp):
    // the quarter work-list kernel
    if (o.kernel == 1 && o.quarters && o.compact && q.tpr_log2 == 5 && q.rows_per_block <= 16 * (kThreads / 32)) {
      k_integrate_q<ZFAST, COLOR, LE_TRUNC, WMODE><<<q.grid, kThreads, 0, st>>>(P, q.rows_per_block, pf | (o.block_test ? 4 : 0));
      CU_LAUNCHED();
      return TSDF_OK;
    }
  }
  if constexpr (NF == 1 && VEC == 4 && MATH == MATH_EXACT) {
    if (o.kernel == 3) {   // dense, and the host has shown every voxel strictly inside the image
      k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC, MATH, WMODE, 3><<<q.grid, kThreads, 0, st>>>(P, q.tpr_log2, q.rows_per_block, pf);
      CU_LAUNCHED();
      return TSDF_OK;
    }
  }
  if (o.kernel == 1)      k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC, MATH, WMODE, 2><<<q.grid, kThreads, 0, st>>>(P, q.tpr_log2, q.rows_per_block, pf | (o.block_test ? 4 : 0));
  else if (o.kernel == 0 || o.kernel == 3) k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC, MATH, WMODE, 0><<<q.grid, kThreads, 0, st>>>(P, q.tpr_log2, q.rows_per_block, pf);
#ifdef TSDF_B200_DEV
  else                  k_integrate<ZFAST, COLOR, LE_TRUNC, NF, VEC, MATH, WMODE, 1><<<q.grid, kThreads, 0, st>>>(P, q.tpr_log2, q.rows_per_block, pf);
#endif
  CU_LAUNCHED();
  return TSDF_OK;
}

template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE, bool INSIDE>
int launch_pair_ti(const GridP& g, const FrameP* frames, const LaunchOpts& o, cudaStream_t st) {
  LaunchP<2> P;
  P.g = g;
  P.f[0] = frames[0];
  P.f[1] = frames[1];
  Geometry q = plan(g.n_fast, g.n_mid * g.n_slow, 4, 1, o.rows_mult, kThreads);
  if (o.balance) {
    static const int occ = [] {
      int n = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_integrate2<ZFAST, COLOR, LE_TRUNC, WMODE, INSIDE>, kThreads, 0) != cudaSuccess) n = 0;
      return n;
    }();
    balance_waves(&q, g.n_mid * g.n_slow, occ * device_sm_count());
  }
  k_integrate2<ZFAST, COLOR, LE_TRUNC, WMODE, INSIDE><<<q.grid, kThreads, 0, st>>>(P, q.tpr_log2, q.rows_per_block, o.prefetch ? 1 : 0);
  CU_LAUNCHED();
  return TSDF_OK;
}
template <bool ZFAST, bool COLOR, bool LE_TRUNC, int WMODE>
int launch_pair_t(const GridP& g, const FrameP* frames, const LaunchOpts& o, cudaStream_t st) {
  return o.kernel == 3 ? launch_pair_ti<ZFAST, COLOR, LE_TRUNC, WMODE, true>(g, frames, o, st)
                       : launch_pair_ti<ZFAST, COLOR, LE_TRUNC, WMODE, false>(g, frames, o, st);
}

int launch_pair(bool zfast, bool color, int wmode, const GridP& g, const FrameP* fr, const LaunchOpts& o, cudaStream_t st) {
  if (zfast) {
    if (color) return wmode == 1 ? launch_pair_t<true, true, false, 1>(g, fr, o, st) : launch_pair_t<true, true, false, 0>(g, fr, o, st);
    return launch_pair_t<true, false, false, 0>(g, fr, o, st);
  }
  if (color) return wmode == 1 ? launch_pair_t<false, true, true, 1>(g, fr, o, st) : launch_pair_t<false, true, true, 0>(g, fr, o, st);
  return launch_pair_t<false, false, true, 0>(g, fr, o, st);
}

// layout (and with it the truncation compare) x colour x integer-weight mode, for one NF/VEC/MATH
template <int NF, int VEC, int MATH>
int launch_sem(bool zfast, bool color, int wmode, const GridP& g, const FrameP* fr, const LaunchOpts& o, cudaStream_t st) {
  if (zfast) {
    if (color) {
      if constexpr (MATH == MATH_EXACT) { if (wmode == 1) return launch_t<true, true, false, NF, VEC, MATH, 1>(g, fr, o, st); }
      return launch_t<true, true, false, NF, VEC, MATH, 0>(g, fr, o, st);
    }
    return launch_t<true, false, false, NF, VEC, MATH, 0>(g, fr, o, st);
  }
  if (color) {
    if constexpr (MATH == MATH_EXACT) { if (wmode == 1) return launch_t<false, true, true, NF, VEC, MATH, 1>(g, fr, o, st); }
    return launch_t<false, true, true, NF, VEC, MATH, 0>(g, fr, o, st);
  }
  return launch_t<false, false, true, NF, VEC, MATH, 0>(g, fr, o, st);
}

}  // namespace

void fill_grid_params(const tsdf_vol* v, bool color, GridP* g) {
  const tsdf_config& c = v->cfg;
  g->n_fast = v->zfast ? v->dzl : c.dim_x;
  g->n_mid = c.dim_y;
  g->n_slow = v->zfast ? c.dim_x : v->dzl;
  g->nx_ = c.dim_x;
  g->nz_ = v->dzl;
  g->z0 = c.z_begin;
  g->ox = c.origin[0]; g->oy = c.origin[1]; g->oz = c.origin[2];
  g->vs = c.voxel_size; g->trunc = c.trunc_margin;
  g->inv_trunc = 1.0f / c.trunc_margin;   // host IEEE division: correctly rounded reciprocal
  g->w_lim = (float)c.im_w - 0.5f;        // exact for W < 2^23
  g->h_lim = (float)c.im_h - 0.5f;
  g->W = c.im_w; g->H = c.im_h;
  g->tsdf = v->d_tsdf; g->weight = v->d_weight; g->color = color ? v->d_color : nullptr;
  g->written = v->d_written + (v->lseq % kWrittenRing);
  g->dbg = v->d_dbg;
}

// Host-side, double precision: do all 8 corner voxels of the slab project at least one pixel
// inside the image, in front of the camera, for every frame?  (Frustum and slab are convex.)
static bool slab_inside_every_frustum(const GridP& g, const FrameP* fr, int nf) {
  for (int f = 0; f < nf; ++f) {
    const float* c = fr[f].c;
    for (int corner = 0; corner < 8; ++corner) {
      // index extents per axis: x in [0,nx), y in [0,ny), z in [z0, z0+nz)
      const int ix = (corner & 1) ? g.nx_ - 1 : 0;
      const int iy = (corner & 2) ? g.n_mid - 1 : 0;
      const int iz = g.z0 + ((corner & 4) ? g.nz_ - 1 : 0);
      const double t0 = ((double)g.ox + (double)ix * g.vs) - c[3];
      const double t1 = ((double)g.oy + (double)iy * g.vs) - c[7];
      const double t2 = ((double)g.oz + (double)iz * g.vs) - c[11];
      const double x = c[0] * t0 + c[4] * t1 + c[8] * t2;
      const double y = c[1] * t0 + c[5] * t1 + c[9] * t2;
      const double z = c[2] * t0 + c[6] * t1 + c[10] * t2;
      if (!(z > 1e-6)) return false;
      const double u = fr[f].fx * (x / z) + fr[f].cx, w = fr[f].fy * (y / z) + fr[f].cy;
      if (!(u >= 0.5 && u <= g.W - 1.5 && w >= 0.5 && w <= g.H - 1.5)) return false;
    }
  }
  return true;
}

// The stronger statement the INSIDE kernels need: the test above (all 8 corners at least one pixel inside the
// image, in double precision; slab and frustum are convex, and "u >= a, z > 0" is a half-space, so every voxel is
// inside in exact arithmetic) AND the fp32 error of the kernel's own u / v stays below a quarter pixel, so the
// COMPUTED u, v lie in [0.25, W - 1.25] x [0.25, H - 1.25]: :39 and :43 hold for every voxel, the rounded pixel
// is inside the image.  Error model: a camera coordinate is a sum of three products of a rotation entry
// (|c| <= ~1) and a difference "origin + index * vs - translation", about 12 roundings of quantities bounded by
// M = |origin| + extent + |translation| (1-norms): |err| <= 12 * 2^-24 * M < 7.2e-7 M; E = 1.5e-6 M doubles it.
// u = fx * (x / z) + cx then moves by at most fx * E * (1 + |x/z|) / z_min plus three roundings of values
// below max(W, H) (< 3e-4 px), and |x/z| <= max(W, H) / min(fx, fy) inside the image.
static bool slab_inside_with_margin(const GridP& g, const FrameP* fr, int nf) {
  if (!slab_inside_every_frustum(g, fr, nf)) return false;
  for (int f = 0; f < nf; ++f) {
    const float* c = fr[f].c;
    double rot = 0.0;
    for (int i = 0; i < 12; ++i) if ((i & 3) != 3) rot = std::max(rot, (double)fabsf(c[i]));
    if (!(rot <= 1.0001)) return false;   // not a rotation (a scaled base pose): keep the tested kernel
    const double ext = ((double)g.nx_ + g.n_mid + g.z0 + g.nz_) * g.vs;
    const double M = fabs((double)g.ox) + fabs((double)g.oy) + fabs((double)g.oz) + ext + fabs((double)c[3]) + fabs((double)c[7]) + fabs((double)c[11]);
    double zmin = 1e300;
    for (int corner = 0; corner < 8; ++corner) {
      const int ix = (corner & 1) ? g.nx_ - 1 : 0, iy = (corner & 2) ? g.n_mid - 1 : 0, iz = g.z0 + ((corner & 4) ? g.nz_ - 1 : 0);
      const double t0 = ((double)g.ox + (double)ix * g.vs) - c[3], t1 = ((double)g.oy + (double)iy * g.vs) - c[7],
                   t2 = ((double)g.oz + (double)iz * g.vs) - c[11];
      zmin = std::min(zmin, c[2] * t0 + c[6] * t1 + c[10] * t2);
    }
    const double fmax = std::max(fabs((double)fr[f].fx), fabs((double)fr[f].fy)), fmin = std::min(fabs((double)fr[f].fx), fabs((double)fr[f].fy));
    if (!(zmin >= 1e-3 && zmin < 1e12 && fmin > 0.0 && M < 1e12)) return false;
    const double E = 1.5e-6 * M;
    const double du = fmax * E * (1.0 + std::max(g.W, g.H) / fmin) / zmin + 1e-3;
    if (!(du < 0.25)) return false;
  }
  return true;
}

int launch_publish(tsdf_vol* v, bool accumulate, cudaStream_t st, void* d_frame_to_clear) {
  int* tail = d_frame_to_clear ? reinterpret_cast<int*>(frame_dmax_ptr(d_frame_to_clear, frame_pixels(v->cfg))) : nullptr;
  const int words = tail ? (int)(frame_tail_bytes(v->cfg.im_h, v->cfg.im_w) / 4) : 0;
  k_publish<<<1, tail ? 256 : 32, 0, st>>>(v->d_written + (v->lseq % kWrittenRing), v->hd_written, accumulate ? 1 : 0, tail, words);
  CU_LAUNCHED();
  return TSDF_OK;
}

// nf frames, one launch each (see the note in the body).  wmode 1 = the host guarantees
// integer weights <= 4094 and obs_weight == 1 for every frame (see blend_intw).
int launch_integrate(tsdf_vol* v, const FrameP* frames, int nf, bool color, cudaStream_t st) {
  GridP g;
  fill_grid_params(v, color, &g);
  const bool vec4 = (g.n_fast % 4) == 0;
  const bool literal = (v->variant & 1) != 0;
  // A/B knobs (tsdf_set_variant): bit0 literal math; bits1-2 prefetch 1=L2 2=L1 (default L2);
  // bit5 prefetch off; bits3-4 rows per block x64 / x32 (default x16); bit6/bit7 segment cull off/on;
  // bits8-9 rows per block x8 / x4; bit11 wave balancing off
  LaunchOpts o;
  o.balance = (v->variant & 2048) == 0;
  o.compact = (v->variant & 4096) == 0;      // bit12: old block shape (row-major, up to 256 threads per row)
  o.block_test = (v->variant & 8192) == 0;   // bit13: block-level hull test off
  o.quarters = (v->variant & (1 << 21)) == 0; // bit21: plain sparse kernel instead of the quarter work lists
  o.prefetch = (v->variant & 32) ? 0 : (((v->variant >> 1) & 3) ? ((v->variant >> 1) & 3) : 1);
  o.rows_mult = (v->variant & 8) ? 64 : ((v->variant & 16) ? 32 : ((v->variant & 256) ? 8 : ((v->variant & 512) ? 4 : 16)));
  // cull: forced off (bit 6) / on (bit 7), else on unless the host can show that the whole slab
  // projects inside every frame's image (then no segment can be culled by the frustum and the
  // plain kernel is ~2-5 % faster); a pure performance choice, results are identical
  o.kernel = (v->variant & 1024) ? 2 /* legacy dense kernel, A/B only */
         : (v->variant & 64) ? 0 : ((v->variant & 128) ? 1 : (slab_inside_every_frustum(g, frames, nf) ? 0 : 1));
  // (sparse kernel: 256-row blocks, i.e. x32 with a 256-entry row table, measured no better than
  // x16 once the block test was in place — S-kitti 0.134 vs 0.130 ms, S-real equal — so x16 stays)
  const int wmode = (color && v->wmode_ok) ? 1 : 0;
  // Frames go in frame order.  Two consecutive DENSE frames (the slab projects inside both images, so
  // the branch-free kernel applies) share one pass over the volume (k_integrate2: voxel state read
  // and written once for two updates); anything else is one launch per frame.  The generic NF-frame
  // form of integrate_body (every frame's probe results live across the volume loads) was measured
  // slower than NF passes in round 1 — 4.1-5.7 ms per frame against 3.9 — and is not instantiated.
  const bool pairs_ok = vec4 && !literal && (v->variant & (1 << 14)) == 0 && o.kernel != 2;
  // bit15: keep the image-bounds tests in the dense kernels even where the host can prove them true (A/B)
  const bool inside_ok = vec4 && !literal && (v->variant & ((1 << 15) | 64 | 128 | 1024)) == 0;
  for (int i = 0; i < nf;) {
    int rc;
    // (variant bit 6 forces the dense kernels whatever the geometry, bit 7 the sparse one, bit 14 no pairs)
    if (pairs_ok && i + 1 < nf && !(v->variant & 128) &&
        ((v->variant & 64) || slab_inside_every_frustum(g, frames + i, 2))) {
      LaunchOpts o2 = o;
      o2.kernel = (inside_ok && slab_inside_with_margin(g, frames + i, 2)) ? 3 : 0;
      rc = launch_pair(v->zfast, color, wmode, g, frames + i, o2, st);
      if (rc) return rc;
      i += 2;
      continue;
    }
    LaunchOpts o1 = o;
    o1.tmap = (v->variant & (1 << 20)) ? frames[i].h_tmap : nullptr;   // bit20: staged image tile (A/B)
    if (!(v->variant & (64 | 128 | 1024))) o1.kernel = slab_inside_every_frustum(g, frames + i, 1) ? 0 : 1;
    if (o1.kernel == 0 && inside_ok && slab_inside_with_margin(g, frames + i, 1)) o1.kernel = 3;
#ifdef TSDF_B200_DEV
    if (literal) rc = vec4 ? launch_sem<1, 4, MATH_LITERAL>(v->zfast, color, 0, g, frames + i, o1, st)
                           : launch_sem<1, 1, MATH_LITERAL>(v->zfast, color, 0, g, frames + i, o1, st);
    else
#endif
                 rc = vec4 ? launch_sem<1, 4, MATH_EXACT>(v->zfast, color, wmode, g, frames + i, o1, st)
                           : launch_sem<1, 1, MATH_EXACT>(v->zfast, color, wmode, g, frames + i, o1, st);
    if (rc) return rc;
    ++i;
  }
  return TSDF_OK;
}

// All instances of a batch share semantics and colour mode; the grid is sized for the
// largest instance.  descs: n_active device descriptors.
int launch_integrate_batch(const BatchDesc* d_descs, int n_active, int max_fast, int max_rows,
                           bool zfast, bool color, bool vec4, int wmode, cudaStream_t st) {
  Geometry q = plan(max_fast, max_rows, vec4 ? 4 : 1, n_active);
  if (vec4 && max_fast >= kSegQ) {
    // compact tiles + quarter work lists (see k_integrate_q); an instance smaller than the grid's
    // largest leaves its surplus blocks at the first test
    Geometry qq = plan(max_fast, max_rows, 4, n_active, 16, 32);
    if (qq.tpr_log2 == 5 && qq.rows_per_block <= 16 * (kThreads / 32)) {
#define LQ(ZF, CO, LE, WM) k_integrate_q_batch<ZF, CO, LE, WM><<<qq.grid, kThreads, 0, st>>>(d_descs, qq.rows_per_block, 1 | 4)
      if (zfast) { if (color) { if (wmode == 1) LQ(true, true, false, 1); else LQ(true, true, false, 0); } else LQ(true, false, false, 0); }
      else       { if (color) { if (wmode == 1) LQ(false, true, true, 1); else LQ(false, true, true, 0); } else LQ(false, false, true, 0); }
#undef LQ
      CU_LAUNCHED();
      return TSDF_OK;
    }
  }
#define LB(ZF, CO, LE, V, WM) k_integrate_batch<ZF, CO, LE, V, WM><<<q.grid, kThreads, 0, st>>>(d_descs, q.tpr_log2, q.rows_per_block, 1 /* L2 prefetch */ | 4 /* block hull test */)
  if (zfast) {
    if (color) { if (wmode == 1) { if (vec4) LB(true, true, false, 4, 1); else LB(true, true, false, 1, 1); }
                 else            { if (vec4) LB(true, true, false, 4, 0); else LB(true, true, false, 1, 0); } }
    else       { if (vec4) LB(true, false, false, 4, 0); else LB(true, false, false, 1, 0); }
  } else {
    if (color) { if (wmode == 1) { if (vec4) LB(false, true, true, 4, 1); else LB(false, true, true, 1, 1); }
                 else            { if (vec4) LB(false, true, true, 4, 0); else LB(false, true, true, 1, 0); } }
    else       { if (vec4) LB(false, false, true, 4, 0); else LB(false, false, true, 1, 0); }
  }
#undef LB
  CU_LAUNCHED();
  return TSDF_OK;
}
