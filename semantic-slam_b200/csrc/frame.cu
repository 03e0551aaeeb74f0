// frame.cu — frame upload companions: turns the raw upload (depth fp32|u16 + rgb u8x3)
// into the "prepared frame" the integrate kernel gathers from, plus small utility kernels.
//
// Prepared frame:
//   colour on : float4 per pixel {depth, c0, c1, c2} -> ONE 128-bit gather per voxel, and
//               the colour channels arrive as exact floats (no per-voxel unpacking)
//   colour off: float per pixel
// Depth is canonicalised here: a sample failing the reference's validity test
// `d <= 0 || d > max_depth` (src/tsdf.cu:46) is stored as 0.0f, so the hot loop needs a
// single compare.  NaN fails neither test in the reference and is kept.
// The upstream fold floor(c2*65536 + c1*256 + c0) followed by its floor/divide unfold returns
// c2, c1, c0 unchanged (each < 256), so the channels are stored directly.
// Optional per-instance mask: depth * (mask/255), i.e. depth kept where mask >= 128
// (src/Engine.cpp:192-193; OpenCV 8-bit divide rounds to nearest).
#include "common.cuh"

namespace {

__device__ __forceinline__ unsigned ld_flag_volatile(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// One 256-pixel chunk per block iteration (grid-stride over chunks: the peer-pull form runs with two
// blocks per SM so that only ~300 threads poll the root's flag over NVLink).
__global__ void k_frame_prep(const void* __restrict__ depth, int fmt, float scale, int sub_c,
                             int sub_r, const uint8_t* __restrict__ rgb,
                             const uint8_t* __restrict__ mask, int H, int W, float max_depth,
                             bool with_color, float4 intr /* fx, cx, fy, cy */, void* __restrict__ out,
                             int* __restrict__ dmax_bits, const PrepSync sy) {
  // ---- head: frame distribution flags (comm.cu) --------------------------------------------
  if (sy.ready_flag && blockIdx.x == 0 && threadIdx.x == 0) {
    // root: the raw frame this kernel reads was written by the H2D copies before it in stream order
    __threadfence_system();
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(sy.ready_flag), "r"(sy.ready_value));
  }
  if (sy.wait_flag) {
    // peer: the source is the ROOT's raw slot (NVLink); wait until the root says frame wait_target
    // is in it.  On timeout nothing is read or written and the rank's error words are raised: the
    // integrate launches behind this frame see FrameP::skip and leave the volume untouched.
    __shared__ int timed_out;
    if (threadIdx.x == 0) {
      int bad = 0;
      if (sy.wait_target > 0) {
        const unsigned long long t0 = globaltimer_ns();
        unsigned ns = 100;
        while ((int)ld_flag_volatile(sy.wait_flag) < sy.wait_target) {
          __nanosleep(ns);
          if (ns < 2000) ns *= 2;
          if (globaltimer_ns() - t0 > sy.timeout_ns || *reinterpret_cast<volatile int*>(sy.d_err)) { bad = 1; break; }
        }
      }
      if (bad) {
        *reinterpret_cast<volatile int*>(sy.d_err) = 1;
        *reinterpret_cast<volatile int*>(sy.h_err) = 1;
        __threadfence_system();
      }
      timed_out = bad;
      // a block that gives up still counts as arrived, so that the last-block detection of the tail stays
      // armed for the next launch even when only SOME blocks of this one timed out; `done` is not raised
      if (bad && sy.done_flag && atomicAdd(sy.arrive, 1u) == gridDim.x - 1) *sy.arrive = 0u;
    }
    __syncthreads();
    if (timed_out) return;   // block-uniform
  }
  const int n = H * W;
  const int lane = threadIdx.x & 31;
  int bmax = 0;
  for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
    const int i = base + threadIdx.x;
    float d = 0.0f;
    if (i < n) {
      if (fmt == TSDF_DEPTH_U16 || fmt == TSDF_DEPTH_RANGE_U16) d = (float)reinterpret_cast<const uint16_t*>(depth)[i] * scale;
      else                                                      d = reinterpret_cast<const float*>(depth)[i] * scale;
      if (fmt >= TSDF_DEPTH_RANGE_F32) {  // lidar range -> camera-z depth, src/DoN.cpp:90-98
        const int r = i / W, c = i - r * W;
        const float x = ((float)c - intr.y) / intr.x;
        const float y = ((float)r - intr.w) / intr.z;
        const float rim = sqrtf(x * x + y * y + 1.0f);
        d = d / rim;
      }
      if (sub_c > 1 || sub_r > 1) {  // examples/label_instance_rgbd.cpp:89-97
        int r = i / W, c = i - r * W;
        if ((c % sub_c) != 0 || (r % sub_r) != 0) d = 0.0f;
      }
      if (mask) d = d * (mask[i] >= 128 ? 1.0f : 0.0f);
      if (d <= 0.0f || d > max_depth) d = 0.0f;
      if (with_color) {
        // channels of the folded colour floor(c2*65536 + c1*256 + c0) as exact floats
        float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
        if (rgb) { c0 = (float)rgb[3 * i + 0]; c1 = (float)rgb[3 * i + 1]; c2 = (float)rgb[3 * i + 2]; }
        reinterpret_cast<float4*>(out)[i] = make_float4(d, c0, c1, c2);
      } else {
        reinterpret_cast<float*>(out)[i] = d;
      }
    }
    // Largest depth of the frame and of its 16x16 / 64x64 pixel tiles, for the integrate kernel's
    // conservative segment cull.  Stored depths are >= +0 or NaN; a NaN (which the reference
    // integrates) counts as +inf so that it disables the cull.  Non-negative floats order like
    // their bits, so integer atomicMax works; lanes of a warp that share a tile are reduced first.
    const int b = (d != d) ? 0x7f800000 : __float_as_int(d);
    const int r = i < n ? i / W : 0, c = i < n ? i - r * W : 0;
    const int tw0 = tiles_across(W, kTile0Log2), tw1 = tiles_across(W, kTile1Log2);
    const int t0 = i < n ? (r >> kTile0Log2) * tw0 + (c >> kTile0Log2) : -1;
    const int t1 = i < n ? (r >> kTile1Log2) * tw1 + (c >> kTile1Log2) : -1;
    int* tiles0 = dmax_bits + kTailHeadWords;
    int* tiles1 = tiles0 + tw0 * tiles_across(H, kTile0Log2);
    unsigned m = __match_any_sync(0xffffffffu, t0);
    int mx = __reduce_max_sync(m, b);
    if (t0 >= 0 && mx > 0 && lane == __ffs(m) - 1) atomicMax(tiles0 + t0, mx);
    m = __match_any_sync(0xffffffffu, t1);
    mx = __reduce_max_sync(m, b);
    if (t1 >= 0 && mx > 0 && lane == __ffs(m) - 1) atomicMax(tiles1 + t1, mx);
    bmax = max(bmax, b);
  }
  bmax = __reduce_max_sync(0xffffffffu, bmax);
  __shared__ int wmax[8];
  if (lane == 0) wmax[threadIdx.x >> 5] = bmax;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) bmax = max(bmax, wmax[k]);
    if (bmax > 0) atomicMax(dmax_bits, bmax);
    // ---- tail: "this rank has consumed the root's raw slot" ---------------------------------
    if (sy.done_flag) {
      __threadfence();
      if (atomicAdd(sy.arrive, 1u) == gridDim.x - 1) {
        *sy.arrive = 0u;                 // re-arm for the next launch (stream-ordered after this one)
        // the root may only overwrite this slot after it has seen done >= wait_target from every peer;
        // if its `ready` already names the frame `ring` later, its wait gave up (timeout) and the
        // slot was rewritten while this kernel read it: the prepared frame is torn
        if (sy.ring > 0 && (int)ld_flag_volatile(sy.wait_flag) >= sy.wait_target + sy.ring) {
          *reinterpret_cast<volatile int*>(sy.d_err) = 3;
          *reinterpret_cast<volatile int*>(sy.h_err) = 3;
        }
        __threadfence_system();
        // some blocks of this launch gave up on the wait: the frame is incomplete, the slot is not "consumed"
        if (*reinterpret_cast<volatile int*>(sy.d_err) == 0)
          asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(sy.done_flag), "r"(sy.done_value));
      }
    }
  }
}

__global__ void k_fill(float* __restrict__ p, float val, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // 128-bit body, scalar tail
  int64_t n4 = n >> 2;
  float4 v4 = make_float4(val, val, val, val);
  for (int64_t k = i; k < n4; k += stride) reinterpret_cast<float4*>(p)[k] = v4;
  for (int64_t k = (n4 << 2) + i; k < n; k += stride) p[k] = val;
}

// Per-instance origin (K9): min over valid masked pixels of the back-projected x, y, z.
// src/Object.cpp:37-49.  fmin of floats is order independent, so a parallel reduction is
// bit-identical to the reference's serial scan.
__global__ void k_instance_origin(const float* __restrict__ depth, const uint8_t* __restrict__ mask,
                                  int H, int W, float cx, float cy, float ifx, float ify,
                                  float* __restrict__ out3) {
  float mx = 1000.0f, my = 1000.0f, mz = 1000.0f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H * W; i += gridDim.x * blockDim.x) {
    float z = depth[i];
    if (mask) z = z * (mask[i] >= 128 ? 1.0f : 0.0f);
    if (z <= 0.0f) continue;
    int r = i / W, c = i - r * W;
    float x = ((float)c - cx) * z * ifx;
    float y = ((float)r - cy) * z * ify;
    mx = x < mx ? x : mx;
    my = y < my ? y : my;
    mz = z < mz ? z : mz;
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = fminf(mx, __shfl_down_sync(0xffffffffu, mx, o));
    my = fminf(my, __shfl_down_sync(0xffffffffu, my, o));
    mz = fminf(mz, __shfl_down_sync(0xffffffffu, mz, o));
  }
  if ((threadIdx.x & 31) == 0) {
    // positive-float ordering == signed-int ordering only for non-negative values; use the
    // sign-aware trick: atomicMin on int for >=0, atomicMax on uint for <0
    auto amin = [](float* addr, float v) {
      if (v >= 0.0f) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
      else atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
    };
    amin(out3 + 0, mx);
    amin(out3 + 1, my);
    amin(out3 + 2, mz);
  }
}

}  // namespace

int launch_frame_prep(const void* d_depth, int depth_fmt, float depth_scale, int sub_c, int sub_r,
                      const uint8_t* d_rgb, const uint8_t* d_mask, int H, int W, float max_depth,
                      bool with_color, const float intr[4], void* d_frame, cudaStream_t st, bool clear_tail,
                      const PrepSync* sync) {
  int n = H * W;
  const float4 k4 = intr ? make_float4(intr[0], intr[1], intr[2], intr[3]) : make_float4(1.f, 0.f, 1.f, 0.f);
  int* d_dmax = reinterpret_cast<int*>(frame_dmax_ptr(d_frame, (size_t)n));
  // the tail must be zero before the kernel's atomicMax: either cleared here or already cleared by
  // the k_publish behind the launch that last consumed this slot (a stale tail would only weaken
  // the cull, never change a result: the maxima can only be too large)
  if (clear_tail) CU_CHECK(cudaMemsetAsync(d_dmax, 0, frame_tail_bytes(H, W), st));
  const PrepSync sy = sync ? *sync : PrepSync();
  const int chunks = (n + 255) / 256;
  const int grid = sy.wait_flag ? (chunks < 296 ? chunks : 296) : chunks;
  k_frame_prep<<<grid, 256, 0, st>>>(d_depth, depth_fmt, depth_scale, sub_c, sub_r, d_rgb,
                                     d_mask, H, W, max_depth, with_color, k4, d_frame, d_dmax, sy);
  CU_LAUNCHED();
  return TSDF_OK;
}

int launch_fill(float* p, float val, int64_t n, cudaStream_t st) {
  if (n <= 0) return TSDF_OK;
  k_fill<<<148 * 8, 256, 0, st>>>(p, val, n);
  CU_LAUNCHED();
  return TSDF_OK;
}

int launch_instance_origin(const float* d_depth, const uint8_t* d_mask, int H, int W, float fx,
                           float cx, float fy, float cy, float* d_out3, cudaStream_t st) {
  const float init[3] = {1000.0f, 1000.0f, 1000.0f};
  CU_CHECK(cudaMemcpyAsync(d_out3, init, sizeof(init), cudaMemcpyHostToDevice, st));
  k_instance_origin<<<148, 256, 0, st>>>(d_depth, d_mask, H, W, cx, cy, 1.0f / fx, 1.0f / fy, d_out3);
  CU_LAUNCHED();
  return TSDF_OK;
}
