// selftest.cu — on-device brute-force check that the exactmath.cuh forms reproduce the IEEE
// operators bit for bit (the claim the parity of the hot loop rests on).  Driven by
// tests/test_exactmath_gpu.py through tsdf_selftest_exactmath; not used by the product path.
#include "common.cuh"
#include "exactmath.cuh"

namespace {

__device__ __forceinline__ uint64_t splitmix(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// float with random mantissa, exponent uniformly in [emin, emax], random sign if allowed
__device__ __forceinline__ float rnd_float(uint64_t& s, int emin, int emax, bool neg, int flavour) {
  uint64_t r = splitmix(s);
  uint32_t man = (uint32_t)r & 0x7FFFFF;
  switch (flavour & 7) {  // adversarial mantissas: all ones, near 1.0, near 2.0, few bits
    case 1: man = 0x7FFFFF - ((uint32_t)(r >> 23) & 7); break;
    case 2: man = (uint32_t)(r >> 23) & 7; break;
    case 3: man &= 0x7FF000; break;
    case 4: man |= 0x7FF000; break;
    default: break;
  }
  int e = emin + (int)((r >> 32) % (uint64_t)(emax - emin + 1));
  uint32_t bits = ((uint32_t)(e + 127) << 23) | man;
  if (neg && (r >> 63)) bits |= 0x80000000u;
  return __uint_as_float(bits);
}

// fails[0]: rcp, [1]: div, [2]: round, [3]: blend_exact, [4]: blend_intw
__global__ void k_selftest(int64_t iters, uint64_t seed, unsigned long long* fails) {
  uint64_t s = seed ^ ((uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0xD1342543DE82EF95ull);
  unsigned f0 = 0, f1 = 0, f2 = 0, f3 = 0;
  for (int64_t it = 0; it < iters; ++it) {
    const int fl = (int)(splitmix(s) & 63);
    // 1+2: reciprocal and quotient; b in the guarded window, a anywhere the quotient is normal
    float b = rnd_float(s, -60, 60, false, fl);
    float y = em::rcp_rn(b);
    if (__float_as_uint(y) != __float_as_uint(__frcp_rn(b))) ++f0;
    float a = rnd_float(s, -40, 40, true, fl >> 3);
    if ((fl & 0x30) == 0x10) a = b * (float)(int)(splitmix(s) & 1023);        // exact quotients
    if ((fl & 0x30) == 0x20) a = __uint_as_float(__float_as_uint(b) ^ 1u);    // a = b -/+ 1 ulp
    float q = em::div_rn(a, b, y);
    if (__float_as_uint(q) != __float_as_uint(__fdiv_rn(a, b))) ++f1;
    // 3: rounding.  u in (-0.5, 2^21), many near k + 0.5
    uint64_t r = splitmix(s);
    float u = (float)((double)(r >> 11) * (1.0 / 9007199254740992.0) * 2097151.0);
    if (r & 1) u = (float)(int)(r & 8191) + 0.5f;
    if ((r & 6) == 2) u = __uint_as_float(__float_as_uint(u) + (int)((r >> 3) & 3) - 1);
    if ((r & 24) == 8) u = -0.49999997f * (float)((r >> 5) & 1);
    if (u > -0.5f && u < 2097151.0f) {
      if (em::round_half_away_nonneg_int(u) != (int)roundf(u)) ++f2;
    }
    // 4: general colour blend vs the literal spelling
    uint64_t c = splitmix(s);
    float oc = (float)(uint32_t)(c & 0xFFFFFF);
    float n0 = (float)((c >> 24) & 255), n1 = (float)((c >> 32) & 255), n2 = (float)((c >> 40) & 255);
    float w0 = rnd_float(s, -8, 14, false, fl >> 2);
    if (c >> 63) w0 = (float)((c >> 48) & 4095) * 0.25f;
    float obs = rnd_float(s, -6, 6, false, 0);
    if ((c >> 62) & 1) obs = 1.0f;
    float w1 = w0 + obs;
    float lit = em::blend_literal(oc, n0, n1, n2, w0, w1);
    float ex = em::blend_exact(oc, n0, n1, n2, w0, w1, em::rcp_rn(w1));
    if (__float_as_uint(lit) != __float_as_uint(ex)) ++f3;
  }
  if (f0) atomicAdd(fails + 0, (unsigned long long)f0);
  if (f1) atomicAdd(fails + 1, (unsigned long long)f1);
  if (f2) atomicAdd(fails + 2, (unsigned long long)f2);
  if (f3) atomicAdd(fails + 3, (unsigned long long)f3);
}

// Exhaustive: every old channel o, new channel n in [0,255] and every integer weight
// w0 in [0, 4093] (w1 = w0 + 1 <= 4094): blend_intw == blend_literal.
__global__ void k_selftest_intw(unsigned long long* fails) {
  const int w0i = blockIdx.x;                       // 0..4093
  const float w0 = (float)w0i, w1 = w0 + 1.0f;
  const float yw = em::rcp_rn(w1);
  unsigned f = 0;
  for (int on = threadIdx.x; on < 65536; on += blockDim.x) {
    const int o = on >> 8, n = on & 255;
    // the three channels carry (o,n), (255-o, n), (o, 255-n) to exercise all byte lanes
    float oc = (float)((o << 16) | ((255 - o) << 8) | o);
    float n2 = (float)n, n1 = (float)n, n0 = (float)(255 - n);
    float lit = em::blend_literal(oc, n0, n1, n2, w0, w1);
    float fast = em::blend_intw(oc, n0, n1, n2, w0, yw);
    if (__float_as_uint(lit) != __float_as_uint(fast)) ++f;
  }
  if (f) atomicAdd(fails + 4, (unsigned long long)f);
}

}  // namespace

extern "C" int tsdf_selftest_exactmath(int device, int64_t iters_per_thread, uint64_t seed, int64_t fails_out[5]) {
  REQUIRE(fails_out && iters_per_thread > 0, "bad argument");
  int ndev = 0;
  int rc = tsdf_device_count(&ndev);
  if (rc) return rc;
  DeviceGuard g(device);
  unsigned long long* d = nullptr;
  CU_CHECK(cudaMalloc(&d, 5 * sizeof(unsigned long long)));
  CU_CHECK(cudaMemset(d, 0, 5 * sizeof(unsigned long long)));
  k_selftest<<<148 * 8, 256>>>(iters_per_thread, seed, d);
  k_selftest_intw<<<4094, 256>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  unsigned long long h[5] = {0, 0, 0, 0, 0};
  if (e == cudaSuccess) e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) { tsdf_set_error("selftest: %s", cudaGetErrorString(e)); return TSDF_E_CUDA; }
  for (int i = 0; i < 5; ++i) fails_out[i] = (int64_t)h[i];
  return TSDF_OK;
}
