// Synthetic inputs generated on device: Philox4x32-10 keyed on
// (seed, global row, column), so a shard of rows is identical whatever the GPU count.
// The normal deviate is an 8-term Irwin-Hall sum of 16-bit uniforms: integer arithmetic plus
// one exact int->float conversion and one fp32 multiply, so oracle/synth.py reproduces it
// bit-for-bit on the CPU.
#include "common.cuh"

#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

struct u32x4 {
  uint32_t x, y, z, w;
};

__device__ __forceinline__ u32x4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(PHILOX_M0, c0), lo0 = PHILOX_M0 * c0;
    uint32_t hi1 = __umulhi(PHILOX_M1, c2), lo1 = PHILOX_M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += PHILOX_W0; k1 += PHILOX_W1;
  }
  return {c0, c1, c2, c3};
}

// 1 / sqrt(8 * (65536^2 - 1) / 12) rounded to fp32
#define IH_SCALE 0x1.3988e2p-16f

__device__ __forceinline__ float ih_normal(u32x4 r) {
  int s = (int)(r.x & 0xffffu) + (int)(r.x >> 16) + (int)(r.y & 0xffffu) + (int)(r.y >> 16) +
          (int)(r.z & 0xffffu) + (int)(r.z >> 16) + (int)(r.w & 0xffffu) + (int)(r.w >> 16);
  return __fmul_rn((float)(s - 262140), IH_SCALE);
}

__device__ __forceinline__ float znorm(uint64_t seed, uint64_t row, uint32_t col, uint32_t stream) {
  return ih_normal(philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), col, stream, (uint32_t)seed,
                                 (uint32_t)(seed >> 32)));
}
__device__ __forceinline__ uint32_t urand32(uint64_t seed, uint64_t row, uint32_t col, uint32_t stream) {
  return philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), col, stream, (uint32_t)seed, (uint32_t)(seed >> 32)).x;
}

__global__ void synth_mixture_kernel(float* __restrict__ X, int64_t n, int64_t row0, int d, int64_t ld, uint64_t seed,
                                     int ncomp, float cscale, const float* __restrict__ col_scale) {
  int64_t total = n * ld;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / ld;
    int c = (int)(i - r * ld);
    float v = 0.f;
    if (c < d) {
      uint64_t grow = (uint64_t)(row0 + r);
      uint32_t comp = urand32(seed, grow, 0u, 1u) % (uint32_t)ncomp;
      float mu = __fmul_rn(cscale, znorm(seed, comp, (uint32_t)c, 2u));
      v = __fadd_rn(mu, znorm(seed, grow, (uint32_t)c, 0u));
      if (col_scale) v = __fmul_rn(v, col_scale[c]);
    }
    X[i] = v;
  }
}

// one warp per row: coalesced row writes, margins by warp reduction
__global__ void synth_logistic_kernel(float* __restrict__ X, float* __restrict__ y, int64_t n, int64_t row0, int d,
                                      int64_t ld, uint64_t seed, const float* __restrict__ col_scale,
                                      const float* __restrict__ w_true, const float* __restrict__ b_true, int K) {
  int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int KM = (K == 2) ? 1 : K;
  for (int64_t r = warp; r < n; r += nwarps) {
    uint64_t grow = (uint64_t)(row0 + r);
    float m[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) m[k] = 0.f;
    for (int c = lane; c < (int)ld; c += 32) {
      float v = 0.f;
      if (c < d) {
        v = znorm(seed, grow, (uint32_t)c, 0u);
        if (col_scale) v = __fmul_rn(v, col_scale[c]);
#pragma unroll
        for (int k = 0; k < 32; ++k)
          if (k < KM) m[k] = fmaf(v, w_true[(size_t)k * d + c], m[k]);
      }
      X[r * ld + c] = v;
    }
    float u = (float)(urand32(seed, grow, 0u, 3u) >> 8) * (1.0f / 16777216.0f);
    float label = 0.f;
    if (KM == 1) {
      float mm = warp_sum(m[0]) + b_true[0];
      label = (u < 1.f / (1.f + expf(-mm))) ? 1.f : 0.f;
    } else {
      float mx = -INFINITY;
#pragma unroll
      for (int k = 0; k < 32; ++k)
        if (k < KM) {
          m[k] = warp_sum(m[k]) + b_true[k];
          mx = fmaxf(mx, m[k]);
        }
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 32; ++k)
        if (k < KM) {
          m[k] = expf(m[k] - mx);
          s += m[k];
        }
      float acc = 0.f, t = u * s;
      int pick = KM - 1;
      bool done = false;
#pragma unroll
      for (int k = 0; k < 32; ++k)
        if (k < KM && !done) {
          acc += m[k];
          if (t < acc) {
            pick = k;
            done = true;
          }
        }
      label = (float)pick;
    }
    if (lane == 0) y[r] = label;
  }
}

extern "C" int b2f_synth_mixture(float* X, int64_t n, int64_t row0, int d, int64_t ld, uint64_t seed, int ncomp,
                                 float center_scale, const float* col_scale, b2f_stream_t stream) {
  B2F_REQUIRE(X && n >= 0 && d > 0 && ld >= d && ncomp > 0, "b2f_synth_mixture: bad arguments");
  if (n == 0) return B2F_OK;
  int grid = b2f_sm_count() * 16;
  synth_mixture_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(X, n, row0, d, ld, seed, ncomp, center_scale, col_scale);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_synth_logistic(float* X, float* y, int64_t n, int64_t row0, int d, int64_t ld, uint64_t seed,
                                  const float* col_scale, const float* w_true, const float* b_true, int n_classes,
                                  b2f_stream_t stream) {
  B2F_REQUIRE(X && y && n >= 0 && d > 0 && ld >= d && w_true && b_true, "b2f_synth_logistic: bad arguments");
  B2F_REQUIRE(n_classes >= 2 && n_classes <= 32, "b2f_synth_logistic: n_classes must be in [2, 32]");
  if (n == 0) return B2F_OK;
  int grid = b2f_sm_count() * 8;
  synth_logistic_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(X, y, n, row0, d, ld, seed, col_scale, w_true, b_true,
                                                          n_classes);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
