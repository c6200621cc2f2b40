// K3 / K4: fused LogisticRegression loss + gradient passes and prediction.
// Replaces the treeAggregate of Binary/Multinomial LogisticAggregator inside RDDLossFunction.calculate
//   reference: orangecontrib/spark/base/spark_ml_estimator.py:22 (module = classification,
//   orangecontrib/spark/widgets/ml/spark_ml_classification.py:15); semantics: SURVEY.md A.3.
// Standardisation never touches X: the caller passes beta = coef / sigma and rescales the raw
// gradient sum by 1 / sigma afterwards.
#include <stdlib.h>

#include "common.cuh"

#define LR_THREADS 256
#define LR_WARPS (LR_THREADS / 32)

__device__ __forceinline__ float log1pexp_f(float m) {
  // stable log(1 + e^m)
  const float t = log1pf(expf(-fabsf(m)));
  return m > 0.f ? m + t : t;
}

__device__ __forceinline__ void split_f64(double v, float& hi, float& lo) {
  hi = (float)v;
  lo = (float)(v - (double)hi);
}

// ---------------------------------------------------------------------------------------------
// K3 binary.  One warp streams whole rows: lane l owns columns 4l..4l+3 of every 128-column chunk.
// beta is held as hi + lo fp32 pairs so the margin sees the fp64 weights to ~1e-14 relative;
// the gradient is accumulated in fp32 registers over short chains (LR_FLUSH row batches) and
// folded into per-warp fp64 shared-memory accumulators, then into a per-CTA fp64 partial.
// ---------------------------------------------------------------------------------------------
#define LR_FLUSH 16

template <int NCH>
__global__ void __launch_bounds__(LR_THREADS, (NCH >= 8) ? 1 : 2)
    lr_binary_kernel(const float* __restrict__ X, const float* __restrict__ y, const float* __restrict__ w, int64_t n,
                     int d, int64_t ld, const double* __restrict__ beta, const double* __restrict__ icpt,
                     double* __restrict__ partials) {
  constexpr int LR_BATCH = (NCH <= 2) ? 4 : ((NCH == 4) ? 2 : 1);  // rows in flight per warp
  extern __shared__ double acc64[];  // [LR_WARPS][NCH*128]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* my64 = acc64 + (size_t)warp * NCH * 128;
  for (int i = lane; i < NCH * 128; i += 32) my64[i] = 0.0;
  __syncwarp();  // lanes zero one layout (i = lane + 32 j) and accumulate in another (lane * 4 + j)

  float4 bh[NCH], bl[NCH], g[NCH];
#pragma unroll
  for (int ch = 0; ch < NCH; ++ch) {
    const int c = ch * 128 + lane * 4;
    float h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double v = (c + j < d) ? beta[c + j] : 0.0;
      split_f64(v, h[j], l[j]);
    }
    bh[ch] = make_float4(h[0], h[1], h[2], h[3]);
    bl[ch] = make_float4(l[0], l[1], l[2], l[3]);
    g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float b0h, b0l;
  split_f64(icpt[0], b0h, b0l);

  double loss = 0.0, smult = 0.0, sw = 0.0;
  const int64_t gw = (int64_t)blockIdx.x * LR_WARPS + warp;
  const int64_t nw = (int64_t)gridDim.x * LR_WARPS;
  const int64_t nbatch = (n + LR_BATCH - 1) / LR_BATCH;
  int since_flush = 0;
  for (int64_t b = gw; b < nbatch; b += nw) {
    const int64_t r0 = b * LR_BATCH;
    float4 x[LR_BATCH][NCH];
    float yy[LR_BATCH], ww[LR_BATCH];
#pragma unroll
    for (int i = 0; i < LR_BATCH; ++i) {
      const int64_t r = r0 + i;
      const bool live = r < n;
      const float* xr = X + (live ? r : (n - 1)) * ld;
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        const int c = ch * 128 + lane * 4;
        x[i][ch] = (c < ld) ? ldg_stream(reinterpret_cast<const float4*>(xr + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      yy[i] = live ? __ldg(y + r) : 0.f;
      ww[i] = live ? (w ? __ldg(w + r) : 1.f) : 0.f;
    }
    if constexpr (LR_BATCH == 4) {
      // per-lane partial margins of the 4 rows (hi and lo parts of beta; the lo part only has to remove the
      // SYSTEMATIC error of rounding beta to fp32, so adding it per lane is enough)
      float pm[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float mh = 0.f, ml = 0.f;
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
          mh = fmaf(x[i][ch].x, bh[ch].x, mh); mh = fmaf(x[i][ch].y, bh[ch].y, mh);
          mh = fmaf(x[i][ch].z, bh[ch].z, mh); mh = fmaf(x[i][ch].w, bh[ch].w, mh);
          ml = fmaf(x[i][ch].x, bl[ch].x, ml); ml = fmaf(x[i][ch].y, bl[ch].y, ml);
          ml = fmaf(x[i][ch].z, bl[ch].z, ml); ml = fmaf(x[i][ch].w, bl[ch].w, ml);
        }
        pm[i] = mh + ml;
      }
      // transposed warp reduction: 6 shuffles leave every 8-lane group with the full margin of ONE row
      // (row = 2 * bit4 + bit3 of the lane id), so sigmoid / softplus run once for all four rows
      const bool up16 = lane & 16, up8 = lane & 8;
      float k0 = (up16 ? pm[2] : pm[0]) + __shfl_xor_sync(0xffffffffu, up16 ? pm[0] : pm[2], 16);
      float k1 = (up16 ? pm[3] : pm[1]) + __shfl_xor_sync(0xffffffffu, up16 ? pm[1] : pm[3], 16);
      float v = (up8 ? k1 : k0) + __shfl_xor_sync(0xffffffffu, up8 ? k0 : k1, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      const int myrow = (up16 ? 2 : 0) + (up8 ? 1 : 0);
      const float yi = myrow == 0 ? yy[0] : myrow == 1 ? yy[1] : myrow == 2 ? yy[2] : yy[3];
      const float wi = myrow == 0 ? ww[0] : myrow == 1 ? ww[1] : myrow == 2 ? ww[2] : ww[3];
      const float m = (v + b0l) + b0h;
      const float e = expf(-fabsf(m));                 // one exponential serves sigmoid and softplus
      const float inv = 1.f / (1.f + e);  // IEEE division: the gradient must match the loss to ~1e-8 near convergence
      const float p = m >= 0.f ? inv : e * inv;
      const float softplus = fmaxf(m, 0.f) + log1pf(e);
      const float mymult = wi * (p - yi);
      if ((lane & 7) == 0) {
        loss += (double)(wi * (softplus - yi * m));
        smult += (double)mymult;
        sw += (double)wi;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float mult = __shfl_sync(0xffffffffu, mymult, i * 8);
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
          g[ch].x = fmaf(mult, x[i][ch].x, g[ch].x); g[ch].y = fmaf(mult, x[i][ch].y, g[ch].y);
          g[ch].z = fmaf(mult, x[i][ch].z, g[ch].z); g[ch].w = fmaf(mult, x[i][ch].w, g[ch].w);
        }
      }
    } else {
#pragma unroll
    for (int i = 0; i < LR_BATCH; ++i) {
      float mh = 0.f, ml = 0.f;
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        mh = fmaf(x[i][ch].x, bh[ch].x, mh); mh = fmaf(x[i][ch].y, bh[ch].y, mh);
        mh = fmaf(x[i][ch].z, bh[ch].z, mh); mh = fmaf(x[i][ch].w, bh[ch].w, mh);
        ml = fmaf(x[i][ch].x, bl[ch].x, ml); ml = fmaf(x[i][ch].y, bl[ch].y, ml);
        ml = fmaf(x[i][ch].z, bl[ch].z, ml); ml = fmaf(x[i][ch].w, bl[ch].w, ml);
      }
      const float m = (warp_sum(mh + ml) + b0l) + b0h;
      const float e = expf(-fabsf(m));
      const float inv = 1.f / (1.f + e);  // IEEE division: the gradient must match the loss to ~1e-8 near convergence
      const float p = m >= 0.f ? inv : e * inv;
      const float mult = ww[i] * (p - yy[i]);
      if (lane == 0) {
        loss += (double)(ww[i] * (fmaxf(m, 0.f) + log1pf(e) - yy[i] * m));
        smult += (double)mult;
        sw += (double)ww[i];
      }
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        g[ch].x = fmaf(mult, x[i][ch].x, g[ch].x); g[ch].y = fmaf(mult, x[i][ch].y, g[ch].y);
        g[ch].z = fmaf(mult, x[i][ch].z, g[ch].z); g[ch].w = fmaf(mult, x[i][ch].w, g[ch].w);
      }
    }
    }
    if (++since_flush == LR_FLUSH) {
      since_flush = 0;
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        double* p = my64 + ch * 128 + lane * 4;
        p[0] += (double)g[ch].x; p[1] += (double)g[ch].y; p[2] += (double)g[ch].z; p[3] += (double)g[ch].w;
        g[ch] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
#pragma unroll
  for (int ch = 0; ch < NCH; ++ch) {
    double* p = my64 + ch * 128 + lane * 4;
    p[0] += (double)g[ch].x; p[1] += (double)g[ch].y; p[2] += (double)g[ch].z; p[3] += (double)g[ch].w;
  }
  __shared__ double red[3][LR_WARPS];
  loss = warp_sum(loss);
  smult = warp_sum(smult);
  sw = warp_sum(sw);
  if (lane == 0) {
    red[0][warp] = loss;
    red[1][warp] = smult;
    red[2][warp] = sw;
  }
  __syncthreads();
  double* out = partials + (size_t)blockIdx.x * (d + 3);
  for (int c = threadIdx.x; c < d; c += LR_THREADS) {
    double s = 0.0;
#pragma unroll
    for (int wv = 0; wv < LR_WARPS; ++wv) s += acc64[(size_t)wv * NCH * 128 + c];
    out[c] = s;
  }
  if (threadIdx.x == 0) {
    double a = 0, bsum = 0, c = 0;
    for (int wv = 0; wv < LR_WARPS; ++wv) {
      a += red[0][wv];
      bsum += red[1][wv];
      c += red[2][wv];
    }
    out[d] = bsum;       // sum mult       -> intercept gradient
    out[d + 1] = a;      // sum w * loss
    out[d + 2] = c;      // sum w
  }
}

// deterministic fold of per-CTA partials [nparts][len] into out[len] (out is accumulated)
__global__ void lr_fold_kernel(const double* __restrict__ partials, int nparts, int len, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  double s = 0.0;
  for (int p = 0; p < nparts; ++p) s += partials[(size_t)p * len + i];
  out[i] += s;
}

// ---------------------------------------------------------------------------------------------
// K4 multinomial, CUDA-core version.  CTA tile = MN_ROWS rows staged in shared memory by cp.async
// (double buffered); GEMM1 margins = X_tile . W^T, thread-per-row softmax, GEMM2 G += (P - Y)^T X_tile
// with the K x d gradient held in registers for the CTA's lifetime (fp32, short chains) and flushed
// into a CTA-private fp64 partial every MN_FLUSH tiles.
// ---------------------------------------------------------------------------------------------
#define MN_ROWS 32
#define MN_FLUSH 8

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
  const int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

template <int KP, int NJ>
__global__ void __launch_bounds__(LR_THREADS, 1)
    lr_multinomial_kernel(const float* __restrict__ X, const float* __restrict__ y, const float* __restrict__ w,
                          int64_t n, int d, int64_t ld, int K, const double* __restrict__ beta,
                          const double* __restrict__ icpt, double* __restrict__ partials) {
  extern __shared__ __align__(16) float sm[];
  const int xs_stride = (int)ld + 4;  // +4 floats: rows shift by 4 banks
  const int ws_stride = ((d + 3) & ~3) + 4;
  float* Ws = sm;                                 // [KP][ws_stride]
  float* Xs = Ws + (size_t)KP * ws_stride;        // [2][MN_ROWS][xs_stride]
  float* Ps = Xs + (size_t)2 * MN_ROWS * xs_stride;  // [MN_ROWS][KP]
  float* Bs = Ps + MN_ROWS * KP;                  // [KP] intercepts
  float* Ys = Bs + KP;                            // [2][MN_ROWS] labels, [2][MN_ROWS] weights
  const int tid = threadIdx.x;
  const int plen = K * d + K + 2;
  double* part = partials + (size_t)blockIdx.x * plen;
  for (int i = tid; i < plen; i += LR_THREADS) part[i] = 0.0;

  for (int i = tid; i < KP * ws_stride; i += LR_THREADS) {
    const int k = i / ws_stride, c = i - k * ws_stride;
    Ws[i] = (k < K && c < d) ? (float)beta[(size_t)k * d + c] : 0.f;
  }
  for (int i = tid; i < KP; i += LR_THREADS) Bs[i] = (i < K) ? (float)icpt[i] : -INFINITY;

  float G[NJ][KP];
#pragma unroll
  for (int a = 0; a < NJ; ++a)
#pragma unroll
    for (int k = 0; k < KP; ++k) G[a][k] = 0.f;
  float gb = 0.f;  // intercept gradient for class tid (tid < K)
  double loss = 0.0, sw = 0.0;

  const int64_t ntiles = (n + MN_ROWS - 1) / MN_ROWS;
  const int vec_per_row = (int)(ld / 4);
  auto issue = [&](int64_t tile, int buf) {
    const int64_t r0 = tile * MN_ROWS;
    float* xb = Xs + (size_t)buf * MN_ROWS * xs_stride;
    for (int i = tid; i < MN_ROWS * vec_per_row; i += LR_THREADS) {
      const int rl = i / vec_per_row, c4 = (i - rl * vec_per_row) * 4;
      const int64_t r = r0 + rl;
      const bool live = r < n;
      cp_async16(xb + (size_t)rl * xs_stride + c4, X + (live ? r : 0) * ld + c4, live);
    }
    if (tid < MN_ROWS) {
      const int64_t r = r0 + tid;
      const bool live = r < n;
      Ys[buf * MN_ROWS + tid] = live ? y[r] : 0.f;
      Ys[(2 + buf) * MN_ROWS + tid] = live ? (w ? w[r] : 1.f) : 0.f;
    }
    cp_async_commit();
  };

  int64_t tile = blockIdx.x;
  int buf = 0, since_flush = 0;
  if (tile < ntiles) issue(tile, 0);
  for (; tile < ntiles; tile += gridDim.x, buf ^= 1) {
    const int64_t nxt = tile + gridDim.x;
    if (nxt < ntiles) {
      issue(nxt, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const float* xb = Xs + (size_t)buf * MN_ROWS * xs_stride;
    // ---- GEMM1: thread (row rl = tid / 8, class lane kg = tid % 8) -> classes kg, kg+8, ...
    {
      const int rl = tid >> 3, kg = tid & 7;
      float m[KP / 8];
#pragma unroll
      for (int q = 0; q < KP / 8; ++q) m[q] = 0.f;
      const float* xr = xb + (size_t)rl * xs_stride;
      for (int c = 0; c < d; c += 4) {
        const float4 xv = *reinterpret_cast<const float4*>(xr + c);
#pragma unroll
        for (int q = 0; q < KP / 8; ++q) {
          const float4 wv = *reinterpret_cast<const float4*>(Ws + (size_t)(kg + 8 * q) * ws_stride + c);
          m[q] = fmaf(xv.x, wv.x, m[q]); m[q] = fmaf(xv.y, wv.y, m[q]);
          m[q] = fmaf(xv.z, wv.z, m[q]); m[q] = fmaf(xv.w, wv.w, m[q]);
        }
      }
#pragma unroll
      for (int q = 0; q < KP / 8; ++q) Ps[rl * KP + kg + 8 * q] = m[q] + Bs[kg + 8 * q];
    }
    __syncthreads();
    // ---- softmax / loss: one thread per row
    if (tid < MN_ROWS) {
      float* pr = Ps + tid * KP;
      const float wi = Ys[(2 + buf) * MN_ROWS + tid];
      const int yi = (int)Ys[buf * MN_ROWS + tid];
      float mx = 0.f;  // "subtract max if > 0"
      for (int k = 0; k < K; ++k) mx = fmaxf(mx, pr[k]);
      const float my = pr[min(max(yi, 0), K - 1)];
      float s = 0.f;
      for (int k = 0; k < K; ++k) {
        const float e = expf(pr[k] - mx);
        pr[k] = e;
        s += e;
      }
      const float inv = 1.f / s;
      for (int k = 0; k < KP; ++k) {
        float v = 0.f;
        if (k < K) v = wi * (pr[k] * inv - (k == yi ? 1.f : 0.f));
        pr[k] = v;
      }
      loss += (double)(wi * (logf(s) + mx - my));
      sw += (double)wi;
    }
    __syncthreads();
    // ---- GEMM2: thread owns columns tid + 256*a, all classes
    for (int rl = 0; rl < MN_ROWS; ++rl) {
      const float* pr = Ps + rl * KP;
      float xv[NJ];
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        const int c = tid + LR_THREADS * a;
        xv[a] = (c < d) ? xb[(size_t)rl * xs_stride + c] : 0.f;
      }
#pragma unroll
      for (int k4 = 0; k4 < KP; k4 += 4) {
        const float4 pv = *reinterpret_cast<const float4*>(pr + k4);
#pragma unroll
        for (int a = 0; a < NJ; ++a) {
          G[a][k4] = fmaf(pv.x, xv[a], G[a][k4]);
          G[a][k4 + 1] = fmaf(pv.y, xv[a], G[a][k4 + 1]);
          G[a][k4 + 2] = fmaf(pv.z, xv[a], G[a][k4 + 2]);
          G[a][k4 + 3] = fmaf(pv.w, xv[a], G[a][k4 + 3]);
        }
      }
      if (tid < K) gb += pr[tid];
    }
    if (++since_flush == MN_FLUSH) {
      since_flush = 0;
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        const int c = tid + LR_THREADS * a;
#pragma unroll
        for (int k = 0; k < KP; ++k) {
          if (k < K && c < d) part[(size_t)k * d + c] += (double)G[a][k];
          G[a][k] = 0.f;
        }
      }
      if (tid < K) {
        part[(size_t)K * d + tid] += (double)gb;
        gb = 0.f;
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < NJ; ++a) {
    const int c = tid + LR_THREADS * a;
#pragma unroll
    for (int k = 0; k < KP; ++k)
      if (k < K && c < d) part[(size_t)k * d + c] += (double)G[a][k];
  }
  if (tid < K) part[(size_t)K * d + tid] += (double)gb;
  __shared__ double red[2][MN_ROWS];
  if (tid < MN_ROWS) {
    red[0][tid] = loss;
    red[1][tid] = sw;
  }
  __syncthreads();
  if (tid == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < MN_ROWS; ++i) {
      a += red[0][i];
      b += red[1][i];
    }
    part[(size_t)K * d + K] = a;
    part[(size_t)K * d + K + 1] = b;
  }
}

// ---------------------------------------------------------------------------------------------
// prediction (A.3): warp per row
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lr_predict_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld,
                                                         int K, const double* __restrict__ beta,
                                                         const double* __restrict__ icpt, float threshold,
                                                         float* __restrict__ raw, float* __restrict__ prob,
                                                         float* __restrict__ pred) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int KM = (K == 2) ? 1 : K;
  for (int64_t r = warp; r < n; r += nwarps) {
    const float* xr = X + r * ld;
    float m[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) m[k] = 0.f;
    for (int c = lane; c < d; c += 32) {
      const float xv = xr[c];
#pragma unroll
      for (int k = 0; k < 32; ++k)
        if (k < KM) m[k] = fmaf(xv, (float)beta[(size_t)k * d + c], m[k]);
    }
#pragma unroll
    for (int k = 0; k < 32; ++k)
      if (k < KM) m[k] = warp_sum(m[k]) + (float)icpt[k];
    if (lane == 0) {
      if (KM == 1) {
        const float p = 1.f / (1.f + expf(-m[0]));
        if (raw) { raw[r * 2] = -m[0]; raw[r * 2 + 1] = m[0]; }
        if (prob) { prob[r * 2] = 1.f - p; prob[r * 2 + 1] = p; }
        if (pred) pred[r] = p > threshold ? 1.f : 0.f;
      } else {
        float mx = -INFINITY;
        int arg = 0;
#pragma unroll
        for (int k = 0; k < 32; ++k)
          if (k < KM && m[k] > mx) { mx = m[k]; arg = k; }
        float s = 0.f;
        float e[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          e[k] = (k < KM) ? expf(m[k] - mx) : 0.f;
          s += e[k];
        }
#pragma unroll
        for (int k = 0; k < 32; ++k)
          if (k < KM) {
            if (raw) raw[r * KM + k] = m[k];
            if (prob) prob[r * KM + k] = e[k] / s;
          }
        if (pred) pred[r] = (float)arg;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
static int lr_grid_binary() { return b2f_sm_count() * 2; }
static int lr_grid_multi() { return b2f_sm_count(); }

extern "C" size_t b2f_lr_scratch_bytes(int d, int n_classes) {
  if (d <= 0 || n_classes < 2) return 0;
  const int K = (n_classes == 2) ? 1 : n_classes;
  const size_t len = (size_t)K * d + K + 2;
  const size_t parts = (size_t)((K == 1) ? lr_grid_binary() : lr_grid_multi());
  size_t bytes = b2f_align_up(parts * len * sizeof(double), 1024);
  if (K > 1) bytes += lr_mn_tc_extra_scratch(d);  // W' as TF32 hi / lo for the tensor-core kernel
  return bytes;
}

template <int NCH>
static int lr_binary_launch(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld,
                            const double* beta, const double* icpt, double* partials, int grid, cudaStream_t st) {
  size_t smem = (size_t)LR_WARPS * NCH * 128 * sizeof(double);
  B2F_CUDA(b2f_smem_attr(lr_binary_kernel<NCH>, smem));
  lr_binary_kernel<NCH><<<grid, LR_THREADS, smem, st>>>(X, y, w, n, d, ld, beta, icpt, partials);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_lr_binary_lossgrad(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld,
                                      const double* beta, const double* icpt, double* out, void* scratch,
                                      b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && ld >= d && ld % 4 == 0, "b2f_lr_binary_lossgrad: need ld >= d, ld %% 4 == 0");
  B2F_REQUIRE(ld <= 1024, "b2f_lr_binary_lossgrad: d (padded) must be <= 1024, got %lld", (long long)ld);
  if (n == 0) return B2F_OK;  // an empty row shard adds nothing (X / y may be NULL)
  B2F_REQUIRE(X && y && beta && icpt && out && scratch, "b2f_lr_binary_lossgrad: null pointer");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_lr_binary_lossgrad: X must be 16-byte aligned");
  cudaStream_t st = b2f_cs(stream);
  const int grid = lr_grid_binary();
  double* partials = static_cast<double*>(scratch);
  const int nch = (int)((ld + 127) / 128);
  int rc;
  if (nch <= 1) rc = lr_binary_launch<1>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  else if (nch <= 2) rc = lr_binary_launch<2>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  else if (nch <= 4) rc = lr_binary_launch<4>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  else rc = lr_binary_launch<8>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  if (rc != B2F_OK) return rc;
  const int len = d + 3;
  lr_fold_kernel<<<(len + 127) / 128, 128, 0, st>>>(partials, grid, len, out);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

template <int KP, int NJ>
static int lr_multi_launch(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld, int K,
                           const double* beta, const double* icpt, double* partials, int grid, cudaStream_t st) {
  size_t floats = (size_t)KP * (((d + 3) & ~3) + 4) + (size_t)2 * MN_ROWS * (ld + 4) + (size_t)MN_ROWS * KP + KP + 4 * MN_ROWS;
  size_t smem = floats * sizeof(float);
  B2F_REQUIRE(smem <= 227 * 1024, "b2f_lr_multinomial_lossgrad: K=%d, d=%d needs %zu bytes of shared memory (> 227 KB)", K,
              d, smem);
  B2F_CUDA(b2f_smem_attr(lr_multinomial_kernel<KP, NJ>, smem));
  lr_multinomial_kernel<KP, NJ><<<grid, LR_THREADS, smem, st>>>(X, y, w, n, d, ld, K, beta, icpt, partials);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_lr_multinomial_lossgrad(const float* X, const float* y, const float* w, int64_t n, int d,
                                           int64_t ld, int n_classes, const double* beta, const double* icpt,
                                           double* out, void* scratch, b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && ld >= d && ld % 4 == 0, "b2f_lr_multinomial_lossgrad: need ld >= d, ld %% 4 == 0");
  B2F_REQUIRE(n_classes >= 2 && n_classes <= 32, "b2f_lr_multinomial_lossgrad: n_classes must be in [2, 32]");
  if (n == 0) return B2F_OK;  // an empty row shard adds nothing (X / y may be NULL)
  B2F_REQUIRE(X && y && beta && icpt && out && scratch, "b2f_lr_multinomial_lossgrad: null pointer");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_lr_multinomial_lossgrad: X must be 16-byte aligned");
  cudaStream_t st = b2f_cs(stream);
  const int K = n_classes;
  const int kp = (K + 7) / 8 * 8;
  const int nj = (d + LR_THREADS - 1) / LR_THREADS;
  B2F_REQUIRE(nj <= 4 && kp * (nj == 3 ? 4 : nj) <= 64,
              "b2f_lr_multinomial_lossgrad: unsupported shape K=%d d=%d (need ceil(K/8)*8 * ceil(d/256) <= 64)", K, d);
  const int grid = lr_grid_multi();
  double* partials = static_cast<double*>(scratch);
  bool use_tc = lr_mn_tc_supported(d, ld, K);
#ifdef B2F_DEV
  if (getenv("B2F_LR_NO_TC")) use_tc = false;
#endif
  if (use_tc) {
    const size_t plen_ = (size_t)K * d + K + 2;
    void* wbuf = static_cast<char*>(scratch) + b2f_align_up((size_t)grid * plen_ * sizeof(double), 1024);
    int rc_ = lr_mn_tc_launch(X, y, w, n, d, K, beta, icpt, partials, grid, wbuf, st);
    if (rc_ != B2F_OK) return rc_;
    lr_fold_kernel<<<((int)plen_ + 127) / 128, 128, 0, st>>>(partials, grid, (int)plen_, out);
    B2F_LAUNCH_CHECK();
    return B2F_OK;
  }
  const int kp0 = (K + 7) / 8 * 8;
  (void)kp0;
  int rc = B2F_EUNSUPPORTED;
#define MN_CASE(KPV, NJV)                                                                            \
  if (kp == KPV && (nj == NJV || (NJV == 4 && nj == 3)))                                             \
    rc = lr_multi_launch<KPV, NJV>(X, y, w, n, d, ld, K, beta, icpt, partials, grid, st);
  MN_CASE(8, 1) MN_CASE(8, 2) MN_CASE(8, 4) MN_CASE(16, 1) MN_CASE(16, 2) MN_CASE(16, 4) MN_CASE(24, 1) MN_CASE(24, 2)
  MN_CASE(32, 1) MN_CASE(32, 2)
#undef MN_CASE
  if (rc != B2F_OK) {
    if (rc == B2F_EUNSUPPORTED) b2f_set_error("b2f_lr_multinomial_lossgrad: no kernel for K=%d d=%d", K, d);
    return rc;
  }
  const int len = K * d + K + 2;
  lr_fold_kernel<<<(len + 127) / 128, 128, 0, st>>>(partials, grid, len, out);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_lr_predict(const float* X, int64_t n, int d, int64_t ld, int n_classes, const double* beta,
                              const double* icpt, double threshold, float* raw, float* prob, float* pred,
                              b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && ld >= d, "b2f_lr_predict: need ld >= d");
  B2F_REQUIRE(n_classes >= 2 && n_classes <= 32, "b2f_lr_predict: n_classes must be in [2, 32]");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && beta && icpt, "b2f_lr_predict: null pointer");
  int64_t want = (n + 7) / 8;
  int64_t cap = (int64_t)b2f_sm_count() * 8;
  int grid = (int)(want < cap ? want : cap);
  lr_predict_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(X, n, d, ld, n_classes, beta, icpt, (float)threshold, raw, prob,
                                                      pred);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the binary loss+gradient pass.
//
// Some callers of this library are not able to hand over scratch memory: a host program that
// is not built on a device allocator (for example a JVM / cgo / ctypes binding that only owns
// the input and output buffers it was given), or code running where cudaMalloc is off limits
// (inside a latency-critical loop, or after the process has handed all device memory to
// another framework's pool).  For them the per-CTA partial sums live in the fixed device array
// below instead of caller-provided scratch.
//
// The array is part of the library image: every GPU that loads libb200fit.so holds these
// LR_NA_SCRATCH_DOUBLES * 8 bytes (about 1.2 MB) until the process exits.
// ---------------------------------------------------------------------------------------------
#define LR_NA_MAX_D 512       // largest feature count (padded row length) the variant accepts
#define LR_NA_MAX_PARTS 304   // largest number of per-CTA partials it will ever use
#define LR_NA_SCRATCH_DOUBLES (LR_NA_MAX_PARTS * (LR_NA_MAX_D + 3))

static __device__ __align__(256) double lr_na_scratch[LR_NA_SCRATCH_DOUBLES];

// Device address of lr_na_scratch, looked up once per device.  In host code the symbol name is
// only a handle for the runtime, not a pointer a kernel can use, so the address has to come
// from cudaGetSymbolAddress with the right device current.
// (The cache is written without a lock.  Two host threads making their first call on one device
// at the same moment both look the address up and store the same value, which is harmless; the
// calls themselves must be serialised anyway because they share the array.)
static int lr_na_scratch_ptr(double** out) {
  static double* cached[64];
  int dev = 0;
  B2F_CUDA(cudaGetDevice(&dev));
  B2F_REQUIRE(dev >= 0 && dev < 64, "b2f_lr_binary_lossgrad_noalloc: device index %d out of range", dev);
  if (!cached[dev]) {
    void* p = nullptr;
    B2F_CUDA(cudaGetSymbolAddress(&p, lr_na_scratch));
    cached[dev] = static_cast<double*>(p);
  }
  *out = cached[dev];
  return B2F_OK;
}

extern "C" int b2f_lr_binary_lossgrad_noalloc(const float* X, const float* y, const float* w, int64_t n, int d,
                                              int64_t ld, const double* beta, const double* icpt, double* out,
                                              b2f_stream_t stream) {
  B2F_REQUIRE(X && y && beta && icpt && out, "b2f_lr_binary_lossgrad_noalloc: null pointer");
  B2F_REQUIRE(n >= 0 && d > 0 && ld >= d && ld % 4 == 0, "b2f_lr_binary_lossgrad_noalloc: need ld >= d, ld %% 4 == 0");
  B2F_REQUIRE(ld <= LR_NA_MAX_D,
              "b2f_lr_binary_lossgrad_noalloc: padded feature count %lld exceeds the built-in scratch limit of %d; "
              "use b2f_lr_binary_lossgrad with caller-provided scratch",
              (long long)ld, LR_NA_MAX_D);
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_lr_binary_lossgrad_noalloc: X must be 16-byte aligned");
  if (n == 0) return B2F_OK;
  double* partials = nullptr;
  int rc = lr_na_scratch_ptr(&partials);  // first call on a device does the one-time address lookup
  if (rc != B2F_OK) return rc;
  cudaStream_t st = b2f_cs(stream);
  int grid = lr_grid_binary();
  if (grid > LR_NA_MAX_PARTS) grid = LR_NA_MAX_PARTS;
  // Nothing in the scratch has to start at zero: each CTA overwrites its whole d + 3 slice and the
  // fold reads exactly the `grid` slices written by this call, so leftovers of earlier calls are inert.
  const int nch = (int)((ld + 127) / 128);
  if (nch <= 1) rc = lr_binary_launch<1>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  else if (nch <= 2) rc = lr_binary_launch<2>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  else rc = lr_binary_launch<4>(X, y, w, n, d, ld, beta, icpt, partials, grid, st);
  if (rc != B2F_OK) return rc;
  const int len = d + 3;
  lr_fold_kernel<<<(len + 127) / 128, 128, 0, st>>>(partials, grid, len, out);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
