// Evaluators: squared-Euclidean silhouette as Spark's ClusteringEvaluator
// computes it -- from per-cluster statistics, no pairwise distances:
//   N_c = sum w,  S_c = sum w x,  Psi_c = sum w ||x||^2            (pass 1, sil_stats_kernel)
//   mean squared distance of x to cluster c:  m_c(x) = ||x||^2 + Psi_c / N_c - 2 x . S_c / N_c
//   a = m_own * N_own / (N_own - 1)   (own point excluded),   b = min over the other non-empty clusters of m_c
//   s = 0 if N_own == 1;  1 - a/b if a < b;  b/a - 1 if a > b;  0 otherwise       (pass 2, sil_score_kernel)
// silhouette = sum w s / sum w.  Replaces pyspark.ml.evaluation.ClusteringEvaluator.evaluate;
//   reference call site: orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:42-55 (module = evaluation, :7,19).
#include "common.cuh"

namespace {

// one warp per row: fp64 atomics into stats[c][0..d-1] = S_c, [d] = Psi_c, [d+1] = N_c
__global__ void sil_stats_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld,
                                 const int32_t* __restrict__ labels, const float* __restrict__ w, int k,
                                 double* __restrict__ stats, int32_t* __restrict__ bad_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    const int c = labels[r];
    if (c < 0 || c >= k) {
      if (lane == 0 && bad_flag) *bad_flag = 1;
      continue;
    }
    const double wt = w ? (double)w[r] : 1.0;
    double* row = stats + (size_t)c * (d + 2);
    double psi = 0.0;
    for (int t = lane; t < d; t += 32) {
      const double v = (double)X[r * ld + t];
      psi += v * v;
      atomicAdd(row + t, wt * v);
    }
    psi = warp_sum(psi);
    if (lane == 0) {
      atomicAdd(row + d, wt * psi);
      atomicAdd(row + d + 1, wt);
    }
  }
}

#define SIL_THREADS 256
#define SIL_TILE_FLOATS 8192  // cluster means of a tile in shared memory (32 KB)

// one thread per row; cluster means mu_c = S_c / N_c (fp32) pass through shared memory in tiles
__global__ void __launch_bounds__(SIL_THREADS)
    sil_score_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld, const int32_t* __restrict__ labels,
                     const float* __restrict__ w, int k, const float* __restrict__ mu, const float* __restrict__ q,
                     const double* __restrict__ stats, double* __restrict__ out, int tile_k) {
  extern __shared__ __align__(16) float sil_smem[];
  const int ds = (d + 3) & ~3;
  double acc = 0.0, wsum = 0.0;
  const int64_t nblocks = (n + SIL_THREADS - 1) / SIL_THREADS;
  for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
    const int64_t r = blk * SIL_THREADS + threadIdx.x;
    const bool live = r < n;
    const float* xr = X + (live ? r : 0) * ld;
    const int own = live ? labels[r] : -1;
    float xn = 0.f;
    if (live)
      for (int t = 0; t < d; ++t) xn = fmaf(xr[t], xr[t], xn);
    float a = 0.f, b = INFINITY;
    for (int j0 = 0; j0 < k; j0 += tile_k) {
      const int tk = min(tile_k, k - j0);
      __syncthreads();
      for (int i = threadIdx.x; i < tk * ds; i += SIL_THREADS) {
        const int j = i / ds, t = i - j * ds;
        sil_smem[i] = (t < d) ? mu[(size_t)(j0 + j) * d + t] : 0.f;
      }
      __syncthreads();
      if (!live) continue;
      for (int j = 0; j < tk; ++j) {
        const int c = j0 + j;
        const double nc = stats[(size_t)c * (d + 2) + d + 1];
        if (!(nc > 0.0)) continue;  // empty cluster
        const float* m = sil_smem + j * ds;
        float dot = 0.f;
        for (int t = 0; t < d; ++t) dot = fmaf(xr[t], m[t], dot);
        const float val = fmaxf(xn + q[c] - 2.f * dot, 0.f);
        if (c == own) a = (nc > 1.0) ? val * (float)(nc / (nc - 1.0)) : -1.f;  // -1: singleton cluster
        else b = fminf(b, val);
      }
    }
    if (live && own >= 0 && own < k) {
      float s = 0.f;
      if (a >= 0.f && b < INFINITY) s = (a < b) ? 1.f - a / b : (a > b) ? b / a - 1.f : 0.f;
      const double wt = w ? (double)w[r] : 1.0;
      acc += wt * (double)s;
      wsum += wt;
    }
  }
  acc = warp_sum(acc);
  wsum = warp_sum(wsum);
  if ((threadIdx.x & 31) == 0 && wsum > 0.0) {
    atomicAdd(out, acc);
    atomicAdd(out + 1, wsum);
  }
}

// mu = S / N (fp32), q = Psi / N (fp32) for the non-empty clusters
__global__ void sil_means_kernel(const double* __restrict__ stats, int k, int d, float* __restrict__ mu,
                                 float* __restrict__ q) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k * (d + 1); i += gridDim.x * blockDim.x) {
    const int c = i / (d + 1), t = i - c * (d + 1);
    const double nc = stats[(size_t)c * (d + 2) + d + 1];
    const double v = nc > 0.0 ? stats[(size_t)c * (d + 2) + t] / nc : 0.0;
    if (t < d) mu[(size_t)c * d + t] = (float)v;
    else q[c] = (float)v;
  }
}

}  // namespace

extern "C" int b2f_silhouette_stats(const float* X, int64_t n, int d, int64_t ld, const int32_t* labels,
                                    const float* w, int k, double* stats, int32_t* bad_flag, b2f_stream_t stream) {
  B2F_REQUIRE(d > 0 && k > 0 && n >= 0 && ld >= d, "b2f_silhouette_stats: bad shape");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && labels && stats, "b2f_silhouette_stats: null pointer");
  int64_t want = (n + 7) / 8;
  int64_t cap = (int64_t)b2f_sm_count() * 16;
  sil_stats_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, b2f_cs(stream)>>>(X, n, d, ld, labels, w, k, stats,
                                                                                   bad_flag);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" size_t b2f_silhouette_scratch_bytes(int k, int d) {
  return (k > 0 && d > 0) ? ((size_t)k * d + k) * sizeof(float) : 0;
}

extern "C" int b2f_silhouette_score(const float* X, int64_t n, int d, int64_t ld, const int32_t* labels,
                                    const float* w, int k, const double* stats, void* scratch, double* out,
                                    b2f_stream_t stream) {
  B2F_REQUIRE(d > 0 && k > 0 && n >= 0 && ld >= d, "b2f_silhouette_score: bad shape");
  const int ds = (d + 3) & ~3;
  B2F_REQUIRE(ds <= SIL_TILE_FLOATS, "b2f_silhouette_score: d = %d above %d", d, SIL_TILE_FLOATS);
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && labels && stats && scratch && out, "b2f_silhouette_score: null pointer");
  float* mu = static_cast<float*>(scratch);
  float* q = mu + (size_t)k * d;
  sil_means_kernel<<<(k * (d + 1) + 255) / 256, 256, 0, b2f_cs(stream)>>>(stats, k, d, mu, q);
  B2F_LAUNCH_CHECK();
  int tile_k = SIL_TILE_FLOATS / ds;
  if (tile_k > k) tile_k = k;
  int64_t want = (n + SIL_THREADS - 1) / SIL_THREADS;
  int64_t cap = (int64_t)b2f_sm_count() * 4;
  sil_score_kernel<<<(unsigned)(want < cap ? want : cap), SIL_THREADS, (size_t)tile_k * ds * sizeof(float),
                     b2f_cs(stream)>>>(X, n, d, ld, labels, w, k, mu, q, stats, out, tile_k);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
