// K1 loader: ordered feature columns -> row-major fp32 tiles, plus column statistics.
// Replaces VectorAssembler.transform (reference: orangecontrib/spark/widgets/ml/spark_ml_dataset.py:575-576)
// and the Summarizer pass behind StandardScaler / LogisticRegression (SURVEY.md A.1, A.3 pass 0, A.6).
#include "common.cuh"

#define PK_ROWS 64
#define PK_COLS 32

__device__ __forceinline__ float load_as_f32(const void* p, int dtype, int64_t i) {
  switch (dtype) {
    case B2F_F32: return __ldg(reinterpret_cast<const float*>(p) + i);
    case B2F_F64: return __double2float_rn(__ldg(reinterpret_cast<const double*>(p) + i));
    case B2F_I32: return (float)(double)__ldg(reinterpret_cast<const int32_t*>(p) + i);
    case B2F_I64: return (float)(double)__ldg(reinterpret_cast<const long long*>(p) + i);
    default: return (float)__ldg(reinterpret_cast<const unsigned char*>(p) + i);
  }
}

// Fast path of the loader for fp32 columns (what the Dataset Builder hands over after its casts, and what the BASELINE
// configs hold).  A CTA takes 32 rows x up to 128 columns at a time, so that BOTH sides of the transpose move whole
// 128-byte lines: a lane reads four consecutive rows of one column with one 16-byte load (eight lanes = the 32 rows
// of a column = one line), and the rows leave as contiguous 512-byte pieces.  (The first version, 64 x 32 tiles with
// scalar loads, ran at 0.45 of the HBM peak; 128 x 32 tiles with 16-byte loads but 128-byte row pieces at 0.58.)
// Shared-memory tile [32][128] floats, swizzled at 16-byte granularity with the row group as key: the four scalars a
// lane scatters and the 16-byte reads of the output phase are both conflict-free without padding.
// 64 x 64 tiles (16 KB) whatever the width; a 128-column matrix is two column tiles.  Measured, device columns ->
// row-major, read + write: 32 x 128 tiles 4.2 TB/s at d = 128 and, half of the threads without a column, 2.7 at d = 64;
// 64 x 64 tiles 4.9 / 4.8 (256 contiguous bytes per column and tile instead of 128); 64 x 128 tiles (32 KB, fewer
// resident CTAs) 3.9.  The next tile's loads are issued before the current tile's stores (+4 %).
#define PK4_ROWS 64
#define PK4_TILE (PK4_ROWS * 64)
template <int COLS> __device__ __forceinline__ int pk4_slot(int row, int col) {  // float index of (row, col), swizzled
  return row * COLS + ((((col >> 2) ^ ((row >> 2) & 7)) << 2) | (col & 3));
}
template <int COLS>
__device__ __forceinline__ void pack_columns_f32_body(float* tile, const void* const* __restrict__ cols, int d, int64_t n,
                                                      float* __restrict__ X, int64_t ld, int32_t* bad_flag) {
  constexpr int ROWS = PK4_ROWS;
  constexpr int RG = ROWS / 4;           // groups of four rows: 16 lanes along the rows of one column
  constexpr int CPW = 32 / RG;           // columns one warp instruction covers: 2
  constexpr int LPR = COLS / 4;          // lanes per output row (one float4 each): 32 | 16
  constexpr int NLD = COLS / (8 * CPW);  // 16-byte loads per thread and tile: 8 | 4
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rg = lane % RG, cq = lane / RG;
  const int64_t row_tiles = (n + ROWS - 1) / ROWS;
  const int col_tiles = (int)((ld + COLS - 1) / COLS);
  const int64_t tiles = row_tiles * col_tiles;
  bool bad = false;
  float4 v[NLD];
  auto load_tile = [&](int64_t t) {  // this thread's 16-byte loads of tile t, all in flight together
    const int64_t r = (t / col_tiles) * ROWS + rg * 4;
    const int c0 = (int)(t % col_tiles) * COLS;
#pragma unroll
    for (int cc = 0; cc < NLD; ++cc) {
      const int c = c0 + (warp * NLD + cc) * CPW + cq;
      v[cc] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < d) {
        const float* p = static_cast<const float*>(cols[c]);
        if (r + 3 < n) {
          v[cc] = ldg_stream(reinterpret_cast<const float4*>(p + r));
        } else {
          if (r < n) v[cc].x = __ldg(p + r);
          if (r + 1 < n) v[cc].y = __ldg(p + r + 1);
          if (r + 2 < n) v[cc].z = __ldg(p + r + 2);
        }
      }
    }
  };
  if ((int64_t)blockIdx.x < tiles) load_tile(blockIdx.x);
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t r0 = (t / col_tiles) * ROWS;
    const int c0 = (int)(t % col_tiles) * COLS;
#pragma unroll
    for (int cc = 0; cc < NLD; ++cc) {
      const int cl = (warp * NLD + cc) * CPW + cq;
      bad |= isnan(v[cc].x) | isnan(v[cc].y) | isnan(v[cc].z) | isnan(v[cc].w);
      tile[pk4_slot<COLS>(rg * 4 + 0, cl)] = v[cc].x;
      tile[pk4_slot<COLS>(rg * 4 + 1, cl)] = v[cc].y;
      tile[pk4_slot<COLS>(rg * 4 + 2, cl)] = v[cc].z;
      tile[pk4_slot<COLS>(rg * 4 + 3, cl)] = v[cc].w;
    }
    __syncthreads();
    // the next tile's loads go out before this tile's stores: reads stay in flight through the store phase
    if (t + gridDim.x < tiles) load_tile(t + gridDim.x);
#pragma unroll
    for (int rr = 0; rr < ROWS * LPR / 256; ++rr) {  // a warp writes 512 contiguous bytes (one or two rows) per instruction
      const int rl = (warp * 32 + lane) / LPR + rr * (256 / LPR);
      const int c4 = (lane % LPR) * 4;
      const int64_t row = r0 + rl;
      if (row < n && c0 + c4 < ld)
        *reinterpret_cast<float4*>(X + row * ld + c0 + c4) =
            *reinterpret_cast<const float4*>(tile + pk4_slot<COLS>(rl, c4));
    }
    __syncthreads();
  }
  if (bad_flag && bad) atomicExch(bad_flag, 1);
}

// CTA = 256 threads, tile = 64 rows x 32 columns.  Reads run along N inside each column (coalesced),
// the tile is transposed through padded shared memory, and rows leave as 16-byte stores.
__global__ void __launch_bounds__(256) pack_columns_kernel(const void* const* __restrict__ cols,
                                                           const int32_t* __restrict__ dtypes, int d, int64_t n,
                                                           float* __restrict__ X, int64_t ld, int32_t* bad_flag) {
  __shared__ __align__(16) float tile[PK4_TILE];  // 32 KB; the generic path uses it as [64][33]
  // all columns fp32 and 16-byte aligned -> the 16-byte-load path (decided per CTA, the same everywhere)
  int ok = 1;
  for (int c = threadIdx.x; c < d; c += 256)
    ok &= (dtypes[c] == B2F_F32) && ((reinterpret_cast<uintptr_t>(cols[c]) & 15) == 0);
  if (__syncthreads_and(ok)) {
    pack_columns_f32_body<64>(tile, cols, d, n, X, ld, bad_flag);
    return;
  }
  float (*gt)[PK_COLS + 1] = reinterpret_cast<float (*)[PK_COLS + 1]>(tile);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row_tiles = (n + PK_ROWS - 1) / PK_ROWS;
  const int col_tiles = (int)((ld + PK_COLS - 1) / PK_COLS);
  const int64_t tiles = row_tiles * col_tiles;
  bool bad = false;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t r0 = (t / col_tiles) * PK_ROWS;
    const int c0 = (int)(t % col_tiles) * PK_COLS;
#pragma unroll
    for (int cc = 0; cc < PK_COLS / 8; ++cc) {
      const int cl = warp + cc * 8, c = c0 + cl;
      const void* p = nullptr;
      int dt = 0;
      if (c < d) {
        p = cols[c];
        dt = dtypes[c];
      }
#pragma unroll
      for (int rr = 0; rr < PK_ROWS / 32; ++rr) {
        const int rl = lane + rr * 32;
        const int64_t r = r0 + rl;
        float v = 0.f;
        if (p && r < n) {
          v = load_as_f32(p, dt, r);
          bad |= isnan(v);  // upstream's assembler rejects NaN (and null) under handleInvalid = "error"; Inf passes
        }
        gt[rl][cl] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < PK_ROWS / 32; ++rr) {
      const int rl = (threadIdx.x >> 3) + rr * 32;
      const int c4 = (threadIdx.x & 7) * 4;
      const int64_t r = r0 + rl;
      if (r < n && c0 + c4 < ld) {
        float4 o = make_float4(gt[rl][c4], gt[rl][c4 + 1], gt[rl][c4 + 2], gt[rl][c4 + 3]);
        *reinterpret_cast<float4*>(X + r * ld + c0 + c4) = o;
      }
    }
    __syncthreads();
  }
  if (bad_flag && bad) atomicExch(bad_flag, 1);
}

// grid.y = chunks of 256 columns.  A CTA's 256 threads are laid out as RL row lanes x G float4 column groups, with
// G = the number of column groups this chunk really has (16 at d = 64) and RL = 256 / G: every thread has a column
// to read whatever d is (with a fixed 64 x 4 layout three quarters of the threads idled at d = 64 and the pass ran
// at 12 % of the HBM peak).  Four rows in flight per thread.
__global__ void __launch_bounds__(256) colstats_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld,
                                                       const float* __restrict__ w, const float* __restrict__ shift,
                                                       double* __restrict__ out) {
  const int col0 = blockIdx.y * 256;
  const int G = min(64, (d - col0 + 3) >> 2);
  const int RL = 256 / G;
  const int tx = threadIdx.x % G, ty = threadIdx.x / G;
  const int c = col0 + tx * 4;
  const bool active = ty < RL;
  float sh[4] = {0.f, 0.f, 0.f, 0.f};
  if (active && shift)
    for (int j = 0; j < 4; ++j)
      if (c + j < d) sh[j] = shift[c + j];
  double s1[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0}, sw = 0, sw2 = 0;
  const int64_t stride = (int64_t)gridDim.x * RL;
  if (active) {
    for (int64_t r0 = (int64_t)blockIdx.x * RL + ty; r0 < n; r0 += 4 * stride) {
      float4 v[4];
      float wi[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t r = r0 + u * stride;
        v[u] = make_float4(sh[0], sh[1], sh[2], sh[3]);  // contributes exactly zero
        wi[u] = 0.f;
        if (r < n) {
          v[u] = ldg_stream(reinterpret_cast<const float4*>(X + r * ld + c));
          wi[u] = w ? w[r] : 1.f;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float xs[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double t = (double)xs[j] - (double)sh[j];
          const double wt = (double)wi[u] * t;
          s1[j] += wt;
          s2[j] += wt * t;
        }
        if (blockIdx.y == 0 && tx == 0) {
          sw += (double)wi[u];
          sw2 += (double)wi[u] * (double)wi[u];
        }
      }
    }
  }
  __shared__ double red[256][8];  // [ty * G + tx]
  __shared__ double redw[256][2];
  if (active) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      red[ty * G + tx][j] = s1[j];
      red[ty * G + tx][4 + j] = s2[j];
    }
    if (tx == 0) {
      redw[ty][0] = sw;
      redw[ty][1] = sw2;
    }
  }
  __syncthreads();
  if (ty == 0) {
    double t1[4] = {0, 0, 0, 0}, t2[4] = {0, 0, 0, 0};
    for (int y = 0; y < RL; ++y)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        t1[j] += red[y * G + tx][j];
        t2[j] += red[y * G + tx][4 + j];
      }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (c + j < d) {
        atomicAdd(out + c + j, t1[j]);
        atomicAdd(out + d + c + j, t2[j]);
      }
  }
  if (blockIdx.y == 0 && threadIdx.x == 0) {
    double a = 0, q = 0;
    for (int y = 0; y < RL; ++y) {
      a += redw[y][0];
      q += redw[y][1];
    }
    atomicAdd(out + 2 * d, a);
    atomicAdd(out + 2 * d + 1, q);
  }
}

__global__ void __launch_bounds__(256) label_hist_kernel(const float* __restrict__ y, const float* __restrict__ w,
                                                         int64_t n, int nbins, double* __restrict__ hist,
                                                         int32_t* bad_flag) {
  extern __shared__ double sh[];
  for (int i = threadIdx.x; i < nbins; i += blockDim.x) sh[i] = 0.0;
  __syncthreads();
  bool bad = false;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = y[i];
    const int c = (int)v;
    if (!(v >= 0.f) || (float)c != v || c >= nbins) {
      bad = true;
      continue;
    }
    atomicAdd(&sh[c], w ? (double)w[i] : 1.0);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    if (sh[i] != 0.0) atomicAdd(hist + i, sh[i]);
  if (bad_flag && bad) atomicExch(bad_flag, 1);
}

extern "C" int b2f_pack_columns(const void* const* col_ptrs_dev, const int32_t* col_dtypes_dev, int d, int64_t n,
                                float* X, int64_t ld, int32_t* bad_flag, b2f_stream_t stream) {
  B2F_REQUIRE(d > 0 && n >= 0 && ld >= d && ld % 4 == 0, "b2f_pack_columns: need d > 0, ld >= d, ld %% 4 == 0");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(col_ptrs_dev && col_dtypes_dev && X, "b2f_pack_columns: null pointer");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_pack_columns: X must be 16-byte aligned");
  int64_t tiles = ((n + PK_ROWS - 1) / PK_ROWS) * ((ld + PK_COLS - 1) / PK_COLS);
  int64_t cap = (int64_t)b2f_sm_count() * 8;
  int grid = (int)(tiles < cap ? tiles : cap);
  pack_columns_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(col_ptrs_dev, col_dtypes_dev, d, n, X, ld, bad_flag);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_colstats(const float* X, int64_t n, int d, int64_t ld, const float* w, const float* shift,
                            double* out, b2f_stream_t stream) {
  B2F_REQUIRE(d > 0 && n >= 0 && ld >= d && ld % 4 == 0, "b2f_colstats: need d > 0, ld >= d, ld %% 4 == 0");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && out, "b2f_colstats: null pointer");
  dim3 grid;
  grid.y = (d + 255) / 256;
  const int g0 = d >= 256 ? 64 : (d + 3) / 4;        // column groups of the first chunk -> row lanes per CTA
  const int rows_per_sweep = (256 / g0) * 4;          // four rows in flight per thread
  int64_t want = (n + rows_per_sweep - 1) / rows_per_sweep;
  int64_t cap = (int64_t)b2f_sm_count() * 8 / grid.y + 1;
  grid.x = (unsigned)(want < cap ? want : cap);
  colstats_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(X, n, d, ld, w, shift, out);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------
// StandardScalerModel.transform (A.6): Y = (X - shift) * scale per column; one HBM pass, 16-byte accesses.
// Y may be X itself.  Padding columns d..ld-1 are written as zeros.
// ---------------------------------------------------------------------------------------------
__global__ void scale_columns_kernel(const float* __restrict__ X, float* Y, int64_t n, int d, int64_t ld,
                                     const float* __restrict__ scale, const float* __restrict__ shift) {
  const int64_t ld4 = ld >> 2;
  const int64_t total = n * ld4;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % ld4) * 4;
    float4 v = reinterpret_cast<const float4*>(X)[i];
    float sc[4], sh[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      sc[u] = (c + u < d) ? (scale ? __ldg(scale + c + u) : 1.f) : 0.f;
      sh[u] = (shift && c + u < d) ? __ldg(shift + c + u) : 0.f;
    }
    v.x = (v.x - sh[0]) * sc[0];
    v.y = (v.y - sh[1]) * sc[1];
    v.z = (v.z - sh[2]) * sc[2];
    v.w = (v.w - sh[3]) * sc[3];
    reinterpret_cast<float4*>(Y)[i] = v;
  }
}

extern "C" int b2f_scale_columns(const float* X, float* Y, int64_t n, int d, int64_t ld, const float* scale,
                                 const float* shift, b2f_stream_t stream) {
  B2F_REQUIRE(d > 0 && n >= 0 && ld >= d && ld % 4 == 0, "b2f_scale_columns: need d > 0, ld >= d, ld %% 4 == 0");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && Y, "b2f_scale_columns: null pointer");
  B2F_REQUIRE(((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(Y)) & 15) == 0,
              "b2f_scale_columns: X and Y must be 16-byte aligned");
  if (n == 0) return B2F_OK;
  const int64_t want = (n * (ld >> 2) + 255) / 256;
  const int64_t cap = (int64_t)b2f_sm_count() * 16;
  scale_columns_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, b2f_cs(stream)>>>(X, Y, n, d, ld, scale, shift);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_label_hist(const float* y, const float* w, int64_t n, int n_bins, double* hist, int32_t* bad_flag,
                              b2f_stream_t stream) {
  B2F_REQUIRE(n_bins > 0 && n_bins <= 4096 && n >= 0, "b2f_label_hist: n_bins must be in [1, 4096]");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(y && hist, "b2f_label_hist: null pointer");
  int64_t want = (n + 255) / 256;
  int64_t cap = (int64_t)b2f_sm_count() * 4;
  int grid = (int)(want < cap ? want : cap);
  label_hist_kernel<<<grid, 256, n_bins * sizeof(double), b2f_cs(stream)>>>(y, w, n, n_bins, hist, bad_flag);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
