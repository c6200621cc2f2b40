// K2 KMeans: centre preparation, the CUDA-core (any k / d) assignment + accumulate kernel, fold,
// finalize and predict, and the dispatch into the tcgen05 kernel (kmeans_tc.cu).
// Replaces the JVM work behind KMeans.fit / KMeansModel.transform
//   reference: orangecontrib/spark/base/spark_ml_estimator.py:22 (module = clustering,
//   orangecontrib/spark/widgets/ml/spark_ml_clustering.py:14) and
//   orangecontrib/spark/widgets/ml/spark_ml_model.py:53;  MLlib semantics: SURVEY.md A.2.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "common.cuh"

// ---------------------------------------------------------------------------------------------
// prepare: fp64 master centres -> kernel operands
// ---------------------------------------------------------------------------------------------
__global__ void km_prepare_kernel(const double* __restrict__ C, int k, int d, float* __restrict__ c32,
                                  float* __restrict__ ch, float* __restrict__ cl, float* __restrict__ cn,
                                  double* __restrict__ c64, __nv_bfloat16* __restrict__ cb1,
                                  __nv_bfloat16* __restrict__ cb2, double* __restrict__ cn64,
                                  unsigned int* __restrict__ misc) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int j = warp; j < k; j += nwarps) {
    double nn = 0.0;
    float amax = 0.f;
    for (int c = lane; c < d; c += 32) {
      const size_t i = (size_t)j * d + c;
      const double v = C[i];
      nn += v * v;
      amax = fmaxf(amax, __double2float_ru(fabs(v)));
      c32[i] = (float)v;
      c64[i] = v;
      const double m2 = -2.0 * v;
      const float hi = to_tf32((float)m2);
      ch[i] = hi;
      cl[i] = to_tf32((float)(m2 - (double)hi));
      const __nv_bfloat16 b1 = __float2bfloat16_rn((float)m2);
      cb1[i] = b1;
      cb2[i] = __float2bfloat16_rn((float)(m2 - (double)__bfloat162float(b1)));
    }
    nn = warp_sum(nn);
    if (lane == 0) {
      cn[j] = (float)nn;
      cn64[j] = nn;
      // max_j ||c_j|| rounded up: non-negative floats order like their bit patterns
      atomicMax(misc + KM_MISC_CMAX, __float_as_uint(__fsqrt_ru((float)nn * 1.0000002f)));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0 && amax > 0.f && amax < INFINITY) atomicMax(misc + KM_MISC_CABS, __float_as_uint(amax));
  }
}

// Second prepare kernel: fp16 operand plane of the single-MMA filter kernel.  S = 2^e is chosen so that
// max |-2 c_jl| * S lies in [2^11, 2^12]: centre entries keep full fp16 precision down to 2^-26 of the largest one
// and data entries up to ~16x the largest |-2c| entry stay finite (rows beyond that are sent to the re-check).
__global__ void km_prepare_h16_kernel(const double* __restrict__ C, int k, int d, __half* __restrict__ ch16,
                                      const double* __restrict__ cn64, __half* __restrict__ cnh,
                                      float* __restrict__ misc, float xmax) {
  const float cabs = misc[KM_MISC_CABS];
  const float cmax = misc[KM_MISC_CMAX];
  // The key window is sized for rows up to 8 max||c||: a few outliers far beyond every centre must not coarsen the
  // quantisation for everybody else (one key unit is R / 2^21); longer rows are flagged by the split warps and go to
  // the re-check like rows whose entries overflow fp16.
  xmax = fminf(xmax, 8.f * cmax);
  int ea = 0, eb = 0;  // S_a = 2^ea (x operand), S_b = 2^eb (centre operand)
  bool global = false;
  if (cabs > 0.f) {
    int ex;
    frexpf(2.f * cabs, &ex);  // 2 cabs = m * 2^ex, m in [0.5, 1)
    ea = eb = max(-60, min(60, 12 - ex));  // per-row mode: both operands on the centres' scale
    if (xmax > 0.f && xmax < INFINITY && cmax < INFINITY) {
      // global mode: sc = 2^es with R * sc < 2^21, R >= |score| of every row; split es between the two operands so
      // that xmax * S_a and 2 cabs * S_b come out alike (their product is below 2^21, so neither can overflow)
      const float R = fmaf(2.f * xmax, cmax, cmax * cmax) * 1.0001f;
      int er, ex_x;
      frexpf(R, &er);          // R < 2^er
      frexpf(xmax, &ex_x);     // xmax < 2^ex_x
      const int es = 21 - er;  // sc = 2^es
      const int a = (es + ex - ex_x) >> 1;  // floor((es + log 2cabs - log xmax) / 2)
      if (a > -60 && a < 60 && es - a > -60 && es - a < 60) {
        ea = a;
        eb = es - a;
        global = true;
      }
    }
  }
  const float Sb = exp2f((float)eb);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    misc[KM_MISC_SCALE] = exp2f((float)ea);
    misc[KM_MISC_SCALE_B] = Sb;
    misc[KM_MISC_INV_SCALE2] = exp2f((float)(-(ea + eb)));
    misc[KM_MISC_XMAX] = global ? xmax : 0.f;
    misc[KM_MISC_SC] = global ? exp2f((float)(ea + eb)) : 0.f;
  }
  const float S = Sb;
  const size_t kd = (size_t)k * d;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < kd; i += (size_t)gridDim.x * blockDim.x)
    ch16[i] = __float2half_rn((float)(-2.0 * C[i]) * S);
  // ||c||^2 * S_a S_b / 2^15 = h1 + h2 + h3 (fp16 pieces; per-row mode: |c_jl| S <= 2^11 and d <= 128 keep h below
  // 2^14; global mode: ||c||^2 sc < 2^21, h < 2^6); column 3: the centre-side half of the magic constant
  const double s2 = exp2((double)(ea + eb)) / (double)KM_CNH_E1;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)k * KM_CNH_COLS;
       i += (size_t)gridDim.x * blockDim.x) {
    const int j = (int)(i / KM_CNH_COLS), c = (int)(i % KM_CNH_COLS);
    __half out = __float2half_rn(0.f);
    if (c < 3) {
      const double h = cn64[j] * s2;
      const __half h1 = __float2half_rn((float)h);
      const double r1 = h - (double)__half2float(h1);
      const __half h2 = __float2half_rn((float)r1);
      const double r2 = r1 - (double)__half2float(h2);
      out = c == 0 ? h1 : c == 1 ? h2 : __float2half_rn((float)r2);
    } else if (c == 3 && global) {
      out = __float2half_rn(KM_CNH_MB);
    }
    cnh[i] = out;
  }
}

// ---------------------------------------------------------------------------------------------
// generic step: one thread per row, centres streamed through shared memory in tiles.
// Distances use the direct form sum (x - c)^2 in fp32 (no cancellation), ascending centre order
// with strict '<' so ties keep the lowest index (findClosest semantics).
// ---------------------------------------------------------------------------------------------
#define KG_THREADS 256
#define KG_TILE_FLOATS 8192   // centre tile in shared memory (32 KB)
#define KG_TABLE_FLOATS 8192  // per-CTA accumulation table(s) (32 KB)

template <bool ACC>
__global__ void __launch_bounds__(KG_THREADS) km_generic_kernel(const float* __restrict__ X, int64_t n, int d,
                                                                int64_t ld, const float* __restrict__ w,
                                                                const float* __restrict__ C32, int k,
                                                                double* __restrict__ payload, float* __restrict__ rep,
                                                                int32_t* __restrict__ labels,
                                                                float* __restrict__ dist2, int tile_k,
                                                                int table_reps) {
  extern __shared__ __align__(16) float smem[];
  const int ds = (d + 3) & ~3;          // padded centre row in smem
  float* ctile = smem;                  // [tile_k][ds]
  float* table = smem + KG_TILE_FLOATS; // [table_reps][k][d+1] when table_reps > 0
  const int tstride = k * (d + 1);
  if (ACC && table_reps > 0) {
    for (int i = threadIdx.x; i < table_reps * tstride; i += KG_THREADS) table[i] = 0.f;
  }
  double cost = 0.0, wsum = 0.0;
  const int64_t nblocks = (n + KG_THREADS - 1) / KG_THREADS;
  for (int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
    const int64_t r = blk * KG_THREADS + threadIdx.x;
    const bool live = r < n;
    const float* xr = X + (live ? r : 0) * ld;
    float best = INFINITY;
    int bi = 0;
    for (int j0 = 0; j0 < k; j0 += tile_k) {
      const int tk = min(tile_k, k - j0);
      __syncthreads();
      for (int i = threadIdx.x; i < tk * ds; i += KG_THREADS) {
        const int jj = i / ds, c = i - jj * ds;
        ctile[i] = (c < d) ? C32[(size_t)(j0 + jj) * d + c] : 0.f;
      }
      __syncthreads();
      if (live) {
        for (int jj = 0; jj < tk; jj += 4) {
          float4 a0 = make_float4(0, 0, 0, 0), a1 = a0, a2 = a0, a3 = a0;
          const float* c0 = ctile + (size_t)jj * ds;
          const int j1 = min(jj + 1, tk - 1), j2 = min(jj + 2, tk - 1), j3 = min(jj + 3, tk - 1);
          const float* c1 = ctile + (size_t)j1 * ds;
          const float* c2 = ctile + (size_t)j2 * ds;
          const float* c3 = ctile + (size_t)j3 * ds;
          for (int c = 0; c < ds; c += 4) {
            float4 x = *reinterpret_cast<const float4*>(xr + c);
            if (c + 3 >= d) {  // mask padding columns (caller's padding need not be zero)
              if (c + 1 >= d) x.y = 0.f;
              if (c + 2 >= d) x.z = 0.f;
              x.w = 0.f;
            }
            float4 v, t;
            v = *reinterpret_cast<const float4*>(c0 + c);
            t.x = x.x - v.x; t.y = x.y - v.y; t.z = x.z - v.z; t.w = x.w - v.w;
            a0.x = fmaf(t.x, t.x, a0.x); a0.y = fmaf(t.y, t.y, a0.y); a0.z = fmaf(t.z, t.z, a0.z); a0.w = fmaf(t.w, t.w, a0.w);
            v = *reinterpret_cast<const float4*>(c1 + c);
            t.x = x.x - v.x; t.y = x.y - v.y; t.z = x.z - v.z; t.w = x.w - v.w;
            a1.x = fmaf(t.x, t.x, a1.x); a1.y = fmaf(t.y, t.y, a1.y); a1.z = fmaf(t.z, t.z, a1.z); a1.w = fmaf(t.w, t.w, a1.w);
            v = *reinterpret_cast<const float4*>(c2 + c);
            t.x = x.x - v.x; t.y = x.y - v.y; t.z = x.z - v.z; t.w = x.w - v.w;
            a2.x = fmaf(t.x, t.x, a2.x); a2.y = fmaf(t.y, t.y, a2.y); a2.z = fmaf(t.z, t.z, a2.z); a2.w = fmaf(t.w, t.w, a2.w);
            v = *reinterpret_cast<const float4*>(c3 + c);
            t.x = x.x - v.x; t.y = x.y - v.y; t.z = x.z - v.z; t.w = x.w - v.w;
            a3.x = fmaf(t.x, t.x, a3.x); a3.y = fmaf(t.y, t.y, a3.y); a3.z = fmaf(t.z, t.z, a3.z); a3.w = fmaf(t.w, t.w, a3.w);
          }
          const float d0 = (a0.x + a0.y) + (a0.z + a0.w), d1 = (a1.x + a1.y) + (a1.z + a1.w);
          const float d2 = (a2.x + a2.y) + (a2.z + a2.w), d3 = (a3.x + a3.y) + (a3.z + a3.w);
          if (d0 < best) { best = d0; bi = j0 + jj; }
          if (jj + 1 < tk && d1 < best) { best = d1; bi = j0 + jj + 1; }
          if (jj + 2 < tk && d2 < best) { best = d2; bi = j0 + jj + 2; }
          if (jj + 3 < tk && d3 < best) { best = d3; bi = j0 + jj + 3; }
        }
      }
    }
    if (live) {
      if (labels) labels[r] = bi;
      if (dist2) dist2[r] = best;
      if (ACC) {
        const float wi = w ? w[r] : 1.f;
        cost += (double)wi * (double)best;
        wsum += (double)wi;
        if (table_reps > 0) {
          float* t = table + (size_t)((threadIdx.x >> 5) % table_reps) * tstride + (size_t)bi * (d + 1);
          for (int c = 0; c < d; ++c) atomicAdd(t + c, wi * xr[c]);
          atomicAdd(t + d, wi);
        } else {
          float* t = rep + ((size_t)(blockIdx.x % KM_REPLICAS) * k + bi) * d;
          for (int c = 0; c < d; ++c) atomicAdd(t + c, wi * xr[c]);
          atomicAdd(payload + (size_t)k * d + bi, (double)wi);
        }
      }
    }
  }
  if (ACC) {
    __syncthreads();
    if (table_reps > 0) {
      for (int i = threadIdx.x; i < tstride; i += KG_THREADS) {
        double s = 0.0;
        for (int rp = 0; rp < table_reps; ++rp) s += (double)table[(size_t)rp * tstride + i];
        if (s != 0.0) {
          const int j = i / (d + 1), c = i - j * (d + 1);
          if (c < d) atomicAdd(payload + (size_t)j * d + c, s);
          else atomicAdd(payload + (size_t)k * d + j, s);
        }
      }
    }
    cost = warp_sum(cost);
    wsum = warp_sum(wsum);
    __shared__ double red[2][KG_THREADS / 32];
    if ((threadIdx.x & 31) == 0) {
      red[0][threadIdx.x >> 5] = cost;
      red[1][threadIdx.x >> 5] = wsum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double c = 0, s = 0;
      for (int i = 0; i < KG_THREADS / 32; ++i) {
        c += red[0][i];
        s += red[1][i];
      }
      atomicAdd(payload + (size_t)k * d + k, c);
      atomicAdd(payload + (size_t)k * d + k + 1, s);
    }
  }
}

static int km_generic_launch(const float* X, int64_t n, int d, int64_t ld, const float* w, const float* C32, int k,
                             double* payload, float* rep, int32_t* labels, float* dist2, bool acc,
                             cudaStream_t st) {
  const int ds = (d + 3) & ~3;
  B2F_REQUIRE(ds * 4 <= KG_TILE_FLOATS, "b2f_kmeans: generic kernel supports d <= %d", KG_TILE_FLOATS / 4);
  int tile_k = KG_TILE_FLOATS / ds;
  tile_k = (tile_k / 4) * 4;
  if (tile_k > k) tile_k = (k + 3) / 4 * 4;
  int table_reps = 0;
  const size_t tstride = (size_t)k * (d + 1);
  if (acc && tstride <= KG_TABLE_FLOATS) {
    table_reps = (int)(KG_TABLE_FLOATS / tstride);
    if (table_reps > KG_THREADS / 32) table_reps = KG_THREADS / 32;
  }
  size_t smem = (size_t)KG_TILE_FLOATS * 4 + (size_t)table_reps * tstride * 4;
  int64_t nblocks = (n + KG_THREADS - 1) / KG_THREADS;
  int64_t cap = (int64_t)b2f_sm_count() * 3;
  int grid = (int)(nblocks < cap ? nblocks : cap);
  if (grid < 1) grid = 1;
  if (acc) {
    B2F_CUDA(b2f_smem_attr(km_generic_kernel<true>, smem));
    km_generic_kernel<true><<<grid, KG_THREADS, smem, st>>>(X, n, d, ld, w, C32, k, payload, rep, labels, dist2,
                                                            tile_k, table_reps);
  } else {
    B2F_CUDA(b2f_smem_attr(km_generic_kernel<false>, smem));
    km_generic_kernel<false><<<grid, KG_THREADS, smem, st>>>(X, n, d, ld, w, C32, k, payload, rep, labels, dist2,
                                                             tile_k, 0);
  }
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------
// exact re-check: fp64 assignment of the listed rows against every centre, on the fp64 master centres:
//   dist(r, j) = ||x_r||^2 + ||c_j||^2 - 2 x_r . c_j   (all fp64; each dot product is summed in the same order
//   whichever thread computes it, so duplicate centres tie exactly and the lowest index wins: findClosest order).
// A CTA takes KX_ROWS listed rows: warp w owns rows 4w .. 4w+3, lane l owns centre jb + l of the current block of
// 32 centres, whose coordinates pass through a padded shared-memory tile; so the centre table is read from L2 once
// per 32 rows.  Accumulates straight into the fp64 payload.  Only rows that the tensor-core kernels could not
// certify come here (tens per million).
// ---------------------------------------------------------------------------------------------
#define KX_ROWS 32
#define KX_MAX_D 512
__global__ void __launch_bounds__(256) km_exact_kernel(const float* __restrict__ X, int64_t ld, int d,
                                                       const float* __restrict__ colscale,
                                                       const double* __restrict__ C64,
                                                       const double* __restrict__ CN64, int k,
                                                       double* __restrict__ payload, int32_t* __restrict__ labels,
                                                       const int* __restrict__ count, const int32_t* __restrict__ idx,
                                                       int cap, int accumulate) {
  extern __shared__ __align__(16) double kx_smem[];
  const int dpad = (d + 31) & ~31;
  double* xs = kx_smem;                     // [KX_ROWS][dpad]
  double* cs = xs + (size_t)KX_ROWS * dpad; // [32 centres][34]: rows 16-byte aligned, conflict-free column reads
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int used = min(*count, cap);
  double cost = 0.0, rows = 0.0;
  for (int g0 = blockIdx.x * KX_ROWS; g0 < used; g0 += gridDim.x * KX_ROWS) {
    __syncthreads();
    for (int i = threadIdx.x; i < KX_ROWS * dpad; i += 256) {
      const int r = i / dpad, t = i - r * dpad;
      // (scaled pass: the row is fl32(x * scale), the value a materialised StandardScaler copy would hold)
      float xv = 0.f;
      if (g0 + r < used && t < d) {
        xv = X[(int64_t)idx[g0 + r] * ld + t];
        if (colscale) xv = __fmul_rn(xv, colscale[t]);
      }
      xs[i] = (double)xv;
    }
    __syncthreads();
    double xn[4], best[4];
    int bi[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      double a = 0.0;
      for (int t = lane; t < dpad; t += 32) a = fma(xs[(4 * warp + u) * dpad + t], xs[(4 * warp + u) * dpad + t], a);
      xn[u] = warp_sum(a);
      best[u] = INFINITY;
      bi[u] = 0x7fffffff;
    }
    for (int jb = 0; jb < k; jb += 32) {
      double s[4] = {0.0, 0.0, 0.0, 0.0};
      for (int t0 = 0; t0 < dpad; t0 += 32) {
        __syncthreads();
#pragma unroll
        for (int e = threadIdx.x; e < 1024; e += 256) {
          const int c = e >> 5, tt = e & 31;
          cs[c * 34 + tt] = (jb + c < k && t0 + tt < d) ? C64[(size_t)(jb + c) * d + t0 + tt] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int tt = 0; tt < 32; tt += 2) {
          const double2 cv = *reinterpret_cast<const double2*>(cs + lane * 34 + tt);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const double2 xv = *reinterpret_cast<const double2*>(xs + (4 * warp + u) * dpad + t0 + tt);
            s[u] = fma(xv.x, cv.x, s[u]);
            s[u] = fma(xv.y, cv.y, s[u]);
          }
        }
      }
      const int j = jb + lane;
      if (j < k) {
        const double cn = CN64[j];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const double v = fmax(xn[u] + cn - 2.0 * s[u], 0.0);
          if (v < best[u]) {  // a thread sees its centres in ascending order: strict '<' keeps the lowest index
            best[u] = v;
            bi[u] = j;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best[u], o);
        const int oj = __shfl_xor_sync(0xffffffffu, bi[u], o);
        if (ov < best[u] || (ov == best[u] && oj < bi[u])) {
          best[u] = ov;
          bi[u] = oj;
        }
      }
      const int li = g0 + 4 * warp + u;
      if (li < used) {
        const int64_t r = idx[li];
        if (lane == 0 && labels) labels[r] = bi[u];
        if (accumulate) {
          for (int t = lane; t < d; t += 32)
            atomicAdd(payload + (size_t)bi[u] * d + t, xs[(4 * warp + u) * dpad + t]);
          if (lane == 0) {
            atomicAdd(payload + (size_t)k * d + bi[u], 1.0);
            cost += best[u];
            rows += 1.0;
          }
        }
      }
    }
  }
  if (accumulate && lane == 0 && rows > 0.0) {
    atomicAdd(payload + (size_t)k * d + k, cost);
    atomicAdd(payload + (size_t)k * d + k + 1, rows);
  }
}

int km_exact_rows(const float* X, int64_t ld, int d, const float* colscale, const void* prep, int k, void* acc,
                  int32_t* labels, const int* count, const int32_t* idx, int cap, bool accumulate,
                  cudaStream_t stream) {
  B2F_REQUIRE(d <= KX_MAX_D, "exact re-check: d = %d above %d", d, KX_MAX_D);
  KmPrepLayout P = km_prep_layout(k, d);
  const double* c64 = reinterpret_cast<const double*>(static_cast<const char*>(prep) + P.off_c64);
  const double* cn64 = reinterpret_cast<const double*>(static_cast<const char*>(prep) + P.off_cn64);
  double* payload = nullptr;
  if (accumulate) payload = reinterpret_cast<double*>(static_cast<char*>(acc) + km_acc_layout(k, d).off_payload);
  const int dpad = (d + 31) & ~31;
  const size_t smem = ((size_t)KX_ROWS * dpad + 32 * 34) * sizeof(double);
  B2F_CUDA(b2f_smem_attr(km_exact_kernel, smem));
  int grid = (cap + KX_ROWS - 1) / KX_ROWS;
  if (grid > b2f_sm_count() * 4) grid = b2f_sm_count() * 4;
  km_exact_kernel<<<grid, 256, smem, stream>>>(X, ld, d, colscale, c64, cn64, k, payload, labels, count, idx, cap,
                                               accumulate ? 1 : 0);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------
// squared distance of every row to the centre its label names: sum (x - c)^2 in fp32 on the fp32 centres, the
// same direct form the generic kernel reports.  Warp per row, 16-byte loads; the centre table stays in L2.
// Serves computeCost and the k-means|| rounds when the labels come from the tensor-core pipeline.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) km_dist_label_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld,
                                                            const float* __restrict__ colscale,
                                                            const float* __restrict__ C32, int k,
                                                            const int32_t* __restrict__ labels,
                                                            float* __restrict__ dist2) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < n; r += nwarps) {
    const int lbl = labels[r];
    float a = 0.f;
    if (lbl >= 0 && lbl < k) {
      const float* xr = X + r * ld;
      const float* cr = C32 + (size_t)lbl * d;
      for (int c = lane * 4; c < d; c += 128) {
        float4 x = ldg_stream(reinterpret_cast<const float4*>(xr + c));
        if (colscale) {
          x.x = __fmul_rn(x.x, colscale[c]);
          if (c + 1 < d) x.y = __fmul_rn(x.y, colscale[c + 1]);
          if (c + 2 < d) x.z = __fmul_rn(x.z, colscale[c + 2]);
          if (c + 3 < d) x.w = __fmul_rn(x.w, colscale[c + 3]);
        }
        float t = x.x - cr[c];
        a = fmaf(t, t, a);
        if (c + 1 < d) { t = x.y - cr[c + 1]; a = fmaf(t, t, a); }
        if (c + 2 < d) { t = x.z - cr[c + 2]; a = fmaf(t, t, a); }
        if (c + 3 < d) { t = x.w - cr[c + 3]; a = fmaf(t, t, a); }
      }
    }
    a = warp_sum(a);
    if (lane == 0) dist2[r] = a;
  }
}

int km_dist_to_label(const float* X, int64_t n, int d, int64_t ld, const float* colscale, const void* prep, int k,
                     const int32_t* labels, float* dist2, cudaStream_t stream) {
  KmPrepLayout P = km_prep_layout(k, d);
  const float* c32 = reinterpret_cast<const float*>(static_cast<const char*>(prep) + P.off_c32);
  int64_t want = (n + 7) / 8;
  int64_t cap = (int64_t)b2f_sm_count() * 16;
  km_dist_label_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, stream>>>(X, n, d, ld, colscale, c32, k, labels,
                                                                               dist2);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// ---------------------------------------------------------------------------------------------
// fold / finalize
// ---------------------------------------------------------------------------------------------
__global__ void km_fold_kernel(double* __restrict__ payload, float* __restrict__ rep,
                               unsigned long long* __restrict__ cnt, int k, int d,
                               const float* __restrict__ colscale) {
  const size_t kd = (size_t)k * d;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < kd; i += (size_t)gridDim.x * blockDim.x) {
    double s = 0.0;
#pragma unroll
    for (int rp = 0; rp < KM_REPLICAS; ++rp) {
      s += (double)rep[rp * kd + i];
      rep[rp * kd + i] = 0.f;
    }
    if (colscale) s *= (double)colscale[i % d];
    payload[i] += s;
  }
  for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < (size_t)k; j += (size_t)gridDim.x * blockDim.x) {
    payload[kd + j] += (double)cnt[j];
    cnt[j] = 0ull;
  }
}

__global__ void km_finalize_init_kernel(const double* __restrict__ payload, int k, int d, int32_t* status,
                                        double* status_f64) {
  status[0] = 1;
  status[1] = 0;
  status_f64[0] = 0.0;
  status_f64[1] = payload[(size_t)k * d + k];
}

__global__ void km_finalize_kernel(const double* __restrict__ payload, int k, int d, double tol2,
                                   double* __restrict__ C, int32_t* status, double* status_f64) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int j = warp; j < k; j += nwarps) {
    const double cnt = payload[(size_t)k * d + j];
    if (!(cnt > 0.0)) continue;  // empty cluster keeps its centre
    double moved = 0.0;
    for (int c = lane; c < d; c += 32) {
      const size_t i = (size_t)j * d + c;
      const double nc = payload[i] / cnt;
      const double df = nc - C[i];
      moved += df * df;
      C[i] = nc;
    }
    moved = warp_sum(moved);
    if (lane == 0) {
      if (moved > tol2) atomicExch(status, 0);
      atomicMax(reinterpret_cast<unsigned long long*>(status_f64), (unsigned long long)__double_as_longlong(moved));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
extern "C" size_t b2f_kmeans_prep_bytes(int k, int d) { return (k > 0 && d > 0) ? km_prep_layout(k, d).bytes : 0; }
extern "C" size_t b2f_kmeans_acc_bytes(int k, int d) { return (k > 0 && d > 0) ? km_acc_layout(k, d).bytes : 0; }
extern "C" int64_t b2f_kmeans_payload_len(int k, int d) {
  return (k > 0 && d > 0) ? (int64_t)km_acc_layout(k, d).payload_len : 0;
}

extern "C" int b2f_kmeans_prepare(const double* centers, int k, int d, void* prep, b2f_stream_t stream) {
  return b2f_kmeans_prepare_x(centers, k, d, 0.0, prep, stream);
}

// max_i ||x_i|| (of the rows x * col_scale when a factor is given), fp32 rounded up, atomicMax on the float bits
__global__ void __launch_bounds__(256) km_row_norm_max_kernel(const float* __restrict__ X, int64_t n, int d, int64_t ld,
                                                              const float* __restrict__ colscale,
                                                              unsigned int* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float best = 0.f;
  for (int64_t r = warp; r < n; r += nwarps) {
    const float* xr = X + r * ld;
    float a = 0.f;
    for (int c = lane * 4; c < d; c += 128) {
      float4 x = ldg_stream(reinterpret_cast<const float4*>(xr + c));
      if (c + 1 >= d) x.y = 0.f;
      if (c + 2 >= d) x.z = 0.f;
      if (c + 3 >= d) x.w = 0.f;
      if (colscale) {
        x.x = __fmul_rn(x.x, colscale[c]);
        if (c + 1 < d) x.y = __fmul_rn(x.y, colscale[c + 1]);
        if (c + 2 < d) x.z = __fmul_rn(x.z, colscale[c + 2]);
        if (c + 3 < d) x.w = __fmul_rn(x.w, colscale[c + 3]);
      }
      a = fmaf(x.x, x.x, a); a = fmaf(x.y, x.y, a); a = fmaf(x.z, x.z, a); a = fmaf(x.w, x.w, a);
    }
    a = warp_sum(a);
    best = fmaxf(best, a);  // NaN rows drop out here; the filter sends them to the re-check anyway
  }
  if (lane == 0 && best > 0.f && best < INFINITY)
    atomicMax(out, __float_as_uint(__fsqrt_ru(best * 1.00001f)));
}

extern "C" int b2f_row_norm_max(const float* X, int64_t n, int d, int64_t ld, const float* col_scale, float* out,
                                b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && ld >= d && ld % 4 == 0, "b2f_row_norm_max: need ld >= d, ld %% 4 == 0");
  B2F_REQUIRE(out, "b2f_row_norm_max: null output");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X, "b2f_row_norm_max: null pointer");
  int64_t want = (n + 7) / 8;
  int64_t cap = (int64_t)b2f_sm_count() * 16;
  km_row_norm_max_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, b2f_cs(stream)>>>(
      X, n, d, ld, col_scale, reinterpret_cast<unsigned int*>(out));
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_kmeans_prepare_x(const double* centers, int k, int d, double x_norm_max, void* prep,
                                    b2f_stream_t stream) {
  B2F_REQUIRE(centers && prep && k > 0 && d > 0, "b2f_kmeans_prepare: bad arguments");
  B2F_REQUIRE(x_norm_max >= 0.0, "b2f_kmeans_prepare_x: x_norm_max must be >= 0 (0 = unknown)");
  KmPrepLayout L = km_prep_layout(k, d);
  char* p = static_cast<char*>(prep);
  int grid = (k + 7) / 8;
  B2F_CUDA(cudaMemsetAsync(p + L.off_misc, 0, 256, b2f_cs(stream)));
  km_prepare_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(
      centers, k, d, reinterpret_cast<float*>(p + L.off_c32), reinterpret_cast<float*>(p + L.off_ch),
      reinterpret_cast<float*>(p + L.off_cl), reinterpret_cast<float*>(p + L.off_cn),
      reinterpret_cast<double*>(p + L.off_c64), reinterpret_cast<__nv_bfloat16*>(p + L.off_cb1),
      reinterpret_cast<__nv_bfloat16*>(p + L.off_cb2), reinterpret_cast<double*>(p + L.off_cn64),
      reinterpret_cast<unsigned int*>(p + L.off_misc));
  B2F_LAUNCH_CHECK();
  km_prepare_h16_kernel<<<(int)(((size_t)k * d + 255) / 256 > 1184 ? 1184 : ((size_t)k * d + 255) / 256), 256, 0,
                          b2f_cs(stream)>>>(centers, k, d, reinterpret_cast<__half*>(p + L.off_ch16),
                                            reinterpret_cast<const double*>(p + L.off_cn64),
                                            reinterpret_cast<__half*>(p + L.off_cnh),
                                            reinterpret_cast<float*>(p + L.off_misc), (float)x_norm_max);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_kmeans_acc_zero(void* acc, int k, int d, b2f_stream_t stream) {
  B2F_REQUIRE(acc && k > 0 && d > 0, "b2f_kmeans_acc_zero: bad arguments");
  B2F_CUDA(cudaMemsetAsync(acc, 0, km_acc_layout(k, d).bytes, b2f_cs(stream)));
  return B2F_OK;
}

static int km_pick_algo(int algo, int k, int d, int64_t ld, const float* w) {
  if (algo == B2F_KM_AUTO) return (!w && km_tc_supported(k, d, ld)) ? B2F_KM_TC3 : B2F_KM_GENERIC;
  return algo;
}

extern "C" int b2f_kmeans_step(const float* X, int64_t n, int d, int64_t ld, const float* w, const void* prep, int k,
                               void* acc, int32_t* labels, int algo, b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && k > 0 && ld >= d && ld % 4 == 0, "b2f_kmeans_step: need ld >= d, ld %% 4 == 0");
  if (n == 0) return B2F_OK;  // an empty row shard: nothing to add (X may be NULL)
  B2F_REQUIRE(X && prep && acc, "b2f_kmeans_step: null pointer");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_kmeans_step: X must be 16-byte aligned");
  KmPrepLayout P = km_prep_layout(k, d);
  KmAccLayout A = km_acc_layout(k, d);
  const char* pp = static_cast<const char*>(prep);
  char* ap = static_cast<char*>(acc);
  algo = km_pick_algo(algo, k, d, ld, w);
  if (algo == B2F_KM_TC3 || algo == B2F_KM_TC1) {
    B2F_REQUIRE(!w, "b2f_kmeans_step: the tensor-core path does not take row weights");
    B2F_REQUIRE(km_tc_supported(k, d, ld), "b2f_kmeans_step: tensor-core path needs d %% 32 == 0, d <= 256, k %% 128 == 0");
    return km_tc_step(X, n, d, ld, prep, k, acc, labels, algo == B2F_KM_TC3 ? 3 : 1, true, nullptr, nullptr, b2f_cs(stream));
  }
  B2F_REQUIRE(algo == B2F_KM_GENERIC, "b2f_kmeans_step: unknown algo %d", algo);
  return km_generic_launch(X, n, d, ld, w, reinterpret_cast<const float*>(pp + P.off_c32), k,
                           reinterpret_cast<double*>(ap + A.off_payload), reinterpret_cast<float*>(ap + A.off_rep),
                           labels, nullptr, true, b2f_cs(stream));
}

extern "C" size_t b2f_kmeans_ws_bytes(int64_t n, int k, int d) {
  return (n >= 0 && k > 0 && d > 0) ? km_tcb_ws_bytes(n, k, d) : 0;
}

// shared by b2f_kmeans_step_ws / b2f_kmeans_step_scaled
static int km_step_ws_impl(const char* who, const float* X, int64_t n, int d, int64_t ld, const float* colscale,
                           const float* w, const void* prep, int k, void* acc, int32_t* labels, void* ws,
                           size_t ws_bytes, int algo, b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && k > 0 && ld >= d && ld % 4 == 0, "%s: need ld >= d, ld %% 4 == 0", who);
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && prep && acc, "%s: null pointer", who);
  const bool have_ws = labels && ws;
  if (algo == B2F_KM_AUTO && !w && have_ws && km_tcb_supported(k, d, ld)) algo = B2F_KM_TCH;
  // 3xTF32 WITH its certainty test and the fp64 tail (the fallback of the filter pipeline) needs the workspace too
  const bool tc3_checked = algo == B2F_KM_TC3 && !w && have_ws && km_tc_supported(k, d, ld);
  if (algo != B2F_KM_TCB && algo != B2F_KM_TCH && !tc3_checked) {
    B2F_REQUIRE(!colscale, "%s: a column scale needs the filter pipeline (d in {64, 128}, k %% 128 == 0, labels, ws)", who);
    return b2f_kmeans_step(X, n, d, ld, w, prep, k, acc, labels, algo, stream);
  }
  B2F_REQUIRE(!w, "%s: the tensor-core path does not take row weights", who);
  B2F_REQUIRE(have_ws, "%s: the filter pipeline needs labels and a workspace", who);
  B2F_REQUIRE(tc3_checked || km_tcb_supported(k, d, ld), "%s: B2F_KM_TCB / B2F_KM_TCH need d in {64, 128}, k %% 128 == 0 and k <= 65536",
              who);
  B2F_REQUIRE(!colscale || algo == B2F_KM_TCH, "%s: a column scale is only taken by B2F_KM_TCH / B2F_KM_AUTO", who);
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "%s: X must be 16-byte aligned", who);
  const int filter = tc3_checked ? KM_FILTER_NONE : algo == B2F_KM_TCH ? KM_FILTER_FP16 : KM_FILTER_BF16X3;
  return km_tcb_step(X, n, d, ld, colscale, prep, k, acc, labels, ws, ws_bytes, true, filter, b2f_cs(stream));
}

extern "C" int b2f_kmeans_step_ws(const float* X, int64_t n, int d, int64_t ld, const float* w, const void* prep,
                                  int k, void* acc, int32_t* labels, void* ws, size_t ws_bytes, int algo,
                                  b2f_stream_t stream) {
  return km_step_ws_impl("b2f_kmeans_step_ws", X, n, d, ld, nullptr, w, prep, k, acc, labels, ws, ws_bytes, algo, stream);
}

extern "C" int b2f_kmeans_step_scaled(const float* X, int64_t n, int d, int64_t ld, const float* col_scale,
                                      const void* prep, int k, void* acc, int32_t* labels, void* ws, size_t ws_bytes,
                                      b2f_stream_t stream) {
  B2F_REQUIRE(col_scale, "b2f_kmeans_step_scaled: null column scale");
  return km_step_ws_impl("b2f_kmeans_step_scaled", X, n, d, ld, col_scale, nullptr, prep, k, acc, labels, ws, ws_bytes,
                         B2F_KM_TCH, stream);
}

extern "C" int b2f_kmeans_scaled_supported(int k, int d, int64_t ld) { return km_tcb_supported(k, d, ld) ? 1 : 0; }

extern "C" int b2f_kmeans_fold(void* acc, int k, int d, b2f_stream_t stream) {
  B2F_REQUIRE(acc && k > 0 && d > 0, "b2f_kmeans_fold: bad arguments");
  KmAccLayout A = km_acc_layout(k, d);
  char* ap = static_cast<char*>(acc);
  size_t kd = (size_t)k * d;
  int grid = (int)((kd + 255) / 256);
  if (grid > b2f_sm_count() * 8) grid = b2f_sm_count() * 8;
  km_fold_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(reinterpret_cast<double*>(ap + A.off_payload),
                                                   reinterpret_cast<float*>(ap + A.off_rep),
                                                   reinterpret_cast<unsigned long long*>(ap + A.off_cnt), k, d,
                                                   nullptr);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

// used by the scaled filter pass (kmeans_tcb.cu): replicas hold RAW row sums, the payload scaled ones
int km_fold_scaled(void* acc, int k, int d, const float* colscale, cudaStream_t stream) {
  KmAccLayout A = km_acc_layout(k, d);
  char* ap = static_cast<char*>(acc);
  size_t kd = (size_t)k * d;
  int grid = (int)((kd + 255) / 256);
  if (grid > b2f_sm_count() * 8) grid = b2f_sm_count() * 8;
  km_fold_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<double*>(ap + A.off_payload),
                                           reinterpret_cast<float*>(ap + A.off_rep),
                                           reinterpret_cast<unsigned long long*>(ap + A.off_cnt), k, d, colscale);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_kmeans_finalize(const double* payload, int k, int d, double tol, double* centers, int32_t* status,
                                   double* status_f64, b2f_stream_t stream) {
  B2F_REQUIRE(payload && centers && status && status_f64 && k > 0 && d > 0, "b2f_kmeans_finalize: bad arguments");
  km_finalize_init_kernel<<<1, 1, 0, b2f_cs(stream)>>>(payload, k, d, status, status_f64);
  int grid = (k + 7) / 8;
  km_finalize_kernel<<<grid, 256, 0, b2f_cs(stream)>>>(payload, k, d, tol * tol, centers, status, status_f64);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

extern "C" int b2f_kmeans_predict(const float* X, int64_t n, int d, int64_t ld, const void* prep, int k,
                                  int32_t* labels, float* dist2, int algo, b2f_stream_t stream) {
  return b2f_kmeans_predict_ws(X, n, d, ld, nullptr, prep, k, labels, dist2, nullptr, 0, algo, stream);
}

extern "C" int b2f_kmeans_predict_ws(const float* X, int64_t n, int d, int64_t ld, const float* col_scale,
                                     const void* prep, int k, int32_t* labels, float* dist2, void* ws, size_t ws_bytes,
                                     int algo, b2f_stream_t stream) {
  B2F_REQUIRE(n >= 0 && d > 0 && k > 0 && ld >= d && ld % 4 == 0, "b2f_kmeans_predict: need ld >= d, ld %% 4 == 0");
  if (n == 0) return B2F_OK;
  B2F_REQUIRE(X && prep && (labels || dist2), "b2f_kmeans_predict: null pointer");
  KmPrepLayout P = km_prep_layout(k, d);
  const char* pp = static_cast<const char*>(prep);
  const bool have_ws = labels && ws;
  // the certainty pipeline (labels equal to fp64 except fp64-level ties) whenever the shape allows it
  int filter = -1;
  if ((algo == B2F_KM_AUTO || algo == B2F_KM_TCH) && have_ws && km_tcb_supported(k, d, ld)) filter = KM_FILTER_FP16;
  else if (algo == B2F_KM_TCB && have_ws && km_tcb_supported(k, d, ld)) filter = KM_FILTER_BF16X3;
  else if ((algo == B2F_KM_AUTO || algo == B2F_KM_TC3) && have_ws && km_tc_supported(k, d, ld)) filter = KM_FILTER_NONE;
  if (filter >= 0 && (!col_scale || filter == KM_FILTER_FP16)) {
    B2F_REQUIRE((reinterpret_cast<uintptr_t>(X) & 15) == 0, "b2f_kmeans_predict: X must be 16-byte aligned");
    int rc = km_tcb_step(X, n, d, ld, col_scale, prep, k, nullptr, labels, ws, ws_bytes, false, filter, b2f_cs(stream));
    if (rc != B2F_OK) return rc;
    if (dist2) return km_dist_to_label(X, n, d, ld, col_scale, prep, k, labels, dist2, b2f_cs(stream));
    return B2F_OK;
  }
  B2F_REQUIRE(!col_scale, "b2f_kmeans_predict_ws: a column scale needs the filter pipeline (d in {64, 128}, "
                          "k %% 128 == 0, labels, ws)");
  B2F_REQUIRE(algo != B2F_KM_TCB && algo != B2F_KM_TCH, "b2f_kmeans_predict_ws: B2F_KM_TCB / B2F_KM_TCH need labels, a "
                                                         "workspace, d in {64, 128} and k %% 128 == 0");
  algo = km_pick_algo(algo, k, d, ld, nullptr);
  if (algo == B2F_KM_TC3 || algo == B2F_KM_TC1) {
    // no workspace: the 3xTF32 (or 1xTF32) scores decide alone -- NOT certified; callers that need the fp64
    // guarantee pass a workspace.  Distances then come from the labels.
    B2F_REQUIRE(km_tc_supported(k, d, ld), "b2f_kmeans_predict: tensor-core path needs d %% 32 == 0, d <= 128, k %% 128 == 0");
    if (labels) {
      int rc = km_tc_step(X, n, d, ld, prep, k, nullptr, labels, algo == B2F_KM_TC3 ? 3 : 1, false, nullptr, nullptr,
                          b2f_cs(stream));
      if (rc != B2F_OK) return rc;
      if (dist2) return km_dist_to_label(X, n, d, ld, nullptr, prep, k, labels, dist2, b2f_cs(stream));
      return B2F_OK;
    }
  }
  return km_generic_launch(X, n, d, ld, nullptr, reinterpret_cast<const float*>(pp + P.off_c32), k, nullptr, nullptr,
                           labels, dist2, false, b2f_cs(stream));
}
