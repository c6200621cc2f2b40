// K2 fused Lloyd kernel for large k*d on the 5th-generation tensor cores (sm_100a only).
//
//   score[i, j] = ||c_j||^2 - 2 x_i . c_j      (argmin_j score == argmin_j ||x_i - c_j||^2)
//
// The contraction runs as tcgen05.mma.kind::tf32 with operands staged by TMA into 128B-swizzled
// shared memory and fp32 accumulators in TMEM.  NPASS = 3 executes the 3xTF32 split
// (x_hi.c_hi + x_hi.c_lo + x_lo.c_hi, every operand explicitly rounded to TF32 so the result does
// not depend on how the tensor core treats low mantissa bits) which holds ~fp32 accuracy;
// NPASS = 1 is the single-pass filter variant.
//
// Replaces MLlib's findClosest + axpy loop behind KMeans.fit;
//   reference call site: orangecontrib/spark/base/spark_ml_estimator.py:22.
//
// CTA = 512 threads, persistent (one CTA per SM), warp roles:
//   warp 0  X producer   : TMA loads of the 128-row X tile (KC chunks of 128 rows x 32 floats)
//   warp 1  MMA issuer   : one elected thread issues tcgen05.mma, commits to mbarriers
//   warp 2  B producer   : TMA loads of centre operand stages (C_hi / C_lo chunk of 128 centres)
//   warp 3  TMEM allocator
//   warps 4-7 (128 thr)  : epilogue: tcgen05.ld, + ||c||^2, running argmin (thread == row), labels
//   warps 8-11 (128 thr) : split each landed X chunk into explicit TF32 hi / lo operands (tcgen05.st
//                          into tensor memory: the MMA takes A from TMEM)
//   warps 12-15 (128 thr): row -> centroid accumulation of finished tiles (16-byte vector reductions
//                          into fp32 replicas), cost and counts -- decoupled from the argmin and
//                          split warps so nothing stalls the tensor pipe.
#include "tc_common.cuh"

namespace {

struct Smem {
  // lives behind the operand tiles in dynamic shared memory (see km_tc_kernel for the carve-up)
  // per 32-column chunk of the X tile: staging (smem) full/empty, operand (TMEM) full/empty
  uint64_t xs_full[4], xs_empty[4], xa_full[4], xa_empty[4];
  uint64_t b_full[8], b_empty[8];
  uint64_t t_full[2], t_empty[2];
  uint64_t a_full[2], a_empty[2];  // epilogue -> accumulate hand-off (labels / scores of one row tile)
  uint64_t n_full[2], n_empty[2];  // split -> epilogue: squared row norms (only used by the certainty test)
  uint32_t tmem_base;
  int lab[2][TM];                  // label, or -1 for a row handed to the exact re-check
  float score[2][TM];
  float xnorm2[2][TM];
};

}  // namespace

template <int NPASS>
__global__ void __launch_bounds__(NTHREADS, 1)
    km_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_ch,
                 const __grid_constant__ CUtensorMap map_cl, const float* __restrict__ X, int64_t n, int d, int64_t ld,
                 const float* __restrict__ cn, int k, int nstages, double* __restrict__ payload,
                 float* __restrict__ rep, unsigned long long* __restrict__ cnt, int32_t* __restrict__ labels,
                 int accumulate, int dbg, const int* __restrict__ n_dev, KmRecheck rck) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];  // 128B swizzle atoms need 1024-byte alignment
  uint8_t* base = smem_raw;
  if (smem_u32(base) & 1023u) __trap();
  const int KC = d / 32;
  uint8_t* xs = base;  // TMA staging of the raw fp32 X tile (KC chunks of 128 rows x 128 B, 128B swizzle)
  uint8_t* stages = xs + (size_t)KC * CHUNK_BYTES;
  const int stage_bytes = (NPASS == 3 ? 2 : 1) * CHUNK_BYTES;
  Smem* sm = reinterpret_cast<Smem*>(stages + (size_t)nstages * stage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (n_dev) n = min(n, (int64_t)max(*n_dev, 0));  // row count produced on the device (re-check pass)
  const int64_t ntiles = (n + TM - 1) / TM;
  const int NT = k / TN;

  if (threadIdx.x == 0) {
    for (int c = 0; c < 4; ++c) {
      mbar_init(&sm->xs_full[c], 1);
      mbar_init(&sm->xs_empty[c], EPI_THREADS);
      mbar_init(&sm->xa_full[c], EPI_THREADS);
      mbar_init(&sm->xa_empty[c], 1);
    }
    for (int s = 0; s < 8; ++s) {
      mbar_init(&sm->b_full[s], 1);
      mbar_init(&sm->b_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&sm->t_full[b], 1);
      mbar_init(&sm->t_empty[b], EPI_THREADS);
      mbar_init(&sm->a_full[b], EPI_THREADS);
      mbar_init(&sm->a_empty[b], EPI_THREADS);
      mbar_init(&sm->n_full[b], EPI_THREADS);
      mbar_init(&sm->n_empty[b], EPI_THREADS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm->tmem_base)),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  if (warp == 0) {
    // ===================== X producer: one TMA per 32-column chunk, chunk-granular reuse ==========
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        for (int kc = 0; kc < KC; ++kc) {
          mbar_wait(&sm->xs_empty[kc], (it & 1) ^ 1);  // the split warps have read the previous tile's chunk
          mbar_expect_tx(&sm->xs_full[kc], (uint32_t)CHUNK_BYTES);
          tma_load_2d(xs + (size_t)kc * CHUNK_BYTES, &map_x, kc * 32, (int)(tile * TM), &sm->xs_full[kc]);
        }
      }
    }
  } else if (warp == 2) {
    // ===================== B producer =====================
    if (lane == 0) {
      uint32_t s = 0, ph = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int nt = 0; nt < NT; ++nt) {
          for (int kc = 0; kc < KC; ++kc) {
            if ((dbg & 1) && ph) continue;
            mbar_wait(&sm->b_empty[s], ph ^ 1);
            mbar_expect_tx(&sm->b_full[s], (uint32_t)stage_bytes);
            uint8_t* dst = stages + (size_t)s * stage_bytes;
            tma_load_2d(dst, &map_ch, kc * 32, nt * TN, &sm->b_full[s]);
            if (NPASS == 3) tma_load_2d(dst + CHUNK_BYTES, &map_cl, kc * 32, nt * TN, &sm->b_full[s]);
            if (++s == (uint32_t)nstages) {
              s = 0;
              ph ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp stays converged (waits are executed by all lanes); one elected lane issues the
    // tcgen05 instructions, so descriptors live in uniform registers and the issue loop stays short.
    const uint64_t desc_fixed = (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
    const uint32_t st_a = smem_u32(stages) >> 4;
    const uint32_t chunk16 = CHUNK_BYTES >> 4, stage16 = (uint32_t)stage_bytes >> 4;
    uint32_t it = 0, acc_it = 0, stage = 0, bphase = 0;
    const int KB = (KC % 2 == 0) ? 2 : 1;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      for (int nt = 0; nt < NT; ++nt, ++acc_it) {
        const uint32_t buf = acc_it & 1;
        mbar_wait(&sm->t_empty[buf], ((acc_it >> 1) & 1) ^ 1);  // epilogue drained this accumulator
        const uint32_t tmem_d = tmem_base + buf * TN;
        // Issue in batches of KB chunks: the tensor queue is only a few instructions deep, so the fixed
        // cost of one wait / elect / commit round must be amortised over a long run of MMAs.
        for (int kc0 = 0; kc0 < KC; kc0 += KB) {
          uint32_t st_idx[2];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            if (j < KB) {
              if (nt == 0) mbar_wait(&sm->xa_full[kc0 + j], it & 1);  // X chunk (hi / lo) sits in TMEM
              if (!(dbg & 1) || bphase == 0) mbar_wait(&sm->b_full[stage], bphase);
              st_idx[j] = stage;
              if (++stage == (uint32_t)nstages) {
                stage = 0;
                bphase ^= 1;
              }
            }
          }
          tc_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              if (j < KB) {
                const int kc = kc0 + j;
                const uint32_t a_hi = tmem_base + TMEM_A_HI + kc * 32;  // 32 TMEM columns per chunk
                const uint32_t a_lo = tmem_base + TMEM_A_LO + kc * 32;
                const uint64_t b_hi = desc_fixed | (uint64_t)(st_a + st_idx[j] * stage16);
                const uint64_t b_lo = b_hi + chunk16;
#pragma unroll
                for (int k8 = 0; k8 < 4; ++k8) {  // 8 tf32 = 32 bytes = 2 descriptor units along K
                  tc_mma_tf32_ts(tmem_d, a_hi + 8 * k8, b_hi + 2 * k8, IDESC_TF32, (kc | k8) ? 1u : 0u);
                  if (NPASS == 3) {
                    tc_mma_tf32_ts(tmem_d, a_hi + 8 * k8, b_lo + 2 * k8, IDESC_TF32, 1u);
                    tc_mma_tf32_ts(tmem_d, a_lo + 8 * k8, b_hi + 2 * k8, IDESC_TF32, 1u);
                  }
                }
                tc_commit(&sm->b_empty[st_idx[j]]);                  // stage may be refilled once these retire
                if (nt == NT - 1) tc_commit(&sm->xa_empty[kc]);      // TMEM X chunk free for the next tile
              }
            }
            if (kc0 + KB >= KC) tc_commit(&sm->t_full[buf]);        // accumulator complete
          }
          __syncwarp();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ===================== epilogue: running argmin, thread == row =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    uint32_t it = 0, acc_it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      // two independent chains (even / odd columns) keep the min/max dependency chains short
      float b1a = INFINITY, b2a = INFINITY, b1b = INFINITY, b2b = INFINITY;
      int i1a = 0, i1b = 1;
      for (int nt = 0; nt < NT; ++nt, ++acc_it) {
        const int buf = acc_it & 1;
        mbar_wait(&sm->t_full[buf], (acc_it >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + lane_addr + (uint32_t)(buf * TN);
        uint32_t v[2][32];
        tmem_ld32(taddr, v[0]);
#pragma unroll
        for (int c = 0; c < TN / 32; ++c) {
          tmem_ld_wait();
          if (c + 1 < TN / 32) tmem_ld32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
          const float4* cn4 = reinterpret_cast<const float4*>(cn + nt * TN + c * 32);  // warp-uniform, L1 hits
          const int jbase = nt * TN + c * 32;
          if (dbg & 2) continue;
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 cv = __ldg(cn4 + j4);
            const float s0 = __uint_as_float(v[c & 1][j4 * 4 + 0]) + cv.x;
            const float s1 = __uint_as_float(v[c & 1][j4 * 4 + 1]) + cv.y;
            const float s2 = __uint_as_float(v[c & 1][j4 * 4 + 2]) + cv.z;
            const float s3 = __uint_as_float(v[c & 1][j4 * 4 + 3]) + cv.w;
            const int j = jbase + j4 * 4;
            i1a = (s0 < b1a) ? j : i1a;       b2a = fminf(b2a, fmaxf(b1a, s0)); b1a = fminf(b1a, s0);
            i1b = (s1 < b1b) ? j + 1 : i1b;   b2b = fminf(b2b, fmaxf(b1b, s1)); b1b = fminf(b1b, s1);
            i1a = (s2 < b1a) ? j + 2 : i1a;   b2a = fminf(b2a, fmaxf(b1a, s2)); b1a = fminf(b1a, s2);
            i1b = (s3 < b1b) ? j + 3 : i1b;   b2b = fminf(b2b, fmaxf(b1b, s3)); b1b = fminf(b1b, s3);
          }
        }
        tc_fence_before();
        mbar_arrive(&sm->t_empty[buf]);
      }
      // merge the chains: smaller score wins, equal scores keep the lower centre index
      const bool a_wins = (b1a < b1b) || (b1a == b1b && i1a < i1b);
      const float b1 = a_wins ? b1a : b1b;
      const int i1 = a_wins ? i1a : i1b;
      const int64_t row = tile * TM + row_in_tile;
      if (labels && row < n) labels[row] = i1;
      bool certain = true;
      if (rck.count) {
        // certainty test of the 3xTF32 scores: rows whose two best scores are within 2 tau go to the exact
        // (fp64) kernel instead of being accumulated here
        const float b2 = fminf(fmaxf(b1a, b1b), fminf(b2a, b2b));
        mbar_wait(&sm->n_full[it & 1], (it >> 1) & 1);
        // (second term: fp32 roundings of ||c||^2 and of the final add, relevant for rows near the origin)
        const float cmx = *rck.cmax;
        const float tau2 = rck.kappa2 * sqrtf(sm->xnorm2[it & 1][row_in_tile]) * cmx + 3.6e-7f * cmx * cmx;
        mbar_arrive(&sm->n_empty[it & 1]);
        certain = (b2 - b1) > tau2;
        const bool list_it = row < n && !certain;
        const unsigned m = __ballot_sync(0xffffffffu, list_it);
        if (m) {
          int base = 0;
          if (lane == (__ffs(m) - 1)) base = atomicAdd(rck.count, __popc(m));
          base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
          if (list_it) {
            const int slot = base + __popc(m & ((1u << lane) - 1));
            if (slot < rck.cap) rck.idx[slot] = rck.src_idx ? rck.src_idx[row] : (int32_t)row;
          }
        }
      }
      if (accumulate) {
        const int ab = it & 1;
        mbar_wait(&sm->a_empty[ab], ((it >> 1) & 1) ^ 1);
        sm->lab[ab][row_in_tile] = certain ? i1 : -1;
        sm->score[ab][row_in_tile] = b1;
        mbar_arrive(&sm->a_full[ab]);  // release semantics order the two stores before the arrive
      }
    }
  } else if (warp >= 8 && warp < 12) {
    // ===================== split X into explicit TF32 hi / lo operands in tensor memory =====================
    const int w4 = warp - 8;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      // thread == row: read the row's 32 floats of each chunk from the swizzled staging tile
      // (16-byte piece p of row r sits at r*128 + ((p ^ (r & 7)) << 4)), split into explicit TF32
      // hi / lo, and store them as the A operand in tensor memory (lane = row, column = k).
      const int r = w4 * 32 + lane;
      const uint32_t lane_addr = (uint32_t)(w4 * 32) << 16;
      float xn = 0.f;
      for (int kc = 0; kc < KC; ++kc) {
        mbar_wait(&sm->xs_full[kc], it & 1);
        const uint8_t* src = xs + (size_t)kc * CHUNK_BYTES + (size_t)r * 128;
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          const float4 v = *reinterpret_cast<const float4*>(src + ((p ^ (r & 7)) << 4));
          xn = fmaf(v.x, v.x, xn); xn = fmaf(v.y, v.y, xn); xn = fmaf(v.z, v.z, xn); xn = fmaf(v.w, v.w, xn);
          const float h0 = to_tf32(v.x), h1 = to_tf32(v.y), h2 = to_tf32(v.z), h3 = to_tf32(v.w);
          hi[p * 4 + 0] = __float_as_uint(h0); hi[p * 4 + 1] = __float_as_uint(h1);
          hi[p * 4 + 2] = __float_as_uint(h2); hi[p * 4 + 3] = __float_as_uint(h3);
          if (NPASS == 3) {
            lo[p * 4 + 0] = __float_as_uint(to_tf32(v.x - h0)); lo[p * 4 + 1] = __float_as_uint(to_tf32(v.y - h1));
            lo[p * 4 + 2] = __float_as_uint(to_tf32(v.z - h2)); lo[p * 4 + 3] = __float_as_uint(to_tf32(v.w - h3));
          }
        }
        // (proxy fence: generic reads of the chunk before its refill by TMA; see kmeans_tcb.cu)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_arrive(&sm->xs_empty[kc]);                 // staging chunk may be refilled (next tile)
        mbar_wait(&sm->xa_empty[kc], (it & 1) ^ 1);     // previous tile's MMAs are done with this TMEM chunk
        tc_fence_after();
        tmem_st32(tmem_base + lane_addr + TMEM_A_HI + kc * 32, hi);
        if (NPASS == 3) tmem_st32(tmem_base + lane_addr + TMEM_A_LO + kc * 32, lo);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&sm->xa_full[kc]);
      }
      // The slot is read by the epilogue's certainty test AFTER it has released the tile's last accumulator, and
      // nothing else holds the split warps back from being two row tiles ahead by then (one centre tile per row
      // tile: wide open; more: a short window).  Without this wait the norms of tile it - 2 could be overwritten
      // and n_full run two phases ahead of its waiter, whose parity wait then never returns (seen as a time-out
      // at k = 128 with a few row tiles per CTA).
      if (rck.count) mbar_wait(&sm->n_empty[it & 1], ((it >> 1) & 1) ^ 1);
      sm->xnorm2[it & 1][r] = xn;  // read by the epilogue's certainty test after this tile's last centre tile
      mbar_arrive(&sm->n_full[it & 1]);
    }
  } else if (warp >= 12) {
    // ===================== accumulate finished row tiles (own warps: never delays the split) ==========
    const int w4 = warp - 12;
    double cost = 0.0;
    unsigned long long nrows = 0;
    float* myrep = rep ? rep + (size_t)(blockIdx.x % KM_REPLICAS) * k * d : nullptr;
    const int nch = (d + 127) / 128;
    uint32_t it = 0;
    if (accumulate) {
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        // rows of the previous tile: labels / scores come from the epilogue warps, the row values are
        // re-read from global memory (L2 hits: the tile was just streamed in by TMA)
        const uint32_t pit = it;
        const int ab = pit & 1;
        mbar_wait(&sm->a_full[ab], (pit >> 1) & 1);
        const int my_lab = sm->lab[ab][w4 * 32 + lane];
        const float my_sc = sm->score[ab][w4 * 32 + lane];
        mbar_arrive(&sm->a_empty[ab]);
        const int64_t r0 = tile * TM + w4 * 32;
        for (int g = 0; g < 32; g += 8) {
          float4 xv[8];
  #pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int64_t rr = r0 + g + u;
            xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rr < n && lane * 4 < d) xv[u] = *reinterpret_cast<const float4*>(X + rr * ld + lane * 4);
          }
  #pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int64_t rr = r0 + g + u;
            const int lbl = __shfl_sync(0xffffffffu, my_lab, g + u);
            const float sc = __shfl_sync(0xffffffffu, my_sc, g + u);
            if (rr < n && lbl >= 0) {  // warp-uniform
              float xn = 0.f;
              if (lane * 4 < d) {
                xn = fmaf(xv[u].x, xv[u].x, xn); xn = fmaf(xv[u].y, xv[u].y, xn);
                xn = fmaf(xv[u].z, xv[u].z, xn); xn = fmaf(xv[u].w, xv[u].w, xn);
                red_add_v4(myrep + (size_t)lbl * d + lane * 4, xv[u]);
              }
              for (int ch = 1; ch < nch; ++ch) {  // d > 128 (single-pass variant only)
                const int c = ch * 128 + lane * 4;
                if (c < d) {
                  const float4 x2 = *reinterpret_cast<const float4*>(X + rr * ld + c);
                  xn = fmaf(x2.x, x2.x, xn); xn = fmaf(x2.y, x2.y, xn);
                  xn = fmaf(x2.z, x2.z, xn); xn = fmaf(x2.w, x2.w, xn);
                  red_add_v4(myrep + (size_t)lbl * d + c, x2);
                }
              }
              xn = warp_sum(xn);
              if (lane == 0) {
                cost += (double)fmaxf(xn + sc, 0.f);
                atomicAdd(cnt + lbl, 1ull);
                ++nrows;
              }
            }
          }
        }
      }
    }
    if (accumulate && lane == 0) {
      atomicAdd(payload + (size_t)k * d + k, cost);
      atomicAdd(payload + (size_t)k * d + k + 1, (double)nrows);
    }
  }
  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static const int kSmemLimit = 227 * 1024;

static int km_tc_stages(int d, int passes) {
  const int KC = d / 32;
  const int xbytes = KC * CHUNK_BYTES;  // fp32 staging only: the hi / lo operands live in tensor memory
  const int stage = (passes == 3 ? 2 : 1) * CHUNK_BYTES;
  const int left = kSmemLimit - (int)sizeof(Smem) - 64 - xbytes;
  int s = left / stage;
  if (s > 8) s = 8;
  return s;
}

bool km_tc_supported(int k, int d, int64_t ld) {
  if (d % 32 != 0 || d < 32 || d > 128) return false;  // 3xTF32 keeps X hi+lo tiles resident: d <= 128
  if (k % TN != 0 || k < TN) return false;
  if (ld % 4 != 0) return false;
  return km_tc_stages(d, 3) >= 2;
}

int km_tc_step(const float* X, int64_t n, int d, int64_t ld, const void* prep, int k, void* acc, int32_t* labels,
               int passes, bool accumulate, const int* n_dev, const KmRecheck* recheck, cudaStream_t stream) {
  B2F_REQUIRE(passes == 1 || passes == 3, "km_tc_step: passes must be 1 or 3");
  B2F_REQUIRE(n < (1ll << 31) * (int64_t)TM, "km_tc_step: too many rows for one launch");
  const int nstages = km_tc_stages(d, passes);
  B2F_REQUIRE(nstages >= 2, "km_tc_step: d=%d does not fit in shared memory", d);
  KmPrepLayout P = km_prep_layout(k, d);
  const char* pp = static_cast<const char*>(prep);
  const float* ch = reinterpret_cast<const float*>(pp + P.off_ch);
  const float* cl = reinterpret_cast<const float*>(pp + P.off_cl);
  const float* cn = reinterpret_cast<const float*>(pp + P.off_cn);
  CUtensorMap mx, mh, ml;
  int rc;
  if ((rc = make_map_f32(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)ld)) != B2F_OK) return rc;
  if ((rc = make_map_f32(&mh, ch, (uint64_t)k, (uint64_t)d, (uint64_t)d)) != B2F_OK) return rc;
  if ((rc = make_map_f32(&ml, cl, (uint64_t)k, (uint64_t)d, (uint64_t)d)) != B2F_OK) return rc;
  double* payload = nullptr;
  float* rep = nullptr;
  unsigned long long* cnt = nullptr;
  if (accumulate) {
    KmAccLayout A = km_acc_layout(k, d);
    char* ap = static_cast<char*>(acc);
    payload = reinterpret_cast<double*>(ap + A.off_payload);
    rep = reinterpret_cast<float*>(ap + A.off_rep);
    cnt = reinterpret_cast<unsigned long long*>(ap + A.off_cnt);
  }
  const int KC = d / 32;
  const size_t smem = (size_t)KC * CHUNK_BYTES +
                      (size_t)nstages * (passes == 3 ? 2 : 1) * CHUNK_BYTES + sizeof(Smem) + 64;
  const int64_t ntiles = (n + TM - 1) / TM;
  int grid = b2f_sm_count();
#ifdef B2F_DEV  // dev knobs of the timing experiments; not in the product build
  if (const char* g = getenv("B2F_TC_GRID")) grid = atoi(g) > 0 ? atoi(g) : grid;
#endif
  if ((int64_t)grid > ntiles) grid = (int)ntiles;
  KmRecheck rck = {};
  if (recheck) rck = *recheck;
  int dbg = 0;
#ifdef B2F_DEV
  dbg = getenv("B2F_TC_DBG") ? atoi(getenv("B2F_TC_DBG")) : 0;  // results invalid when set
#endif
  if (passes == 3) {
    B2F_CUDA(b2f_smem_attr(km_tc_kernel<3>, smem));
    km_tc_kernel<3><<<grid, NTHREADS, smem, stream>>>(mx, mh, ml, X, n, d, ld, cn, k, nstages, payload, rep, cnt,
                                                      labels, accumulate ? 1 : 0, dbg, n_dev, rck);
  } else {
    B2F_CUDA(b2f_smem_attr(km_tc_kernel<1>, smem));
    km_tc_kernel<1><<<grid, NTHREADS, smem, stream>>>(mx, mh, ml, X, n, d, ld, cn, k, nstages, payload, rep, cnt,
                                                      labels, accumulate ? 1 : 0, dbg, n_dev, rck);
  }
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
