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
// CTA = 256 threads, persistent (one CTA per SM), warp roles:
//   warp 0  X producer   : TMA loads of the 128-row X tile (KC chunks of 128 rows x 32 floats)
//   warp 1  MMA issuer   : one elected thread issues tcgen05.mma, commits to mbarriers
//   warp 2  B producer   : TMA loads of centre operand stages (C_hi / C_lo chunk of 128 centres)
//   warp 3  TMEM allocator
//   warps 4-7 (128 thr)  : split X into hi/lo in place, then epilogue: tcgen05.ld, + ||c||^2,
//                          running argmin (thread == row), labels, cost, and the row -> centroid
//                          accumulation with 16-byte vector reductions into fp32 replicas.
#include <cuda.h>

#include "common.cuh"

namespace {

constexpr int TM = 128;            // rows per tile == UMMA M == TMEM lanes
constexpr int TN = 128;            // centres per accumulator tile == UMMA N
constexpr int CHUNK_BYTES = TM * 128;  // 128 rows x 32 floats, 128B swizzle atom rows
constexpr int NTHREADS = 256;
constexpr int EPI_THREADS = 128;
constexpr uint32_t TMEM_COLS = 256;  // two accumulator buffers of TN columns

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap, never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) {
      printf("b200fit km_tc: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void red_add_v4(float* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// K-major, 128B-swizzled operand tile whose 8-row groups are 1024 B apart (SM100 UMMA descriptor:
// start>>4 | LBO(=1, unused for swizzled K-major)<<16 | SBO(1024>>4)<<32 | version 1<<46 | SWIZZLE_128B(2)<<61)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::tf32, fp32 accumulate, A and B K-major, M = 128, N = TN
constexpr uint32_t IDESC_TF32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

struct Smem {
  // dynamic shared memory, carved manually (1024-byte aligned): XH[KC] | XL[KC] (NPASS==3) | stages
  uint64_t x_full, x_split, x_empty;
  uint64_t b_full[8], b_empty[8];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base;
  float cn[2][TN];
};

}  // namespace

template <int NPASS>
__global__ void __launch_bounds__(NTHREADS, 1)
    km_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_ch,
                 const __grid_constant__ CUtensorMap map_cl, const float* __restrict__ X, int64_t n, int d, int64_t ld,
                 const float* __restrict__ cn, int k, int nstages, double* __restrict__ payload,
                 float* __restrict__ rep, unsigned long long* __restrict__ cnt, int32_t* __restrict__ labels,
                 int accumulate) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int KC = d / 32;
  uint8_t* xh = base;
  uint8_t* xl = xh + (size_t)KC * CHUNK_BYTES;
  uint8_t* stages = (NPASS == 3) ? xl + (size_t)KC * CHUNK_BYTES : xl;
  const int stage_bytes = (NPASS == 3 ? 2 : 1) * CHUNK_BYTES;
  Smem* sm = reinterpret_cast<Smem*>(stages + (size_t)nstages * stage_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntiles = (n + TM - 1) / TM;
  const int NT = k / TN;

  if (threadIdx.x == 0) {
    mbar_init(&sm->x_full, 1);
    mbar_init(&sm->x_split, EPI_THREADS);
    mbar_init(&sm->x_empty, 1);
    for (int s = 0; s < nstages; ++s) {
      mbar_init(&sm->b_full[s], 1);
      mbar_init(&sm->b_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&sm->t_full[b], 1);
      mbar_init(&sm->t_empty[b], EPI_THREADS);
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
    // ===================== X producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        mbar_wait(&sm->x_empty, (it & 1) ^ 1);  // all MMAs of the previous tile have read X
        mbar_expect_tx(&sm->x_full, (uint32_t)(KC * CHUNK_BYTES));
        for (int kc = 0; kc < KC; ++kc)
          tma_load_2d(xh + (size_t)kc * CHUNK_BYTES, &map_x, kc * 32, (int)(tile * TM), &sm->x_full);
      }
    }
  } else if (warp == 2) {
    // ===================== B producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int nt = 0; nt < NT; ++nt) {
          for (int kc = 0; kc < KC; ++kc, ++it) {
            const int s = it % nstages;
            const uint32_t ph = (it / nstages) & 1;
            mbar_wait(&sm->b_empty[s], ph ^ 1);
            mbar_expect_tx(&sm->b_full[s], (uint32_t)stage_bytes);
            uint8_t* dst = stages + (size_t)s * stage_bytes;
            tma_load_2d(dst, &map_ch, kc * 32, nt * TN, &sm->b_full[s]);
            if (NPASS == 3) tma_load_2d(dst + CHUNK_BYTES, &map_cl, kc * 32, nt * TN, &sm->b_full[s]);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      uint32_t it = 0, bit = 0, acc_it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        mbar_wait(&sm->x_split, it & 1);  // X tile landed and was split into hi / lo
        tc_fence_after();
        for (int nt = 0; nt < NT; ++nt, ++acc_it) {
          const int buf = acc_it & 1;
          mbar_wait(&sm->t_empty[buf], ((acc_it >> 1) & 1) ^ 1);  // epilogue drained this accumulator
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(buf * TN);
          for (int kc = 0; kc < KC; ++kc, ++bit) {
            const int s = bit % nstages;
            mbar_wait(&sm->b_full[s], (bit / nstages) & 1);
            tc_fence_after();
            const uint32_t a_hi = smem_u32(xh + (size_t)kc * CHUNK_BYTES);
            const uint32_t a_lo = smem_u32(xl + (size_t)kc * CHUNK_BYTES);
            const uint32_t b_hi = smem_u32(stages + (size_t)s * stage_bytes);
            const uint32_t b_lo = b_hi + CHUNK_BYTES;
#pragma unroll
            for (int k8 = 0; k8 < 4; ++k8) {
              const uint32_t off = k8 * 32;  // 8 tf32 = 32 bytes along K inside the swizzle atom
              tc_mma_tf32(tmem_d, umma_desc_sw128(a_hi + off), umma_desc_sw128(b_hi + off), IDESC_TF32,
                          (kc | k8) ? 1u : 0u);
              if (NPASS == 3) {
                tc_mma_tf32(tmem_d, umma_desc_sw128(a_hi + off), umma_desc_sw128(b_lo + off), IDESC_TF32, 1u);
                tc_mma_tf32(tmem_d, umma_desc_sw128(a_lo + off), umma_desc_sw128(b_hi + off), IDESC_TF32, 1u);
              }
            }
            tc_commit(&sm->b_empty[s]);  // stage may be refilled once these MMAs retire
          }
          tc_commit(&sm->t_full[buf]);  // accumulator complete
        }
        tc_commit(&sm->x_empty);  // X tile free for the next TMA
      }
    }
  } else if (warp >= 4) {
    // ===================== split + epilogue (128 threads, thread == row) =====================
    const int et = threadIdx.x - 128;
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    double cost = 0.0;
    unsigned long long nrows = 0;
    uint32_t it = 0, acc_it = 0;
    float* myrep = rep ? rep + (size_t)(blockIdx.x % KM_REPLICAS) * k * d : nullptr;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      // ---- split x into explicit TF32 hi (in place) and lo
      mbar_wait(&sm->x_full, it & 1);
      {
        float4* h4 = reinterpret_cast<float4*>(xh);
        float4* l4 = reinterpret_cast<float4*>(xl);
        const int nvec = KC * CHUNK_BYTES / 16;
        for (int i = et; i < nvec; i += EPI_THREADS) {
          const float4 v = h4[i];
          float4 hi, lo;
          hi.x = to_tf32(v.x); hi.y = to_tf32(v.y); hi.z = to_tf32(v.z); hi.w = to_tf32(v.w);
          h4[i] = hi;
          if (NPASS == 3) {
            lo.x = to_tf32(v.x - hi.x); lo.y = to_tf32(v.y - hi.y); lo.z = to_tf32(v.z - hi.z); lo.w = to_tf32(v.w - hi.w);
            l4[i] = lo;
          }
        }
        fence_proxy_async();  // generic-proxy writes -> visible to the tensor core (async proxy)
        mbar_arrive(&sm->x_split);
      }
      // ---- running argmin over all centre tiles
      float b1 = INFINITY, b2 = INFINITY;
      int i1 = 0;
      for (int nt = 0; nt < NT; ++nt, ++acc_it) {
        const int buf = acc_it & 1;
        sm->cn[buf][et] = cn[nt * TN + et];
        epi_bar_sync();
        mbar_wait(&sm->t_full[buf], (acc_it >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + lane_addr + (uint32_t)(buf * TN);
#pragma unroll
        for (int c0 = 0; c0 < TN; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(taddr + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float s = __uint_as_float(v[j]) + sm->cn[buf][c0 + j];
            const bool lt = s < b1;
            b2 = fminf(b2, fmaxf(b1, s));
            b1 = fminf(b1, s);
            i1 = lt ? (nt * TN + c0 + j) : i1;
          }
        }
        tc_fence_before();
        mbar_arrive(&sm->t_empty[buf]);
      }
      // ---- labels, cost, centroid accumulation (rows re-read from L2, one warp per row)
      const int64_t row = tile * TM + row_in_tile;
      if (labels && row < n) labels[row] = i1;
      if (accumulate) {
        const int nch = (d + 127) / 128;
        for (int r = 0; r < 32; ++r) {
          const int64_t rr = tile * TM + q * 32 + r;
          if (rr >= n) break;
          const int lbl = __shfl_sync(0xffffffffu, i1, r);
          const float sc = __shfl_sync(0xffffffffu, b1, r);
          float xn = 0.f;
          for (int ch = 0; ch < nch; ++ch) {
            const int c = ch * 128 + lane * 4;
            if (c < d) {
              const float4 xv = *reinterpret_cast<const float4*>(X + rr * ld + c);
              xn = fmaf(xv.x, xv.x, xn); xn = fmaf(xv.y, xv.y, xn); xn = fmaf(xv.z, xv.z, xn); xn = fmaf(xv.w, xv.w, xn);
              red_add_v4(myrep + (size_t)lbl * d + c, xv);
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
    if (accumulate) {
      // one double atomic per warp
      if (lane == 0) {
        atomicAdd(payload + (size_t)k * d + k, cost);
        atomicAdd(payload + (size_t)k * d + k + 1, (double)nrows);
      }
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
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D fp32 row-major [rows, cols] (row pitch `pitch` floats), box = 32 floats x 128 rows, 128B swizzle
static int make_map(CUtensorMap* m, const float* base, uint64_t rows, uint64_t cols, uint64_t pitch) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    b2f_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return B2F_ECUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch * sizeof(float)};
  cuuint32_t box[2] = {32, (cuuint32_t)TM};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b2f_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu pitch=%llu)", (int)r,
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)pitch);
    return B2F_ECUDA;
  }
  return B2F_OK;
}

static const int kSmemLimit = 227 * 1024;

static int km_tc_stages(int d, int passes) {
  const int KC = d / 32;
  const int xbytes = (passes == 3 ? 2 : 1) * KC * CHUNK_BYTES;
  const int stage = (passes == 3 ? 2 : 1) * CHUNK_BYTES;
  const int left = kSmemLimit - 1024 /*align*/ - (int)sizeof(Smem) - 64 - xbytes;
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
               int passes, bool accumulate, cudaStream_t stream) {
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
  if ((rc = make_map(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)ld)) != B2F_OK) return rc;
  if ((rc = make_map(&mh, ch, (uint64_t)k, (uint64_t)d, (uint64_t)d)) != B2F_OK) return rc;
  if ((rc = make_map(&ml, cl, (uint64_t)k, (uint64_t)d, (uint64_t)d)) != B2F_OK) return rc;
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
  const size_t smem = 1024 + (size_t)(passes == 3 ? 2 : 1) * KC * CHUNK_BYTES +
                      (size_t)nstages * (passes == 3 ? 2 : 1) * CHUNK_BYTES + sizeof(Smem) + 64;
  const int64_t ntiles = (n + TM - 1) / TM;
  int grid = b2f_sm_count();
  if ((int64_t)grid > ntiles) grid = (int)ntiles;
  if (passes == 3) {
    B2F_CUDA(cudaFuncSetAttribute(km_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    km_tc_kernel<3><<<grid, NTHREADS, smem, stream>>>(mx, mh, ml, X, n, d, ld, cn, k, nstages, payload, rep, cnt,
                                                      labels, accumulate ? 1 : 0);
  } else {
    B2F_CUDA(cudaFuncSetAttribute(km_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    km_tc_kernel<1><<<grid, NTHREADS, smem, stream>>>(mx, mh, ml, X, n, d, ld, cn, k, nstages, payload, rep, cnt,
                                                      labels, accumulate ? 1 : 0);
  }
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}
