// K4 on the tensor cores: fused multinomial (softmax) LogisticRegression loss + gradient pass
// (sm_100a only).  Per 128-row tile, with K classes padded to N = 32:
//   GEMM1  margins[128 x 32]  = X_tile[128 x d] . W'^T            tcgen05 kind::tf32, 3xTF32 split,
//                                                                 A (x_hi / x_lo) from tensor memory
//   softmax / loss / (P - Y) per row (thread == row), written to shared memory as bf16 pairs
//   GEMM2  G^T[d x 32]       += X_tile^T[d x 128] . (P - Y)[128 x 32]   tcgen05 kind::f16, bf16x3,
//                                                                 A = X^T read MN-major from smem
//          as x1 . [p1 | p2] + x2 . [p1 | p2] (N = 64 instructions: each X piece is read from shared memory once)
// X is streamed from HBM once (GEMM1) and re-read from L2 for GEMM2; G^T stays in tensor memory for the
// CTA's lifetime.
// What bounds it (hand-off trace, tools/lr_trace.py): shared-memory bandwidth.  Per 128 x 512 tile the shared
// memory moves ~1.8 MB (TMA writes 512 KB for the two passes, split-warp loads 512 KB, bf16 stores 256 KB, GEMM2
// operand reads 256 + 96 KB, GEMM1 W reads 192 KB) = 14 600 cycles at 128 B/clk of the 18 600 measured per tile.  Replaces the treeAggregate of MultinomialLogisticBlockAggregator;
//   reference call site: orangecontrib/spark/base/spark_ml_estimator.py:22 with module = classification
//   (orangecontrib/spark/widgets/ml/spark_ml_classification.py:15).
//
// CTA = 512 threads, persistent, one per SM:
//   warp 0  X producer (TMA chunks of 128 rows x 32 floats): GEMM1 chunks of tile t+1 and GEMM2 chunks of t,
//           in the order lt_phase() fixes for all three of producer, split warps and MMA issuer
//   warp 1  MMA issuer (software pipelined: GEMM1(t+1) is issued before / between the blocks of GEMM2(t))
//   warp 2  W producer (TMA, [32 classes x 32 k] of W_hi and W_lo per chunk)
//   warp 3  TMEM allocator
//   warps 4-7   softmax epilogue (thread == row)
//   warps 8-15  split: fp32 chunk -> tf32 hi/lo into tensor memory (GEMM1) or bf16 x1/x2 into the MN-major
//               shared-memory operand tile (GEMM2); two groups of 4 warps (thread == row) alternate chunks
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "tc_common.cuh"

namespace {

constexpr int LT_THREADS = 512;
constexpr int LT_SPLIT_THREADS = 256;
// Both switches are on in the product; -DLT_INTERLEAVE=0 / -DLT_PCAT=0 rebuild the earlier kernel for comparison
// (20M x 512, 20 classes: 12.97 ms with both off, 11.68 interleave only, 12.95 [p1|p2] only, 10.84 with both).
#ifndef LT_INTERLEAVE
#define LT_INTERLEAVE 1  // the last LT_BETWEEN (FB - 1) GEMM1 chunks of tile t+1 go between the GEMM2 blocks of tile t (lt_phase)
#endif
#ifndef LT_BETWEEN
#define LT_BETWEEN 3     // GEMM1 chunks issued after each GEMM2 block but the last (same box, d = 512: 1 -> +10%, 2 -> +2.5%,
                         // 4 -> +1.5% time relative to 3)
#endif
#ifndef LT_PCAT
#define LT_PCAT 1        // GEMM2 reads x1 once against [p1 | p2] (N = 64) instead of twice (N = 32 each)
#endif
// G^T is moved out of tensor memory into the CTA's fp64 partial every LT_FLUSH row tiles.  Measured at 20M x 512
// (135 000 rows per CTA in one fp32 accumulator): 2.7e-4 of the largest gradient entry, i.e. eps/4 per accumulating
// MMA relative to the running sum -- the tensor core TRUNCATES when it adds into the accumulator, so the error is a
// bias that grows with the number of accumulations, not a random walk.  16 accumulations per tile and block:
// LT_FLUSH = 8 keeps the bias near 2e-6.
#ifndef LT_FLUSH
#define LT_FLUSH 8
#endif
constexpr int LT_NP = 32;          // classes padded to the MMA N
#ifndef LT_XS_SLOTS
#define LT_XS_SLOTS 6
#endif
#ifndef LT_W_STAGES
#define LT_W_STAGES 4
#endif
constexpr int LT_XS = LT_XS_SLOTS; // X staging slots (16 KB each); must stay a multiple of the number of split groups (2):
                                   // a slot is then always consumed by the same group and its parity waits cannot alias
constexpr int LT_WS = LT_W_STAGES; // W stages (hi + lo, 8 KB each)
constexpr int LT_WSTAGE = 2 * LT_NP * 128;
// tensor memory columns
constexpr uint32_t LT_TMEM_COLS = 512;
constexpr uint32_t LT_D1 = 0;      // margins, 2 buffers x 32 columns
constexpr uint32_t LT_A1 = 64;     // GEMM1 operand ring: LT_A1_SLOTS slots x (32 hi + 32 lo) columns
constexpr uint32_t LT_A1_SLOTS = LT_PCAT ? 3 : 4;
constexpr uint32_t LT_G = LT_A1 + LT_A1_SLOTS * 64;  // G^T: d/128 blocks x LT_GW columns (features on lanes)
constexpr uint32_t LT_GW = LT_PCAT ? 64 : 32;        // LT_PCAT: columns 0-31 = (x1 + x2).p1, columns 32-63 = (x1 + x2).p2
static_assert(LT_G + 4 * LT_GW <= LT_TMEM_COLS, "tensor memory budget");
// (P - Y) operand tile in shared memory, per buffer 16 KB: bf16 pieces p1, p2, each two K atoms (64 rows of X) of
// 32 class rows x 128 B.  LT_PCAT puts p2's class rows right behind p1's inside an atom so that one N = 64 operand
// covers both
constexpr uint32_t LT_PY_ATOM = LT_PCAT ? 8192 : 4096;  // bytes between the two K atoms of one piece
constexpr uint32_t LT_PY_P2 = LT_PCAT ? 4096 : 8192;    // byte offset of p2 relative to p1
// GEMM1: kind::tf32, A (TMEM) and B K-major, M = 128, N = 32
constexpr uint32_t IDESC_G1 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(LT_NP >> 3) << 17) | (8u << 24);
// GEMM2: kind::f16 bf16, A MN-major (bit 15), B K-major, M = 128, N = 32
constexpr uint32_t IDESC_G2 = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(LT_NP >> 3) << 17) | (8u << 24);
constexpr uint32_t IDESC_G2_N64 = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(64 >> 3) << 17) | (8u << 24);

struct SmemL {
  uint64_t xs_full[LT_XS], xs_empty[LT_XS];
  uint64_t w_full[LT_WS], w_empty[LT_WS];
  uint64_t a1_full[4], a1_empty[4];
  uint64_t d1_full[2], d1_empty[2];
  uint64_t py_full[2], py_empty[2];
  uint64_t a2_full, a2_empty;
  uint64_t g_done;
  uint64_t g_part;   // the GEMM2 blocks of a tile that ends a flush period have completed
  uint32_t tmem_base;
  float icpt[LT_NP];
  double red[4][LT_NP + 2];
};

__device__ __forceinline__ void mma_tf32_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
      "r"(a), "l"(b), "r"(IDESC_G1), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void mma_bf16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
// Work order of one phase, shared by the X producer, the split warps and the MMA issuer: the GEMM1 chunks of one
// tile (has_g1) and the GEMM2 blocks of the tile before it (has_g2).  Plain order: all chunks, then all blocks.
// LT_INTERLEAVE: the last LT_BETWEEN (FB - 1) chunks are held back and issued LT_BETWEEN at a time after each block but the last, so
// that the split warps have work while a block's MMAs drain (the operand tile of GEMM2 is single-buffered), and the
// softmax of the GEMM1 tile still has the whole last block plus the next phase's leading chunks to hide behind.
template <class F1, class F2>
__device__ __forceinline__ void lt_phase(int KC, int FB, bool has_g1, bool has_g2, F1&& g1_chunk, F2&& g2_block) {
  const int pre = (LT_INTERLEAVE && has_g1 && has_g2) ? KC - LT_BETWEEN * (FB - 1) : KC;
  int c = 0;
  if (has_g1)
    for (; c < pre; ++c) g1_chunk(c);
  if (has_g2)
    for (int fb = 0; fb < FB; ++fb) {
      g2_block(fb);
      if (has_g1)
        for (int e = 0; e < LT_BETWEEN && c < KC; ++e, ++c) g1_chunk(c);
    }
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// dev tracing (build with -DLT_TRACE, run with B2F_LR_TRACE=<file>; tools/lr_trace.py prints it): block 0 stamps
// clock64() at the hand-offs of its first LTR_TILES row tiles.  Compiled out of the product library.
#ifdef LT_TRACE
constexpr int LTR_TILES = 48, LTR_SLOTS = 16, LTR_ROLES = 6;  // mma | softmax warp 4 | split group 0 | split group 1 | X producer | inside a chunk (warp 8)
__device__ unsigned long long lt_trace[LTR_ROLES * LTR_TILES * LTR_SLOTS];
#define LTR(role, it, slot)                                                                       \
  do {                                                                                            \
    if (blockIdx.x == 0 && lane == 0 && (it) >= 0 && (it) < LTR_TILES)                            \
      lt_trace[((role) * LTR_TILES + (it)) * LTR_SLOTS + (slot)] = (unsigned long long)clock64(); \
    asm volatile("" ::: "memory");                                                                \
  } while (0)
// LTW(k, stmt): run stmt and add the cycles it took to wait-sum k (warp 8 only); LTW_FLUSH writes the sums of a phase
#define LTW(k, stmt)                                  \
  do {                                                \
    const long long c0_ = clock64();                  \
    stmt;                                             \
    if (w8 == 0) ltw[k] += clock64() - c0_;           \
  } while (0)
#define LTW_FLUSH(it)                                                                              \
  do {                                                                                             \
    if (w8 == 0 && blockIdx.x == 0 && lane == 0 && (it) >= 0 && (it) < LTR_TILES)                  \
      for (int k_ = 0; k_ < 4; ++k_) lt_trace[(5 * LTR_TILES + (it)) * LTR_SLOTS + 9 + k_] = (unsigned long long)ltw[k_]; \
    for (int k_ = 0; k_ < 4; ++k_) ltw[k_] = 0;                                                    \
  } while (0)
#else
#define LTR(role, it, slot) do {} while (0)
#define LTW(k, stmt) stmt
#define LTW_FLUSH(it) do {} while (0)
#endif

}  // namespace

__global__ void __launch_bounds__(LT_THREADS, 1)
    lr_mn_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_wh,
                    const __grid_constant__ CUtensorMap map_wl, const float* __restrict__ y,
                    const float* __restrict__ w, int64_t n, int d, int K, const double* __restrict__ icpt,
                    double* __restrict__ partials) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if (smem_u32(smem_raw) & 1023u) __trap();
  const int KC = d / 32;   // GEMM1 chunks per tile
  const int FB = d / 128;  // GEMM2 feature blocks per tile
  uint8_t* xs = smem_raw;                                  // LT_XS x 16 KB   fp32 staging
  uint8_t* wst = xs + LT_XS * CHUNK_BYTES;                 // LT_WS x 8 KB    W_hi | W_lo chunk
  uint8_t* a2 = wst + LT_WS * LT_WSTAGE;                   // 64 KB           x1 (2 atoms) | x2 (2 atoms), MN-major
  uint8_t* py = a2 + 4 * CHUNK_BYTES;                      // 2 x 16 KB       p1 (2 atoms of 4 KB) | p2
  SmemL* sm = reinterpret_cast<SmemL*>(py + 2 * 16384);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntiles = (n + TM - 1) / TM;
  // tiles of this CTA: blockIdx.x, + gridDim.x, ...
  const int my_tiles = (int)((ntiles > blockIdx.x) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0);

  if (threadIdx.x == 0) {
    for (int i = 0; i < LT_XS; ++i) {
      mbar_init(&sm->xs_full[i], 1);
      mbar_init(&sm->xs_empty[i], EPI_THREADS);
    }
    for (int i = 0; i < LT_WS; ++i) {
      mbar_init(&sm->w_full[i], 1);
      mbar_init(&sm->w_empty[i], 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&sm->a1_full[i], EPI_THREADS);
      mbar_init(&sm->a1_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm->d1_full[i], 1);
      mbar_init(&sm->d1_empty[i], EPI_THREADS);
      mbar_init(&sm->py_full[i], EPI_THREADS);
      mbar_init(&sm->py_empty[i], 1);
    }
    mbar_init(&sm->a2_full, LT_SPLIT_THREADS);
    mbar_init(&sm->a2_empty, 1);
    mbar_init(&sm->g_done, 1);
    mbar_init(&sm->g_part, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < LT_NP) sm->icpt[threadIdx.x] = (threadIdx.x < K) ? (float)icpt[threadIdx.x] : 0.f;
  if (warp == 3) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm->tmem_base)),
                 "r"(LT_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  // Work order shared by the producer, the split warps and the MMA issuer (lt_phase):
  //   G1(0), then for t = 0 .. my_tiles-1:  G1(t+1) (if any) and G2(t)
  // GEMM1 of the next tile is issued before GEMM2 of this one, so two tiles per SM are alive in L2
  if (warp == 0) {
    // ===================== X producer =====================
    if (lane == 0) {
      uint32_t g = 0;  // running chunk counter -> staging slot g % LT_XS
      // GEMM1 reads a tile from HBM and asks L2 to keep it (evict-last); GEMM2 re-reads it a few microseconds
      // later and releases it (evict-first), so the second pass is served by L2 instead of DRAM
      auto load_chunk = [&](int64_t tile, int col0, uint64_t policy) {
        const uint32_t slot = g % LT_XS, ph = (g / LT_XS) & 1;
        mbar_wait(&sm->xs_empty[slot], ph ^ 1);
        mbar_expect_tx(&sm->xs_full[slot], (uint32_t)CHUNK_BYTES);
        tma_load_2d_hint(xs + (size_t)slot * CHUNK_BYTES, &map_x, col0, (int)(tile * TM), &sm->xs_full[slot], policy);
        ++g;
      };
      for (int t = -1; t < my_tiles; ++t) {
        const int64_t tile1 = blockIdx.x + (int64_t)(t + 1) * gridDim.x, tile2 = blockIdx.x + (int64_t)t * gridDim.x;
        lt_phase(
            KC, FB, t + 1 < my_tiles, t >= 0,
            [&](int c) {
              if (c == 0) LTR(4, t + 1, 0);
              load_chunk(tile1, c * 32, L2_EVICT_LAST);
              if (c == KC - 1) LTR(4, t + 1, 1);
            },
            [&](int fb) {
              if (fb == 0) LTR(4, t, 2);
              for (int kc = 0; kc < 4; ++kc) load_chunk(tile2, (fb * 4 + kc) * 32, L2_EVICT_FIRST);
              if (fb == FB - 1) LTR(4, t, 3);
            });
      }
    }
  } else if (warp == 2) {
    // ===================== W producer (GEMM1 chunks only) =====================
    if (lane == 0) {
      uint32_t g = 0;
      for (int t = 0; t < my_tiles; ++t) {
        for (int c = 0; c < KC; ++c, ++g) {
          const uint32_t slot = g % LT_WS, ph = (g / LT_WS) & 1;
          mbar_wait(&sm->w_empty[slot], ph ^ 1);
          mbar_expect_tx(&sm->w_full[slot], (uint32_t)LT_WSTAGE);
          uint8_t* dst = wst + (size_t)slot * LT_WSTAGE;
          tma_load_2d(dst, &map_wh, c * 32, 0, &sm->w_full[slot]);
          tma_load_2d(dst + LT_NP * 128, &map_wl, c * 32, 0, &sm->w_full[slot]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint64_t kdesc = (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);  // K-major, SW128, SBO 1024
    // MN-major, SW128: LBO = 16 KB between 64-feature atoms, SBO = 1 KB between 8-row groups
    const uint64_t mndesc = ((uint64_t)(CHUNK_BYTES >> 4) << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
    const uint32_t w_a = smem_u32(wst) >> 4, a2_a = smem_u32(a2) >> 4, py_a = smem_u32(py) >> 4;
    uint32_t g1 = 0;  // GEMM1 chunk counter (operand ring slot g1 % 4, W stage g1 % LT_WS)
    uint32_t g2 = 0;  // GEMM2 block counter (a2 phase)
    auto gemm1 = [&](int t, int c) {  // one 32-feature chunk of the margins of tile t
      const uint32_t buf = t & 1;
      const uint32_t dd = tmem_base + LT_D1 + buf * LT_NP;
      if (c == 0) {
        LTR(0, t, 0);
        mbar_wait(&sm->d1_empty[buf], ((t >> 1) & 1) ^ 1);
        LTR(0, t, 1);
      }
      const uint32_t slot = g1 % LT_A1_SLOTS, ph = (g1 / LT_A1_SLOTS) & 1;
      const uint32_t ws = g1 % LT_WS, wph = (g1 / LT_WS) & 1;
      mbar_wait(&sm->a1_full[slot], ph);
      mbar_wait(&sm->w_full[ws], wph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t ah = tmem_base + LT_A1 + slot * 64, al = ah + 32;
        const uint64_t bh = kdesc | (uint64_t)(w_a + ws * (LT_WSTAGE >> 4));
        const uint64_t bl = bh + ((LT_NP * 128) >> 4);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          mma_tf32_ts(dd, ah + 8 * s, bh + 2 * s, (c | s) ? 1u : 0u);
          mma_tf32_ts(dd, ah + 8 * s, bl + 2 * s, 1u);
          mma_tf32_ts(dd, al + 8 * s, bh + 2 * s, 1u);
        }
        tc_commit(&sm->a1_empty[slot]);
        tc_commit(&sm->w_empty[ws]);
        if (c == KC - 1) tc_commit(&sm->d1_full[buf]);
      }
      __syncwarp();
      ++g1;
      if (c == KC - 1) LTR(0, t, 2);
    };
    auto gemm2 = [&](int t, int fb) {  // one 128-feature block of the gradient of tile t
      const uint32_t buf = t & 1;
      if (fb == 0) {
        LTR(0, t, 3);
        mbar_wait(&sm->py_full[buf], (t >> 1) & 1);
        LTR(0, t, 4);
      }
      mbar_wait(&sm->a2_full, g2 & 1);
      LTR(0, t, 5 + fb);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t dg = tmem_base + LT_G + fb * LT_GW;
        const uint64_t ax1 = mndesc | (uint64_t)a2_a;                             // x1: atoms 0, 1
        const uint64_t ax2 = mndesc | (uint64_t)(a2_a + 2 * (CHUNK_BYTES >> 4));  // x2: atoms 2, 3
        const uint64_t bp1 = kdesc | (uint64_t)(py_a + buf * (16384 >> 4));
        const uint64_t bp2 = bp1 + (LT_PY_P2 >> 4);
#pragma unroll
        for (int s = 0; s < 8; ++s) {  // 16 rows of K per instruction
          const uint32_t aoff = (uint32_t)s * (2048 >> 4);                            // 16 rows x 128 B
          const uint32_t boff = (uint32_t)(s >> 2) * (LT_PY_ATOM >> 4) + (s & 3) * 2;  // atom of 64 rows, 32 B steps
          if (LT_PCAT) {
            // x1 . [p1 | p2] and x2 . [p1 | p2]: all four products of the two-piece split.  (Until round 2 the
            // second one was x2 . p1 only: the dropped x2 . p2 term, 2^-18 |x||p| per product, and a TRUNCATED
            // x2 put the gradient at ~2e-5 of its largest entry; the N = 64 form costs tensor time only.)
            mma_bf16_ss(dg, ax1 + aoff, bp1 + boff, IDESC_G2_N64, ((t % LT_FLUSH) | s) ? 1u : 0u);
            mma_bf16_ss(dg, ax2 + aoff, bp1 + boff, IDESC_G2_N64, 1u);
          } else {
            mma_bf16_ss(dg, ax1 + aoff, bp1 + boff, IDESC_G2, ((t % LT_FLUSH) | s) ? 1u : 0u);
            mma_bf16_ss(dg, ax1 + aoff, bp2 + boff, IDESC_G2, 1u);
            mma_bf16_ss(dg, ax2 + aoff, bp1 + boff, IDESC_G2, 1u);
          }
        }
        tc_commit(&sm->a2_empty);
        if (fb == FB - 1) tc_commit(&sm->py_empty[buf]);
        // end of a flush period: the softmax warps move G^T out before this warp restarts the accumulators (they do
        // so ahead of arriving on py_full of the next tile, which the first GEMM2 block of that tile waits for)
        if (fb == FB - 1 && (t + 1) % LT_FLUSH == 0 && t + 1 < my_tiles) tc_commit(&sm->g_part);
      }
      __syncwarp();
      ++g2;
      if (fb == FB - 1) LTR(0, t, 9);
    };
    for (int t = -1; t < my_tiles; ++t)
      lt_phase(KC, FB, t + 1 < my_tiles, t >= 0, [&](int c) { gemm1(t + 1, c); }, [&](int fb) { gemm2(t, fb); });
    if (elect_one()) tc_commit(&sm->g_done);
    __syncwarp();
  } else if (warp >= 4 && warp < 8) {
    // ===================== softmax epilogue (thread == row) =====================
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    double loss = 0.0, sw = 0.0;
    float gb[LT_NP];
#pragma unroll
    for (int k = 0; k < LT_NP; ++k) gb[k] = 0.f;
    const int plen = K * d + K + 2;
    double* part = partials + (size_t)blockIdx.x * plen;
    // this thread owns gradient entries (class k, feature fb * 128 + r) of the CTA's fp64 partial: zero them, add
    // G^T in at every flush and once more at the end
    for (int fb = 0; fb < FB; ++fb)
      for (int k = 0; k < K; ++k) part[(size_t)k * d + fb * 128 + r] = 0.0;
    auto flush_g = [&]() {
      for (int fb = 0; fb < FB; ++fb) {
        uint32_t v[32];
        tmem_ld32(tmem_base + lane_addr + LT_G + fb * LT_GW, v);
        tmem_ld_wait();
        const int f = fb * 128 + r;  // lane == feature
        if (LT_PCAT) {
          uint32_t v2[32];
          tmem_ld32(tmem_base + lane_addr + LT_G + fb * LT_GW + 32, v2);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < LT_NP; ++k) v[k] = __float_as_uint(__uint_as_float(v[k]) + __uint_as_float(v2[k]));
        }
        // fire-and-forget fp64 reductions (RED.ADD.F64): a read-add-write would put 20 dependent L2 round trips per
        // block into every flush (measured: +4 ms per pass); nobody else touches these addresses
#pragma unroll
        for (int k = 0; k < LT_NP; ++k)
          if (k < K) atomicAdd(part + (size_t)k * d + f, (double)__uint_as_float(v[k]));
      }
    };
    for (int t = 0; t < my_tiles; ++t) {
      const int buf = t & 1;
      if (t > 0 && t % LT_FLUSH == 0) {
        mbar_wait(&sm->g_part, (uint32_t)(t / LT_FLUSH - 1) & 1u);
        tc_fence_after();
        flush_g();
        tc_fence_before();
      }
      const int64_t row = (blockIdx.x + (int64_t)t * gridDim.x) * TM + r;
      const bool live = row < n;
      const float wi = live ? (w ? w[row] : 1.f) : 0.f;
      const int yi = live ? (int)y[row] : 0;
      if (warp == 4) LTR(1, t, 0);
      mbar_wait(&sm->d1_full[buf], (t >> 1) & 1);
      if (warp == 4) LTR(1, t, 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + lane_addr + LT_D1 + buf * LT_NP, v);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&sm->d1_empty[buf]);
      if (warp == 4) LTR(1, t, 2);
      float m[LT_NP];
      float mx = 0.f;  // "subtract max if > 0"
#pragma unroll
      for (int k = 0; k < LT_NP; ++k) {
        m[k] = __uint_as_float(v[k]) + sm->icpt[k];
        if (k < K) mx = fmaxf(mx, m[k]);
      }
      float s = 0.f, my = 0.f;
#pragma unroll
      for (int k = 0; k < LT_NP; ++k) {
        if (k == yi) my = m[k];
        m[k] = (k < K) ? expf(m[k] - mx) : 0.f;
        s += m[k];
      }
      const float inv = 1.f / s;
      if (live) {
        loss += (double)(wi * (logf(s) + mx - my));
        sw += (double)wi;
      }
      // (P - Y) * w as bf16 pairs p1 + p2 into the GEMM2 B operand: [32 classes x 128 rows], K-major,
      // 2 atoms of 64 rows (128 B per class row), 128B swizzle
      if (warp == 4) LTR(1, t, 3);
      mbar_wait(&sm->py_empty[buf], ((t >> 1) & 1) ^ 1);
      if (warp == 4) LTR(1, t, 4);
      uint8_t* pbase = py + (size_t)buf * 16384 + (size_t)(r >> 6) * LT_PY_ATOM;
      const int cb = (r & 63) * 2;  // byte column of this row inside the 128-byte class row
#pragma unroll
      for (int k = 0; k < LT_NP; ++k) {
        const float pk = wi * (m[k] * inv - ((k == yi) ? 1.f : 0.f));
        const float val = (k < K) ? pk : 0.f;
        gb[k] += val;
        const __nv_bfloat16 h1 = __float2bfloat16_rn(val);
        const __nv_bfloat16 h2 = __float2bfloat16_rn(val - __bfloat162float(h1));
        const int off = k * 128 + ((((cb >> 4) ^ (k & 7)) << 4) | (cb & 15));
        *reinterpret_cast<__nv_bfloat16*>(pbase + off) = h1;
        *reinterpret_cast<__nv_bfloat16*>(pbase + LT_PY_P2 + off) = h2;
      }
      fence_async_smem();
      mbar_arrive(&sm->py_full[buf]);
      if (warp == 4) LTR(1, t, 5);
    }
    // ---- final: G^T from tensor memory, intercept gradient, loss (the row sums first: their registers are dead by the read-out)
#pragma unroll
    for (int k = 0; k < LT_NP; ++k) {
      const double t = warp_sum((double)gb[k]);
      if (lane == 0) sm->red[q][k] = t;
    }
    loss = warp_sum(loss);
    sw = warp_sum(sw);
    if (lane == 0) {
      sm->red[q][LT_NP] = loss;
      sm->red[q][LT_NP + 1] = sw;
    }
    mbar_wait(&sm->g_done, 0);
    tc_fence_after();
    if (my_tiles > 0) flush_g();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (r < K + 2) {
      const int src = (r < K) ? r : LT_NP + (r - K);
      part[(size_t)K * d + r] = sm->red[0][src] + sm->red[1][src] + sm->red[2][src] + sm->red[3][src];
    }
  } else if (warp >= 8) {
    // ===================== split warps: two groups of 4 warps, thread == row, alternate chunks =====================
    // Each chunk is converted by ONE group (all 32 columns per thread), so two chunks are in flight at once and
    // the per-chunk latency chain (wait, LDS, convert, tcgen05.st / STS, fence, arrive) overlaps across groups.
    const int w8 = warp - 8;
    const int q = w8 & 3, grp = w8 >> 2;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    uint32_t g = 0;   // staging chunk counter (same order as the producer); this group takes g % 2 == grp
    uint32_t g1 = 0;  // GEMM1 operand ring counter
    uint32_t g2 = 0;  // GEMM2 block counter
#ifdef LT_TRACE
    long long ltw[4] = {0, 0, 0, 0};  // cycles warp 8 waited in a phase: xs_full (G1) | a1_empty | a2_empty | xs_full (G2)
#endif
    auto split1 = [&](int tt, int c) {  // one GEMM1 chunk of tile tt -> tf32 hi / lo into tensor memory
      if (q == 0 && c == 0) LTR(2 + grp, tt, 0);
      if ((g & 1u) == (uint32_t)grp) {
        const uint32_t slot = g % LT_XS, ph = (g / LT_XS) & 1;
        if (w8 == 0) LTR(5, tt, 0);
        LTW(0, mbar_wait(&sm->xs_full[slot], ph));
        if (w8 == 0) LTR(5, tt, 1);
        const uint8_t* src = xs + (size_t)slot * CHUNK_BYTES + (size_t)r * 128;
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          const float4 v = *reinterpret_cast<const float4*>(src + ((p ^ (r & 7)) << 4));
          // hi = x rounded to TF32 (nearest, ties away) with two full-rate integer ops instead of cvt.rna;
          // lo = x - hi is exact in fp32 and is left unrounded: whatever the tensor core drops of it is
          // below 2^-21 |x|
          const uint32_t h0 = (__float_as_uint(v.x) + 0x1000u) & 0xFFFFE000u, h1 = (__float_as_uint(v.y) + 0x1000u) & 0xFFFFE000u;
          const uint32_t h2 = (__float_as_uint(v.z) + 0x1000u) & 0xFFFFE000u, h3 = (__float_as_uint(v.w) + 0x1000u) & 0xFFFFE000u;
          hi[p * 4 + 0] = h0; hi[p * 4 + 1] = h1; hi[p * 4 + 2] = h2; hi[p * 4 + 3] = h3;
          lo[p * 4 + 0] = __float_as_uint(v.x - __uint_as_float(h0)); lo[p * 4 + 1] = __float_as_uint(v.y - __uint_as_float(h1));
          lo[p * 4 + 2] = __float_as_uint(v.z - __uint_as_float(h2)); lo[p * 4 + 3] = __float_as_uint(v.w - __uint_as_float(h3));
        }
        // The staging slot is about to be refilled by TMA (async proxy) and was just read through the generic
        // proxy: the proxy fence orders those reads before the refill.  Without it the refill could overtake reads
        // still in flight -- ptxas is free to sink the arithmetic that consumes them below the arrive, which only
        // clobbers memory -- and at d = 128 (four chunks per tile, six slots: the producer is always waiting at the
        // slot) it did: gradients differed by 3e-3 from run to run as soon as a CTA had three tiles.
        fence_async_smem();
        mbar_arrive(&sm->xs_empty[slot]);
        if (w8 == 0) LTR(5, tt, 2);
        const uint32_t aslot = g1 % LT_A1_SLOTS, aph = (g1 / LT_A1_SLOTS) & 1;
        LTW(1, mbar_wait(&sm->a1_empty[aslot], aph ^ 1));
        if (w8 == 0) LTR(5, tt, 3);
        tc_fence_after();
        tmem_st32(tmem_base + lane_addr + LT_A1 + aslot * 64, hi);
        tmem_st32(tmem_base + lane_addr + LT_A1 + aslot * 64 + 32, lo);
        tmem_st_wait();
        if (w8 == 0) LTR(5, tt, 4);
        tc_fence_before();
        mbar_arrive(&sm->a1_full[aslot]);
        if (w8 == 0) LTR(5, tt, 5);
      }
      ++g;
      ++g1;
      if (q == 0 && c == KC - 1) LTR(2 + grp, tt, 1);
    };
    auto split2 = [&](int tt, int fb) {  // one GEMM2 block of tile tt: 4 chunks -> bf16 x1 / x2, MN-major smem tile
      if (q == 0 && fb == 0) LTR(2 + grp, tt, 2);
      LTW(2, mbar_wait(&sm->a2_empty, (g2 & 1) ^ 1));
      if (q == 0) LTR(2 + grp, tt, 3 + fb);
      for (int kc = 0; kc < 4; ++kc, ++g) {
        if ((g & 1u) != (uint32_t)grp) continue;
        const uint32_t slot = g % LT_XS, ph = (g / LT_XS) & 1;
        if (w8 == 0) LTR(5, tt, 6);
        LTW(3, mbar_wait(&sm->xs_full[slot], ph));
        if (w8 == 0) LTR(5, tt, 7);
        const uint8_t* src = xs + (size_t)slot * CHUNK_BYTES + (size_t)r * 128;
        // features kc*32 .. +31 of this block -> atom kc/2, 16-byte chunks (kc%2)*4 + {0..3} of the row
        uint8_t* d1 = a2 + (size_t)(kc >> 1) * CHUNK_BYTES + (size_t)r * 128;
        uint8_t* d2 = d1 + 2 * CHUNK_BYTES;
#pragma unroll
        for (int hh = 0; hh < 4; ++hh) {
          const float4 va = *reinterpret_cast<const float4*>(src + (((2 * hh) ^ (r & 7)) << 4));
          const float4 vb = *reinterpret_cast<const float4*>(src + (((2 * hh + 1) ^ (r & 7)) << 4));
          const float f[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
          uint32_t o1[4], o2[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            // x1 = x rounded to bf16 (nearest, ties away) and x2 = (x - x1) rounded the same way (what is left of
            // x is then below 2^-18 |x|; truncating x2 left 2^-16), with full-rate integer ops; __byte_perm packs
            // the two high halves (even k in the low 16 bits)
            const uint32_t ra = (__float_as_uint(f[2 * e]) + 0x8000u) & 0xFFFF0000u;
            const uint32_t rb = (__float_as_uint(f[2 * e + 1]) + 0x8000u) & 0xFFFF0000u;
            const uint32_t la = __float_as_uint(f[2 * e] - __uint_as_float(ra)) + 0x8000u;
            const uint32_t lb = __float_as_uint(f[2 * e + 1] - __uint_as_float(rb)) + 0x8000u;
            o1[e] = __byte_perm(ra, rb, 0x7632);
            o2[e] = __byte_perm(la, lb, 0x7632);
          }
          const int off = (((kc & 1) * 4 + hh) ^ (r & 7)) << 4;
          *reinterpret_cast<uint4*>(d1 + off) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
          *reinterpret_cast<uint4*>(d2 + off) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
        }
        fence_async_smem();  // generic reads of the slot before its refill through the async proxy (see split1)
        mbar_arrive(&sm->xs_empty[slot]);
        if (w8 == 0) LTR(5, tt, 8);
      }
      fence_async_smem();
      mbar_arrive(&sm->a2_full);
      ++g2;
      if (q == 0) LTR(2 + grp, tt, 7 + fb);
    };
    for (int t = -1; t < my_tiles; ++t) {
      lt_phase(KC, FB, t + 1 < my_tiles, t >= 0, [&](int c) { split1(t + 1, c); }, [&](int fb) { split2(t, fb); });
      LTW_FLUSH(t);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(LT_TMEM_COLS)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// W' = beta (fp64, [K, d]) as explicit TF32 hi / lo, padded to 32 class rows
__global__ void lr_mn_prep_kernel(const double* __restrict__ beta, int K, int d, float* __restrict__ wh,
                                  float* __restrict__ wl) {
  const int total = LT_NP * d;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i / d;
    const double v = (k < K) ? beta[i] : 0.0;
    const float hi = to_tf32((float)v);
    wh[i] = hi;
    wl[i] = to_tf32((float)(v - (double)hi));
  }
}

// (d = 128 was off this kernel for part of round 2: see the proxy fence in the split warps.)
#ifndef LT_MIN_D
#define LT_MIN_D 128
#endif
bool lr_mn_tc_supported(int d, int64_t ld, int K) { return d % 128 == 0 && d >= LT_MIN_D && d <= 512 && ld == d && K >= 2 && K <= LT_NP; }

size_t lr_mn_tc_extra_scratch(int d) { return (size_t)2 * LT_NP * d * sizeof(float) + 2048; }

// scratch layout: [ per-CTA fp64 partials (grid x plen) | pad to 1 KB | W_hi | W_lo ]
int lr_mn_tc_launch(const float* X, const float* y, const float* w, int64_t n, int d, int K, const double* beta,
                    const double* icpt, double* partials, int grid, void* wbuf, cudaStream_t st) {
  float* wh = static_cast<float*>(wbuf);
  float* wl = wh + (size_t)LT_NP * d;
  lr_mn_prep_kernel<<<(LT_NP * d + 255) / 256, 256, 0, st>>>(beta, K, d, wh, wl);
  B2F_LAUNCH_CHECK();
  CUtensorMap mx, mh, ml;
  int rc;
  if ((rc = make_map_f32(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)d)) != B2F_OK) return rc;
  if ((rc = make_map_2d(&mh, wh, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, LT_NP, (uint64_t)d, (uint64_t)d, LT_NP)) != B2F_OK)
    return rc;
  if ((rc = make_map_2d(&ml, wl, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, LT_NP, (uint64_t)d, (uint64_t)d, LT_NP)) != B2F_OK)
    return rc;
  const size_t smem = (size_t)LT_XS * CHUNK_BYTES + (size_t)LT_WS * LT_WSTAGE + 4 * CHUNK_BYTES + 2 * 16384 +
                      sizeof(SmemL) + 64;
  B2F_CUDA(b2f_smem_attr(lr_mn_tc_kernel, smem));
  lr_mn_tc_kernel<<<grid, LT_THREADS, smem, st>>>(mx, mh, ml, y, w, n, d, K, icpt, partials);
  B2F_LAUNCH_CHECK();
#ifdef LT_TRACE
  if (const char* path = getenv("B2F_LR_TRACE")) {
    std::vector<unsigned long long> host(LTR_ROLES * LTR_TILES * LTR_SLOTS);
    B2F_CUDA(cudaStreamSynchronize(st));
    B2F_CUDA(cudaMemcpyFromSymbol(host.data(), lt_trace, host.size() * sizeof(unsigned long long)));
    if (FILE* f = fopen(path, "wb")) {
      fwrite(host.data(), sizeof(unsigned long long), host.size(), f);
      fclose(f);
    }
  }
#endif
  return B2F_OK;
}
