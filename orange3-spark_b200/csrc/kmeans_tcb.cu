// K2 fused Lloyd kernel: low-precision filter pass + exact re-check (sm_100a only).
//
// The 3xTF32 kernel (kmeans_tc.cu) is limited by the tensor pipe under the board power cap.  Here the first pass
// runs on cheap operands and only has to decide which rows it can label FOR CERTAIN:
//   fp16 filter (B2F_KM_TCH, default):  score ~= (S x) . (S c') / S^2 + ||c||^2,  c' = -2c, both operands rounded
//       once to fp16 after scaling by a power of two S taken from the centres: ONE kind::f16 MMA per product.
//       |error| <= tau = kappa ||x|| max||c|| + underflow term,  kappa = 4 * 2^-11 (+ accumulation slack).
//   bf16x3 filter (B2F_KM_TCB):  x = x1 + x2, c' = c1 + c2 (bf16), score ~= x1.c1 + x1.c2 + x2.c1 + ||c||^2: three
//       MMAs, kappa = 6 * 2^-16.
// A row whose best and second-best scores are more than 2 tau apart has a CERTAIN label and is accumulated here;
// every other row goes to a list, is gathered into a compact matrix and re-assigned by the 3xTF32 kernel (which has
// its own certainty test) and, failing that, by the fp64 kernel.  Labels, sums and counts therefore equal the exact
// path's.  (fp16 filter: the epilogue keeps class minima instead of a running top two (TB_GRID below); an uncertain
// row with at most four candidate centres within 2 tau of the best goes to a candidate list and is decided by
// fp32 / fp64 distances to those (km_pair_exact_kernel) instead of a pass over all k centres.)  Cost: bf16x3 adds its (zero-mean) score error per certain row; the fp16 variant adds sum ||x||^2 only and
// the centre-dependent remainder sum_j (n_j ||c_j||^2 - 2 c_j . S_j) exactly from the sums (tcb_cost_terms_kernel).
//
// Replaces MLlib's findClosest + axpy loop behind KMeans.fit;
//   reference call site: orangecontrib/spark/base/spark_ml_estimator.py:22.
//
// CTA = 640 threads, persistent, one per SM.  The scheduler prefers the higher warp ids, so latency-critical roles
// sit at the top:
//   warps 0-7   epilogue: two per TMEM lane quarter; bf16x3: each takes every other centre tile whole, running best /
//               second best; fp16: each takes every centre tile of every other ROW tile (TB_SKEW below), integer
//               class minima with 3-input min only; results to shared memory
//   warps 8-11  split: fp32 row -> fp16 (or bf16 pair) operand -> tcgen05.st into tensor memory; row norms
//   warps 12-15 finish (thread == row): (merge the two epilogue partials,) certainty test, labels, candidate and
//               re-check lists, accumulation of certain rows on the copy engine (bulk / tensor reductions into 4 fp32
//               replicas), cost, counts
//   warp 16     X producer (TMA, fp32 staging chunks of 128 rows x 32 floats)
//   warps 17,19 MMA issuers, alternate centre tiles in token order (one batch of 4*KA + 1 or 12*KA tcgen05.mma per
//               centre tile, A operand from tensor memory); warp 19 also allocates the tensor memory
//   warp 18     B producer (TMA, stage = one centre tile: 128 centres x d of the 16-bit operand plane(s))
// Tensor memory: [0,256) two accumulators | [256,384) operand x1, two row tiles | [384,512) bf16x3: x2; fp16: a
// third accumulator.  Registers are moved from the control / finish warpgroups to the epilogue (setmaxnreg).
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <vector>

#include "tc_common.cuh"

namespace {

constexpr uint32_t TB_TMEM_COLS = 512;
// bf16x3: [0,256) two accumulators | [256,384) x1, two row tiles of 64 columns | [384,512) x2, likewise
// fp16:   [0,384) three accumulators | [384,512) the operand, two row tiles of 64 columns  (the constant k-block that
//         multiplies the ||c||^2 pieces of the centre operand is read from shared memory)
constexpr uint32_t TB_A1_BF = 256, TB_A2_BF = 384, TB_A1_H = 384, TB_ABUF = 64;
__device__ __forceinline__ uint32_t tb_acc_col(uint32_t buf) { return buf * TN; }

// fp16 filter epilogue works on INTEGER KEYS.  Per row, scores are scaled by a power of two `sc` such that
// |score| * sc < 2^21 (|score| <= ||c||^2 + 2 ||x|| ||c||), and formed as  t = D * (inv * sc) + 1.5 * 2^23  (D holds
// S^2 * score: the ||c||^2 term is added by the tensor core through one more k-block of the operands):
// the magic constant leaves round(score * sc) in the mantissa, so  key = (int)bits(t) * 129 + column  (wrapping
// multiply; the window of possible products stays inside the int32 range) orders like (score, column) under SIGNED
// compare, and best / second best need min/max only -- no compare + select for the index: 2.5 ALU-pipe operations
// per score instead of 4.5; the two multiply-adds, the integer one included (hence 129, not a shift), run on the FMA
// pipe, and no ||c||^2 table is read.
// Each key is within 1 unit of score * sc; the certainty test pays for it with 2.5 units.
constexpr float TB_MAGIC = 12582912.f;  // 1.5 * 2^23
constexpr int TB_KEYMUL = 129;
__device__ __forceinline__ int tb_key(float d, float arow, int col) {
  return __float_as_int(fmaf(d, arow, TB_MAGIC)) * TB_KEYMUL + col;
}
// value part of a key, relative to the key of score 0 (floor division of a possibly negative number)
__device__ __forceinline__ int tb_key_value(int key) {
  const int base = (int)(0x4B400000u * (uint32_t)TB_KEYMUL);  // wraps to a negative number
  const int rel = key - base + (1 << 22) * TB_KEYMUL;         // >= 0: values span (-2^22, 2^22)
  return rel / TB_KEYMUL;
}
__device__ __forceinline__ int tb_key_col(int key) {
  const int base = (int)(0x4B400000u * (uint32_t)TB_KEYMUL);
  const int rel = key - base + (1 << 22) * TB_KEYMUL;
  return rel % TB_KEYMUL;
}
__device__ __forceinline__ float tb_row_scale(float xn, float cmax) {
  const float R = fmaf(2.f * sqrtf(xn), cmax, cmax * cmax);  // +inf / NaN for rows flagged by the split warps
  int e = (__float_as_int(R) >> 23) & 0xff;                  // R < 2^(e - 126)
  e = 274 - e;                                               // biased exponent of 2^(21 - (e - 126))
  e = max(1, min(254, e));
  return __int_as_float(e << 23);
}

// ---- "grid minima" epilogue of the fp16 filter (TB_GRID) -------------------------------------------------------
// The ALU pipe (min / max, 16 lanes per clock) is what bounds the filter kernel, and a running (best, second best)
// over the scores costs 2.5 of its operations per score plus a multiply-add for the key.  Instead every score falls
// into two classes whose minima are all that is kept:
//   M-class  = the 16 consecutive centres it belongs to (a half of a 32-column accumulator chunk)
//   A-class  = its position in them, column mod 16 (16 running minima per row, over all centre tiles)
// one 3-input min per two scores for each (0.5 + 0.5 operations per score, no key arithmetic per score).  Two
// different centres never share both classes, so a centre that is hidden behind a smaller one in one class is the
// minimum of its class in the other: with the three smallest class minima of either kind,
//   second best score  = min(second smallest M-minimum, second smallest A-minimum),
//   every centre within thr of the best lies in (M-classes with minimum within thr) x (A-classes likewise),
// which certifies a row (one candidate) or hands at most four candidate centres to the fp64 check (third smallest
// minimum of either kind out of reach).  Class ids ride in the low bits of the class minima only (16 + 8 multiply-adds
// per centre tile and row instead of 128); keys are relative to TB_QBASE so that value = key >> bits, no wrap.
#ifndef TB_GRID
#define TB_GRID 1
#endif
// ---- skewed row-tile pairs (TB_SKEW, fp16 filter with the class-minima epilogue) ---------------------------------
// Every CTA streams ALL centre tiles from L2 for each of its row tiles: 45 of the 62 GB of L2 -> SM reads per 20M x 128,
// k = 1024 launch, at 80 % of what this L2 delivers.  Here a centre-tile stage serves TWO row tiles: the CTA's even row
// tiles (slot A: MMA issuer 0, epilogue warps 0-3, operand buffer 0) and its odd ones (slot B: issuer 1, epilogue warps
// 4-7, buffer 1) run NT/2 centre tiles out of phase, slot A taking the centre tiles in the order 0 .. NT-1 and slot B
// in the order NT/2 .. NT-1, 0 .. NT/2-1 -- so at every STEP both work on the same centre tile, while their row tiles
// begin and end half a period apart and the operand of one is rebuilt under the MMAs of the other.  Step s loads
// centre tile s % NT once; batches are issued A(s), B(s) in token order; accumulators are handed out along that
// global order.  An epilogue warp now sees every centre tile of its row tile (no merge of two halves afterwards).
// OFF by default: measured -7 % on the cfg-2 step (7.27 -> 6.76 ms per 20M rows) and green in two full runs of the GPU
// suites, but in the last validation run of round 2 one launch of the assignment-only pipeline (per-row key scales,
// 300 000 x 128, k = 1024: 16 row tiles per CTA) ran into the bounded barrier wait of the centre-tile producer
// (b_empty, stage 0).  Not reproduced and not explained with the GPU time that was left, so the validated schedule
// ships; -DTB_SKEW=1 builds this one.  First suspects: the stage is released by the LAST issuer's commit alone (it
// relies on the tensor pipe completing the two issuers' batches in issue order), and slot B's late-joining warps.
#ifndef TB_SKEW
#define TB_SKEW 0
#endif
constexpr int TB_QBASE = 0x4B000000;  // bit pattern of 2^23: quantised scores are TB_QBASE + [0, 2^23)
__device__ __forceinline__ int tb_min3(int a, int b, int c) { return min(min(a, b), c); }
// smallest of 16 consecutive entries of a chunk (8 operations)
__device__ __forceinline__ int tb_min16(const int* q) {
  const int a = tb_min3(q[0], q[1], q[2]), b = tb_min3(q[3], q[4], q[5]), c = tb_min3(q[6], q[7], q[8]);
  const int e = tb_min3(q[9], q[10], q[11]), f = tb_min3(q[12], q[13], q[14]);
  return min(tb_min3(a, b, c), tb_min3(e, f, q[15]));
}
// three smallest, sorted, of two sorted triples (7 operations)
__device__ __forceinline__ void tb_merge3(int& x1, int& x2, int& x3, int y1, int y2, int y3) {
  const int t3 = min(tb_min3(x3, y3, max(x2, y1)), max(x1, y2));
  const int t2 = tb_min3(max(x1, y1), x2, y2);
  x1 = min(x1, y1);
  x2 = t2;
  x3 = t3;
}
// three smallest, sorted, of eight values (25 operations)
__device__ __forceinline__ void tb_top3_of8(const int* x, int& t1, int& t2, int& t3) {
  int c[2][3];
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    const int a1 = min(x[4 * g], x[4 * g + 1]), a2 = max(x[4 * g], x[4 * g + 1]);
    const int b1 = min(x[4 * g + 2], x[4 * g + 3]), b2 = max(x[4 * g + 2], x[4 * g + 3]);
    const int t = max(a1, b1), u = min(a2, b2);
    c[g][0] = min(a1, b1);
    c[g][1] = min(t, u);
    c[g][2] = max(t, u);
  }
  t1 = c[0][0]; t2 = c[0][1]; t3 = c[0][2];
  tb_merge3(t1, t2, t3, c[1][0], c[1][1], c[1][2]);
}

constexpr int TB_THREADS = 640;
// Rows per centre are counted in shared memory (one int per centre behind SmemB, added to the global counters once at
// the end of the kernel) up to this many centres; beyond it the counters would cost a centre stage.
constexpr int TB_CNT_SMEM_MAX = 1024;
static inline int tcb_cnt_smem_bytes(int k) { return k <= TB_CNT_SMEM_MAX ? k * 4 : 0; }
// RFS ("reduce from staging", d = 64 only): the fp32 staging area holds TB_NX whole row tiles (32 KB each at d = 64)
// instead of one, a tile stays there until its labels exist, and the finish warps add each certain row straight from
// it with two TMA tensor reductions (one per 128-byte chunk; the engine undoes the 128B swizzle) into the replica
// row of its label: no second read of X from L2, no row slots, no copy batches.  At d = 64 a row tile has only two
// centre tiles and the accumulation is most of the step (pipeline without it: 10.9 ms per 100M rows, with it 24-25).
#ifndef TB_NX_V
#define TB_NX_V 4
#endif
constexpr int TB_NX = TB_NX_V;
// Tried on top of it and not kept: a per-CTA [k, d] fp32 table of the sums in shared memory for k * d <= 16K (finish
// warp w owns the centres with label % 4 == w and adds their rows with plain loads / stores, no atomics, nothing per
// row to L2): 11.3 ms per 40M rows at k = 256 against 7.9 ms with the tensor reductions -- the table walk (a ballot
// scan of the tile's 128 labels plus ~32 dependent load-add-store rounds per warp and tile) costs the finish warps
// more than handing the rows to the copy engine does.

// One 32-column chunk of the fp16 filter's integer-key epilogue: running best / second best of two chains.
// GS (global key scale, b2f_kmeans_prepare_x): the tensor core has already added the magic constant, the
// accumulator's bit pattern IS the quantised score and a key costs one integer multiply-add; otherwise the score is
// first scaled into its row's key window (tb_key: one more multiply-add).
template <bool GS>
__device__ __forceinline__ void tb_chunk_keys(const uint32_t (&v)[32], float arow, int col0, int& k1a, int& k2a,
                                              int& k1b, int& k2b);
template <bool GS>
__device__ __forceinline__ void tb_chunk_keys(const uint32_t (&v)[32], float arow, int col0, int& k1a, int& k2a,
                                              int& k1b, int& k2b) {
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const int col = col0 + j4 * 4;
    int q0, q1, q2, q3;
    if (GS) {
      q0 = (int)v[j4 * 4 + 0] * TB_KEYMUL + col;
      q1 = (int)v[j4 * 4 + 1] * TB_KEYMUL + col + 1;
      q2 = (int)v[j4 * 4 + 2] * TB_KEYMUL + col + 2;
      q3 = (int)v[j4 * 4 + 3] * TB_KEYMUL + col + 3;
    } else {
      q0 = tb_key(__uint_as_float(v[j4 * 4 + 0]), arow, col);
      q1 = tb_key(__uint_as_float(v[j4 * 4 + 1]), arow, col + 1);
      q2 = tb_key(__uint_as_float(v[j4 * 4 + 2]), arow, col + 2);
      q3 = tb_key(__uint_as_float(v[j4 * 4 + 3]), arow, col + 3);
    }
    const int loa = min(q0, q1), hia = max(q0, q1), lob = min(q2, q3), hib = max(q2, q3);
    k2a = min(k2a, min(hia, max(k1a, loa)));
    k1a = min(k1a, loa);
    k2b = min(k2b, min(hib, max(k1b, lob)));
    k1b = min(k1b, lob);
  }
}
constexpr int TB_CN_BYTES = 128 * 32;  // the ||c||^2 k-block of a centre tile: 128 centres x 16 halves
#ifndef TB_REG_CTRL
// what the epilogue warpgroups gain, (EPI - 96) * 256, must be given back first by the others:
// (96 - CTRL) * 128 + (96 - FIN) * 128  (the pool is per CTA)
// (the pool is what the CTA was LAUNCHED with -- 640 threads x 96 -- not the SM's register file)
// Measured with the class-minima epilogue (20M x 128, k = 1024, whole step): (CTRL, FIN, EPI) = (48, 80, 128) 7.9-8.2 ms;
// (64, 80, 120), (80, 80, 112), (88, 88, 104), (72, 88, 112) all 7.5: at 48 registers the producer and MMA-issuer
// loops reload spilled loop state from local memory every iteration, and that sits on the critical path; the
// epilogue no longer needs 128.
#define TB_REG_CTRL 80
#define TB_REG_FIN 80
#define TB_REG_EPI 112
#define TB_FIN_ROWS 8
#define TB_FX_ROWS 8
#endif
constexpr int TB_FX_WARPS = 4;    // warps that own row slots (the finish warps)
constexpr int TB_FX_SPAN = 32;    // rows of a tile each of them accumulates
// dev tracing (B2F_TC_TRACE=<file>): block 0 stamps clock64() at the hand-offs of its first TR_TILES row tiles
[[maybe_unused]] constexpr int TR_TILES = 64, TR_SLOTS = 32, TR_ROLES = 12;  // mma | 8 epilogue warps | split | finish | B producer
#define TR(role, slot)                                                                      \
  do {                                                                                      \
    if (TRACE && blockIdx.x == 0 && lane == 0 && it < (uint32_t)TR_TILES)                   \
      trace[((role) * TR_TILES + it) * TR_SLOTS + (slot)] = (unsigned long long)clock64();  \
  } while (0)
// stage = 128 centres x 64 k: c1 atom + c2 atom (bf16x3), or the single scaled fp16 atom (fp16 filter)
template <int MODE> __host__ __device__ constexpr int tb_stage_bytes() { return MODE == KM_FILTER_FP16 ? CHUNK_BYTES : 2 * CHUNK_BYTES; }
// kind::f16, fp32 accumulate, A and B K-major, M = 128, N = TN; input format 1 = bf16, 0 = fp16
constexpr uint32_t IDESC_BF16 =
    (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
constexpr uint32_t IDESC_FP16 = (1u << 4) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);

// warp roles.  The scheduler prefers the higher warp ids, so the warps whose latency everybody else waits on (MMA
// issuer, producers) get the top ids and the throughput-bound epilogue warps the bottom ones.
constexpr int W_EPI0 = 0;    // warps 0-7   epilogue
constexpr int W_SPLIT0 = 8;  // warps 8-11  split
constexpr int W_FIN0 = 12;   // warps 12-15 finish
constexpr int W_XPROD = 16, W_MMA = 17, W_BPROD = 18, W_ALLOC = 19;  // W_ALLOC: TMEM allocator + second MMA issuer

struct SmemB {
  uint64_t xs_full[8], xs_empty[4];  // fp32 staging, per 32-float chunk (RFS: xs_full[slot * KC + kc], xs_empty[slot])
  uint64_t xa_full[2], xa_empty[2];  // TMEM operand, per row-tile buffer
  uint64_t b_full[4], b_empty[4];    // centre-operand stages (a stage = one whole centre tile)
  uint64_t t_full[3], t_empty[3];    // accumulators
  uint64_t a_full[2], a_empty[2];    // epilogue -> finish hand-off
  uint64_t n_full[4], n_empty[4];    // split -> finish: squared row norms of a tile
  uint64_t tok[2];                   // MMA issue token: warp p has put its batch into the tensor queue
  uint64_t fx_full[4][2];            // finish warps (fp16 filter): rows of a batch have landed in their slot
  uint32_t tmem_base;
  float xnorm2[4][TM];
  float p_b1[2][2][TM], p_b2[2][2][TM];  // [slot][epilogue warp of the lane quarter][row]: best, second best
  int p_i1[2][2][TM];                    // index of the best
  int p_b3[2][2][TM];                    // fp16 filter: lower bound of the third best key; index of the second best
  int p_i2[2][2][TM];
  int p_a3[2][2][TM];                    // TB_GRID: third smallest A-class minimum (low 4 bits: class of the smallest)
  alignas(16) float cscale[TM];          // scaled pass: per-column factor (StandardScaler fold), d <= 128 entries
                                         // (read as float4: the 4-byte tmem_base above breaks the natural alignment)
};

// (d = 128, k = 1024, fp16 filter: three centre stages fit beside everything else only up to this size)
static_assert(sizeof(SmemB) <= 15296, "SmemB grew: the fp16 filter loses a centre stage at d = 128");

struct TcbWs {  // header of the caller-provided workspace (the first four ints are documented in b200fit.h)
  int count;    // rows appended to the re-check list (may exceed cap: the excess is dropped)
  int used;     // min(count, cap), written by the gather kernel
  int count2;   // rows the 3xTF32 re-check could not certify either -> exact fp64 kernel (same: excess dropped)
  int cap;      // capacity of each of the two lists above
  int pcount;   // rows appended to the pair list (those beyond its capacity are sent to the re-check list instead)
  int pad[59];
};

__device__ __forceinline__ void tc_mma_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

}  // namespace

// Accumulation of the 32 consecutive rows of a finish warp (first row r0; lane u holds the label of row r0 + u, or -1
// for a row that is not to be added) into the replica table, entirely on the copy engine: bulk copy L2 -> shared
// memory slot, then bulk reduce-add of each row into the replica row of its label.  No registers hold row data and no
// load latency sits in this warp; two batches of TB_FX_ROWS rows are in flight (two slots, `fx_full` their barriers,
// `fxc` the running batch count of this warp).  At 8 rows in flight through registers this loop set the pace of the
// whole kernel (hand-off trace).  Split in two so that the caller can put its list appends -- atomics whose return
// value costs an L2 round trip -- between issuing the first two copies and waiting for them.
struct TbAcc {
  static constexpr int NB = TB_FX_SPAN / TB_FX_ROWS;
  const float* X;
  int64_t n, ld, r0;
  int d, lane;
  float* myrep;
  uint8_t* mybuf;
  uint64_t* fx_full;
  int lbl_b[NB];

  __device__ __forceinline__ void issue_load(uint32_t fxc, int b, bool newest_may_pend) {
    const uint32_t rowbytes = (uint32_t)d * 4u;
    const uint32_t slot = (fxc + b) & 1u;
    const bool merged = ld == d;  // the batch's rows are one contiguous block: one copy instead of TB_FX_ROWS
    const int64_t rb = r0 + b * TB_FX_ROWS;
    const int nlive = (int)max((int64_t)0, min((int64_t)TB_FX_ROWS, n - rb));
    const bool act = lane < TB_FX_ROWS && lbl_b[b] >= 0;
    const unsigned m = __ballot_sync(0xffffffffu, act);
    // the reduction that read this slot two batches ago must be done with it
    if (newest_may_pend) bulk_wait_read<1>();
    else bulk_wait_read<0>();
    if (lane == 0) {
      const uint32_t bytes = merged ? (m ? (uint32_t)nlive * rowbytes : 0u) : (uint32_t)__popc(m) * rowbytes;
      if (bytes) mbar_expect_tx(&fx_full[slot], bytes);
      else mbar_arrive(&fx_full[slot]);
      if (merged && bytes)
        bulk_load_hint(mybuf + (size_t)slot * TB_FX_ROWS * rowbytes, X + rb * ld, bytes, &fx_full[slot], L2_EVICT_FIRST);
    }
    __syncwarp();
    if (act && !merged)
      bulk_load_hint(mybuf + (size_t)(slot * TB_FX_ROWS + lane) * rowbytes, X + (rb + lane) * ld, rowbytes,
                     &fx_full[slot], L2_EVICT_FIRST);
  }
  // labels -> per-batch labels, copies of the first two batches on their way
  __device__ __forceinline__ void begin(uint32_t fxc, int my_lab) {
    const int u = lane & (TB_FX_ROWS - 1);
#pragma unroll
    for (int b = 0; b < NB; ++b) lbl_b[b] = __shfl_sync(0xffffffffu, my_lab, b * TB_FX_ROWS + u);
    issue_load(fxc, 0, true);
    issue_load(fxc, 1, false);
  }
  __device__ __forceinline__ void finish(uint32_t& fxc) {
    const uint32_t rowbytes = (uint32_t)d * 4u;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const uint32_t slot = (fxc + b) & 1u;
      mbar_wait_ns(&fx_full[slot], ((fxc + b) >> 1) & 1u, 64);
      if (lane < TB_FX_ROWS && lbl_b[b] >= 0)
        bulk_reduce_add_f32(myrep + (size_t)lbl_b[b] * d, mybuf + (size_t)(slot * TB_FX_ROWS + lane) * rowbytes, rowbytes);
      bulk_commit();
      if (b + 2 < NB) issue_load(fxc, b + 2, false);
    }
    fxc += NB;
  }
};

template <int MODE, bool TRACE, bool SCALED, bool RFS>
__global__ void __launch_bounds__(TB_THREADS, 1)
    km_tcb_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c1,
                  const __grid_constant__ CUtensorMap map_c2, const __grid_constant__ CUtensorMap map_rep,
                  const float* __restrict__ X, int64_t n, int d,
                  int64_t ld, const float* __restrict__ colscale, const float* __restrict__ cn,
                  const float* __restrict__ misc, float kappa2, float kappa0, int k, int nstages,
                  double* __restrict__ payload, float* __restrict__ rep, unsigned long long* __restrict__ cnt,
                  int32_t* __restrict__ labels, int accumulate, TcbWs* __restrict__ ws,
                  int32_t* __restrict__ amb_idx, int cap, int4* __restrict__ pairs, int pcap,
                  unsigned long long* __restrict__ trace) {
  constexpr int TB_STAGE_BYTES = tb_stage_bytes<MODE>();  // per 64-k atom
  constexpr bool H16 = MODE == KM_FILTER_FP16;
  // a stage holds one centre tile: KA atoms (+ the ||c||^2 block)
  const int TB_TILE_BYTES = (d / 64) * TB_STAGE_BYTES + (H16 ? TB_CN_BYTES : 0);
  constexpr uint32_t XBUFS = 2;  // row-tile buffers of the TMEM operand
  constexpr uint32_t TB_A1 = H16 ? TB_A1_H : TB_A1_BF, TB_A2 = TB_A2_BF;
  constexpr uint32_t NBUF = H16 ? 3 : 2;  // accumulator buffers in tensor memory
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if (smem_u32(smem_raw) & 1023u) __trap();
  const int KC = d / 32;  // fp32 staging chunks
  const int KA = d / 64;  // 64-k atoms of the 16-bit operands
  uint8_t* xs = smem_raw;
  uint8_t* stages = xs + (size_t)(RFS ? TB_NX : 1) * KC * CHUNK_BYTES;
  // fp16: constant x-side k-block (128 rows x 16 halves, 32-byte swizzle; halves 0..2 of every row = 2^15)
  uint8_t* aconst = stages + (size_t)nstages * TB_TILE_BYTES;
  // fp16: row buffers of the finish warps: 4 warps x 2 slots x TB_FX_ROWS rows of d floats
  uint8_t* fxbuf = aconst + (H16 ? TB_CN_BYTES : 0);
  SmemB* sm = reinterpret_cast<SmemB*>(fxbuf + ((H16 && !RFS) ? TB_FX_WARPS * 2 * TB_FX_ROWS * d * 4 : 0));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntiles = (n + TM - 1) / TM;
  const int NT = k / TN;
  int* cnt_s = reinterpret_cast<int*>(sm + 1);  // k <= TB_CNT_SMEM_MAX: per-centre row counts of this CTA
  const bool cnt_smem = k <= TB_CNT_SMEM_MAX;
  if (cnt_smem)
    for (int i = threadIdx.x; i < k; i += TB_THREADS) cnt_s[i] = 0;
  // k <= 128 * nstages: every centre tile has a stage of its own -- loaded once, never released, instead of once per
  // row tile (at k = 256, d = 64 that was 40 KB of L2 reads per 32 KB row tile)
  const bool bres = NT <= nstages;
  // TB_SKEW schedule: slot A = this CTA's row tiles 0, 2, 4, ..., slot B = 1, 3, 5, ...; steps [0, endA) and [begB, endB)
  constexpr bool SKEW = H16 && TB_GRID && TB_SKEW;
  const int my_tiles = (int64_t)blockIdx.x < ntiles ? (int)((ntiles - 1 - blockIdx.x) / gridDim.x) + 1 : 0;
  const int nA = (my_tiles + 1) >> 1, nB = my_tiles >> 1;
  const int begB = NT >> 1, endA = nA * NT, endB = begB + nB * NT;
  const int nsteps = nB ? max(endA, endB) : endA;
  // position of a batch in the global issue order A(0), [B(0)], A(1), [B(1)], ...
  auto skew_gA = [&](int s) { return s + min(max(s - begB, 0), nB * NT); };
  auto skew_gB = [&](int s) { return min(s + 1, endA) + (s - begB); };

  if (threadIdx.x == 0) {
    for (int c = 0; c < 4; ++c) {
      mbar_init(&sm->xs_full[c], 1);
      mbar_init(&sm->xs_full[4 + c], 1);
      mbar_init(&sm->xs_empty[c], EPI_THREADS);
      mbar_init(&sm->b_full[c], 1);
      mbar_init(&sm->b_empty[c], 1);
      mbar_init(&sm->n_full[c], EPI_THREADS);
      mbar_init(&sm->n_empty[c], EPI_THREADS);
    }
    for (int c = 0; c < 3; ++c) {
      mbar_init(&sm->t_full[c], 1);
      mbar_init(&sm->t_empty[c], EPI_THREADS);
    }
    for (int c = 0; c < 8; ++c) mbar_init(&sm->fx_full[c >> 1][c & 1], 1);
    for (int c = 0; c < 2; ++c) {
      mbar_init(&sm->tok[c], 1);
      mbar_init(&sm->xa_full[c], EPI_THREADS);
      mbar_init(&sm->xa_empty[c], 1);
      mbar_init(&sm->a_full[c], SKEW ? EPI_THREADS : 2 * EPI_THREADS);  // (TB_SKEW: one epilogue warpgroup per row tile)
      mbar_init(&sm->a_empty[c], EPI_THREADS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (H16) {
    for (int i = threadIdx.x; i < TB_CN_BYTES / 16; i += TB_THREADS) {  // 16-byte chunks: row = i / 2
      const int r = i >> 1, pc = i & 1;
      uint4 val = make_uint4(0u, 0u, 0u, 0u);
      if (pc == ((r >> 2) & 1)) {  // 32-byte swizzle: address bit 4 ^= bit 7, i.e. chunk ^= (row >> 2) & 1
        // halves 0..2 multiply the three ||c||^2 pieces, half 3 the centre-side half of the magic constant (column 3
        // of CNH, zero unless the planes were prepared with a global key scale)
        const __half2 ee = __floats2half2_rn(KM_CNH_E1, KM_CNH_E1), e0 = __floats2half2_rn(KM_CNH_E1, KM_CNH_MA);
        val.x = *reinterpret_cast<const uint32_t*>(&ee);
        val.y = *reinterpret_cast<const uint32_t*>(&e0);
      }
      reinterpret_cast<uint4*>(aconst)[i] = val;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor-core reads
  }
  if (SCALED && threadIdx.x < TM) sm->cscale[threadIdx.x] = (int)threadIdx.x < d ? colscale[threadIdx.x] : 0.f;
  if (warp == W_ALLOC) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm->tmem_base)),
                 "r"(TB_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  // Register budget (launch: 640 threads x 96).  The epilogue keeps 64 accumulator values in flight per thread; at
  // 96 registers it spills its loop state to local memory, and with all of L1 carved out for shared memory each
  // reload is an L2 round trip (~1000 idle cycles per centre tile in the hand-off trace).  The control and finish warpgroups
  // give registers back, the two epilogue warpgroups take them (setmaxnreg works per 4-warp group).
  if (warp >= W_XPROD) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(TB_REG_CTRL));
  }
  if (warp == W_XPROD) {
    // ===================== X producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        if (RFS) {  // ring of TB_NX whole tiles; a slot comes back from the finish warps (their reductions have read it)
          const uint32_t slot = it % TB_NX;
          mbar_wait_ns(&sm->xs_empty[slot], ((it / TB_NX) & 1) ^ 1, 128);
          for (int kc = 0; kc < KC; ++kc) {
            mbar_expect_tx(&sm->xs_full[slot * KC + kc], (uint32_t)CHUNK_BYTES);
            tma_load_2d_hint(xs + (size_t)(slot * KC + kc) * CHUNK_BYTES, &map_x, kc * 32, (int)(tile * TM),
                             &sm->xs_full[slot * KC + kc], L2_EVICT_FIRST);
          }
          continue;
        }
        for (int kc = 0; kc < KC; ++kc) {
          mbar_wait_ns(&sm->xs_empty[kc], (it & 1) ^ 1, 128);
          mbar_expect_tx(&sm->xs_full[kc], (uint32_t)CHUNK_BYTES);
          // the finish warps read the tile again ~2 row tiles later: ask L2 to keep it until then
          tma_load_2d_hint(xs + (size_t)kc * CHUNK_BYTES, &map_x, kc * 32, (int)(tile * TM), &sm->xs_full[kc],
                           accumulate ? L2_EVICT_LAST : L2_EVICT_FIRST);
        }
      }
    }
  } else if (warp == W_BPROD) {
    // ===================== B producer =====================
    if (SKEW && lane == 0) {
      const uint32_t it = 0;  // (tracer slot)
      (void)it;
      const int nload = bres ? min(nsteps, NT) : nsteps;  // resident centre tiles: one load each for the whole kernel
      for (int st = 0; st < nload; ++st) {
        const int nt = st % NT;
        const uint32_t s = bres ? (uint32_t)nt : (uint32_t)(st % nstages);
        if (!bres) mbar_wait_ns(&sm->b_empty[s], (((uint32_t)st / (uint32_t)nstages) & 1u) ^ 1u, 64);
        mbar_expect_tx(&sm->b_full[s], (uint32_t)(KA * TB_STAGE_BYTES + TB_CN_BYTES));
        for (int a = 0; a < KA; ++a)
          tma_load_2d(stages + (size_t)s * TB_TILE_BYTES + (size_t)a * TB_STAGE_BYTES, &map_c1, a * 64, nt * TN,
                      &sm->b_full[s]);
        tma_load_2d(stages + (size_t)s * TB_TILE_BYTES + KA * TB_STAGE_BYTES, &map_c2, 0, nt * TN, &sm->b_full[s]);
      }
    } else if (lane == 0) {
      uint32_t s = 0, ph = 0, it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        if (bres && it > 0) break;  // resident centre tiles: one load each for the whole kernel
        for (int nt = 0; nt < NT; ++nt) {
          if (!bres) mbar_wait_ns(&sm->b_empty[s], ph ^ 1, 64);
          if (nt < 8) TR(11, nt);
          mbar_expect_tx(&sm->b_full[s], (uint32_t)(KA * TB_STAGE_BYTES + (H16 ? TB_CN_BYTES : 0)));
          for (int a = 0; a < KA; ++a) {
            uint8_t* dst = stages + (size_t)s * TB_TILE_BYTES + (size_t)a * TB_STAGE_BYTES;
            tma_load_2d(dst, &map_c1, a * 64, nt * TN, &sm->b_full[s]);
            if (!H16) tma_load_2d(dst + CHUNK_BYTES, &map_c2, a * 64, nt * TN, &sm->b_full[s]);
          }
          // fp16: map_c2 is the ||c||^2 plane (16 halves per centre, three of them non-zero; 32-byte swizzle)
          if (H16) tma_load_2d(stages + (size_t)s * TB_TILE_BYTES + KA * TB_STAGE_BYTES, &map_c2, 0, nt * TN, &sm->b_full[s]);
          if (++s == (uint32_t)nstages) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == W_MMA || warp == W_ALLOC) {
    // ===================== MMA issuers: one batch (4*KA + 1 or 12*KA instructions) per centre tile ============
    // An issuing warp sits in the (shallow) tensor queue for about as long as its batch executes, and only then can
    // it look at the barriers of a next tile.  So TWO warps take alternate centre tiles: while one issues, the other
    // has its operand stage and accumulator checked.  An issue token keeps the batches in tile order in the queue,
    // hence completing in tile order: the epilogue warps' parity waits rely on that (a warp that has seen tile g+1
    // complete must be able to conclude that tile g has), and a commit covers the other warp's earlier tiles too.
    const uint32_t mp = (warp == W_MMA) ? 0u : 1u;
    const uint64_t desc_fixed = (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
    const uint64_t desc_sw32 = (1ull << 16) | (16ull << 32) | (1ull << 46) | (6ull << 61);
    const uint32_t st_a = smem_u32(stages) >> 4;
    const uint32_t chunk16 = CHUNK_BYTES >> 4, stage16 = (uint32_t)TB_STAGE_BYTES >> 4;
    if (SKEW) {
      // Issuer mp owns slot mp: its r-th row tile is the CTA's row tile 2r + mp, operand buffer mp, steps
      // off + r*NT .. off + (r+1)*NT - 1 with off = 0 (A) or NT/2 (B); step s works on centre tile s % NT.
      const int nMine = mp ? nB : nA, off = mp ? begB : 0;
      const uint32_t xb = mp;
      const uint32_t a1 = tmem_base + TB_A1 + xb * TB_ABUF;
      // slot B's first batch comes after A(0) .. A(begB): phases of the token this warp has to see one by one
      if (mp == 1 && nMine > 0)
        for (int ph = 0; ph < begB; ++ph) mbar_wait_ns(&sm->tok[0], (uint32_t)ph & 1u, 20);
      for (int r = 0; r < nMine; ++r) {
        const uint32_t it = (uint32_t)(2 * r) + mp;  // (tracer slot)
        (void)it;
        for (int j = 0; j < NT; ++j) {
          const int st = off + r * NT + j;
          const int nt = st % NT;
          const uint32_t g = (uint32_t)(mp ? skew_gB(st) : skew_gA(st));
          const uint32_t cbuf = g % NBUF, cround = (g / NBUF) & 1u;
          const uint32_t cstage = bres ? (uint32_t)nt : (uint32_t)(st % nstages);
          const uint32_t cbphase = bres ? 0u : (((uint32_t)st / (uint32_t)nstages) & 1u);
          mbar_wait_ns(&sm->b_full[cstage], cbphase, 20);
          mbar_wait_ns(&sm->t_empty[cbuf], cround ^ 1u, 20);
          // the batch before this one in the global order, if it is the other slot's: A(s) before B(s), B(s-1) before A(s)
          if (mp == 1) {
            if (st < endA) mbar_wait_ns(&sm->tok[0], (uint32_t)st & 1u, 20);
          } else {
            if (st - 1 >= begB && st - 1 < endB) mbar_wait_ns(&sm->tok[1], (uint32_t)(st - 1 - begB) & 1u, 20);
          }
          if (j == 0) mbar_wait_ns(&sm->xa_full[xb], (uint32_t)r & 1u, 20);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t tmem_d = tmem_base + tb_acc_col(cbuf);
            const uint64_t bd = desc_fixed | (uint64_t)(st_a + cstage * ((uint32_t)TB_TILE_BYTES >> 4));
#pragma unroll
            for (int a = 0; a < 2; ++a) {
              if (a < KA) {
                const uint64_t b1 = bd + a * stage16;
#pragma unroll
                for (int s4 = 0; s4 < 4; ++s4)
                  tc_mma_f16_ts(tmem_d, a1 + a * 32 + 8 * s4, b1 + 2 * s4, IDESC_FP16, (a | s4) ? 1u : 0u);
              }
            }
            tc_mma_f16_ss(tmem_d, desc_sw32 | (uint64_t)(smem_u32(aconst) >> 4),
                          desc_sw32 | (uint64_t)(st_a + cstage * ((uint32_t)TB_TILE_BYTES >> 4) + KA * stage16),
                          IDESC_FP16, 1u);
            mbar_arrive(&sm->tok[mp]);
            // the stage goes back after its LAST user of this step: B(s) when slot B is active at s, else A(s)
            // (the tensor pipe runs in issue order: B's commit covers A's batch)
            const bool b_active = st >= begB && st < endB;
            if (!bres && (mp == 1 || !b_active)) tc_commit(&sm->b_empty[cstage]);
            if (j == NT - 1) tc_commit(&sm->xa_empty[xb]);
            tc_commit(&sm->t_full[cbuf]);
          }
          __syncwarp();
        }
      }
    }
    uint32_t it = 0, buf = 0, round = 0, stage = 0, bphase = 0, gpar = 0;
    uint32_t mine_n = 0;  // centre tiles this warp has issued
    for (int64_t tile = blockIdx.x; !SKEW && tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t xb = it % XBUFS;  // row-tile buffer of the TMEM operand
      const uint32_t a1 = tmem_base + TB_A1 + xb * TB_ABUF;  // 32 columns = 64 halves of k per atom
      const uint32_t a2 = tmem_base + TB_A2 + xb * TB_ABUF;
      bool first = true;
      for (int nt = 0; nt < NT; ++nt) {
        const bool mine = gpar == mp;
        const bool very_first = (it == 0 && nt == 0);
        gpar ^= 1u;
        // (resident centre tiles: tile nt sits in stage nt, whose barrier completed its only phase long ago)
        const uint32_t cbuf = buf, cround = round, cstage = bres ? (uint32_t)nt : stage, cbphase = bres ? 0u : bphase;
        if (++stage == (uint32_t)nstages) {
          stage = 0;
          bphase ^= 1;
        }
        if (++buf == NBUF) {
          buf = 0;
          round ^= 1;
        }
        if (!mine) continue;
        if (mp == 0 && nt < 8) TR(0, 3 * nt);
        mbar_wait_ns(&sm->b_full[cstage], cbphase, 20);
        if (mp == 0 && nt < 8) TR(0, 3 * nt + 1);
        mbar_wait_ns(&sm->t_empty[cbuf], cround ^ 1, 20);
        // the other warp has issued the previous centre tile: its k-th tile is global tile 2k + (1 - mp)
        if (!very_first) mbar_wait_ns(&sm->tok[mp ^ 1u], (mp ? mine_n : mine_n - 1u) & 1u, 20);
        // Only now the operand buffer: a warp that owns no tile of some row tile (k == 128) never looks at that row
        // tile's phase of xa_full, and a parity wait is only meaningful one phase ahead of what the waiter has seen.
        // After the token every earlier row tile's phase is known to be complete.  (The stage and accumulator
        // barriers cannot be skipped that way: their previous use belongs to a tile at least two back, which the
        // token chain has already ordered.)
        if (first) mbar_wait_ns(&sm->xa_full[xb], (it / XBUFS) & 1, 20);
        first = false;
        tc_fence_after();
        if (mp == 0 && nt < 8) TR(0, 24 + nt);
        if (elect_one()) {
          const uint32_t tmem_d = tmem_base + tb_acc_col(cbuf);
          const uint64_t bd = desc_fixed | (uint64_t)(st_a + cstage * ((uint32_t)TB_TILE_BYTES >> 4));
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            if (a < KA) {
              const uint64_t b1 = bd + a * stage16;
              const uint64_t b2 = b1 + chunk16;
#pragma unroll
              for (int s = 0; s < 4; ++s) {  // 16 halves of k = 32 bytes = 2 descriptor units, 8 TMEM columns
                if (H16) {
                  tc_mma_f16_ts(tmem_d, a1 + a * 32 + 8 * s, b1 + 2 * s, IDESC_FP16, (a | s) ? 1u : 0u);
                } else {
                  tc_mma_f16_ts(tmem_d, a1 + a * 32 + 8 * s, b1 + 2 * s, IDESC_BF16, (a | s) ? 1u : 0u);
                  tc_mma_f16_ts(tmem_d, a1 + a * 32 + 8 * s, b2 + 2 * s, IDESC_BF16, 1u);
                  tc_mma_f16_ts(tmem_d, a2 + a * 32 + 8 * s, b1 + 2 * s, IDESC_BF16, 1u);
                }
              }
            }
          }
          // + ||c||^2:  (2^15, 2^15, 2^15, 0, ...) . (h1, h2, h3, 0, ...), the first 16-k slice of the third atom
          // (operands of this one: 128 x 16 halves, 32-byte swizzle, 8-row groups 256 B apart)
          if (H16)
            tc_mma_f16_ss(tmem_d, desc_sw32 | (uint64_t)(smem_u32(aconst) >> 4),
                          desc_sw32 | (uint64_t)(st_a + cstage * ((uint32_t)TB_TILE_BYTES >> 4) + KA * stage16), IDESC_FP16, 1u);
          mbar_arrive(&sm->tok[mp]);  // batch is in the queue: the other warp may issue the next tile
          if (!bres) tc_commit(&sm->b_empty[cstage]);
          if (nt == NT - 1) tc_commit(&sm->xa_empty[xb]);
          tc_commit(&sm->t_full[cbuf]);
        }
        __syncwarp();
        ++mine_n;
        if (mp == 0 && nt < 8) TR(0, 3 * nt + 2);
      }
    }
  } else if (warp < W_SPLIT0) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(TB_REG_EPI));
    // ===================== epilogue: best / second best per row =====================
    // Two warps share a TMEM lane quarter; warp `half` takes every other centre tile, whole, so the two warps of a
    // scheduler run out of step and each one's TMEM-load and barrier latencies are covered by the other's arithmetic.
    const int q = warp & 3;
    const uint32_t half = (uint32_t)(warp - W_EPI0) >> 2;
    const int row_in_tile = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    // fp16 filter: the accumulator holds S^2 * (x . -2c)
    const float inv = H16 ? misc[KM_MISC_INV_SCALE2] : 1.f;
    const float cmax_e = misc[KM_MISC_CMAX];
    const bool gs = H16 && misc[KM_MISC_SC] > 0.f;  // global key scale: keys come straight out of the accumulator
    uint32_t buf = 0, round = 0, par = 0;
    // TB_SKEW: this warp takes EVERY centre tile of the row tiles of its slot (slot = half), in its slot's order
    const uint32_t nrt = SKEW ? (uint32_t)(half ? nB : nA) : (uint32_t)my_tiles;
    for (uint32_t rr = 0; rr < nrt; ++rr) {
      const uint32_t it = SKEW ? 2 * rr + half : rr;
      float b1a = INFINITY, b2a = INFINITY, b1b = INFINITY, b2b = INFINITY;
      int i1a = 0, i1b = 1;
      // integer-key path (fp16 filter): running best / second best keys and the centre tile of the best
      int B1 = 0x7fffffff, B2 = 0x7fffffff, B3 = 0x7fffffff;  // B3: lower bound of the third best key
      int T1 = 0, T2 = 0;                                     // centre tiles of the best and the second best
      bool got = false;  // this warp had a centre tile of this row tile (not a given at k = 128)
      int A[16];  // TB_GRID: running minimum of every A-class (column mod 16) over this warp's centre tiles
#pragma unroll
      for (int p = 0; p < 16; ++p) A[p] = 0x7fffffff;
      float arow = 1.f;
      if (H16 && !gs) {
        // the split warps finished this row tile before its first MMA could start: the norms are there
        mbar_wait_ns(&sm->n_full[it & 3], (it >> 2) & 1, 32);
        arow = inv * tb_row_scale(sm->xnorm2[it & 3][row_in_tile], cmax_e);
      }
      for (int jt = 0; jt < NT; ++jt) {
        int nt = jt;
        uint32_t cbuf = buf, cround = round;
        bool mine = par == half;
        if (SKEW) {
          const int st = (half ? begB : 0) + (int)rr * NT + jt;
          const uint32_t g = (uint32_t)(half ? skew_gB(st) : skew_gA(st));
          nt = st % NT;
          cbuf = g % NBUF;
          cround = (g / NBUF) & 1u;
          mine = true;
          // Slot B's first accumulator is not the first use of its buffer, and a parity wait only means something one
          // phase ahead of what the waiter has seen: watch the earlier uses (slot A's) go by first.
          if (half == 1 && rr == 0 && jt == 0)
            for (uint32_t u = 0; u < g / NBUF; ++u) mbar_wait_ns(&sm->t_full[cbuf], u & 1u, 20);
        }
        par ^= 1;
        if (++buf == NBUF) {
          buf = 0;
          round ^= 1;
        }
        if (!mine) continue;
        got = true;
        if (nt < 8) TR(1 + warp, 3 * nt);
        mbar_wait_ns(&sm->t_full[cbuf], cround, 20);
        tc_fence_after();
        if (nt < 8) TR(1 + warp, 3 * nt + 1);
        const uint32_t taddr = tmem_base + lane_addr + tb_acc_col(cbuf);
        uint32_t v[2][32];
        tmem_ld32(taddr, v[0]);
        int k1a = 0x7fffffff, k2a = 0x7fffffff, k1b = 0x7fffffff, k2b = 0x7fffffff;
        int hm[8];  // TB_GRID: minima of the tile's eight M-classes (16 consecutive centres each)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          tmem_ld_wait();
          if (c + 1 < 4) tmem_ld32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
          if (c == 3) {
            // The last 32 columns are in registers: hand the accumulator back NOW, a quarter of the tile's
            // arithmetic earlier.  The MMA that refills it is what the OTHER epilogue warp of this lane quarter
            // waits for three centre tiles on (hand-off trace: ~14 % of an epilogue warp's time went into t_full).
            tc_fence_before();
            mbar_arrive(&sm->t_empty[cbuf]);
          }
          if (H16 && TB_GRID) {
            // quantised scores as integers: the accumulator's own bit pattern (global key scale: the tensor core has
            // added the magic constant), else one multiply-add into this row's key window
            int* qv = reinterpret_cast<int*>(v[c & 1]);
            if (!gs) {
#pragma unroll
              for (int j = 0; j < 32; ++j) qv[j] = __float_as_int(fmaf(__int_as_float(qv[j]), arow, TB_MAGIC));
            }
#pragma unroll
            for (int p = 0; p < 16; ++p) A[p] = tb_min3(A[p], qv[p], qv[p + 16]);
            hm[2 * c] = tb_min16(qv);
            hm[2 * c + 1] = tb_min16(qv + 16);
          } else if (H16) {  // the accumulator holds S_a S_b (x . -2c + ||c||^2), in global mode + the magic constant
            if (gs) tb_chunk_keys<true>(v[c & 1], arow, c * 32, k1a, k2a, k1b, k2b);
            else tb_chunk_keys<false>(v[c & 1], arow, c * 32, k1a, k2a, k1b, k2b);
          }
          const float4* cn4 = reinterpret_cast<const float4*>(cn + nt * TN + c * 32);
          const int jbase = nt * TN + c * 32;
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
#ifdef TB_EXP_NOMATH  // timing experiment: results invalid
            if (j4 > 0) continue;
#endif
            if (H16) continue;  // (integer keys: handled per chunk above)
            const float4 cv = __ldg(cn4 + j4);
            const float s0 = fmaf(__uint_as_float(v[c & 1][j4 * 4 + 0]), inv, cv.x);  // inv is a power of two
            const float s1 = fmaf(__uint_as_float(v[c & 1][j4 * 4 + 1]), inv, cv.y);
            const float s2 = fmaf(__uint_as_float(v[c & 1][j4 * 4 + 2]), inv, cv.z);
            const float s3 = fmaf(__uint_as_float(v[c & 1][j4 * 4 + 3]), inv, cv.w);
            const int j = jbase + j4 * 4;
            i1a = (s0 < b1a) ? j : i1a;       b2a = fminf(b2a, fmaxf(b1a, s0)); b1a = fminf(b1a, s0);
            i1b = (s1 < b1b) ? j + 1 : i1b;   b2b = fminf(b2b, fmaxf(b1b, s1)); b1b = fminf(b1b, s1);
            i1a = (s2 < b1a) ? j + 2 : i1a;   b2a = fminf(b2a, fmaxf(b1a, s2)); b1a = fminf(b1a, s2);
            i1b = (s3 < b1b) ? j + 3 : i1b;   b2b = fminf(b2b, fmaxf(b1b, s3)); b1b = fminf(b1b, s3);
          }
        }
        if (H16) {  // fold the tile's two chains into the running triple; an earlier tile keeps equal keys
          int k1 = min(k1a, k1b), k2 = min(max(k1a, k1b), min(k2a, k2b));
          // The tile's third smallest key is >= l3: if its two smallest sit in different chains it IS min(k2a, k2b);
          // if one chain holds both, that chain's third is unknown and min(k2a, k2b) = k2 is all that can be said.
          int l3 = min(k2a, k2b);
          if (TB_GRID) {  // keys of the eight M-class minima: value * 8 + class, then their three smallest (exact)
#pragma unroll
            for (int j = 0; j < 8; ++j)
              hm[j] = (int)((uint32_t)hm[j] * 8u + ((uint32_t)j - (uint32_t)TB_QBASE * 8u));
            tb_top3_of8(hm, k1, k2, l3);
          }
          // third smallest of the two sorted triples (B1, B2, B3) and (k1, k2, l3)
          B3 = min(min(B3, l3), min(max(B2, k1), max(B1, k2)));
          const int lose = max(B1, k1);             // the larger of the two firsts ...
          const int tl = (k1 < B1) ? T1 : nt;       // ... and the tile it came from
          const int nB2 = min(lose, min(B2, k2));
          T2 = (nB2 == B2) ? T2 : ((nB2 == lose) ? tl : nt);
          B2 = nB2;
          T1 = (k1 < B1) ? nt : T1;
          B1 = min(B1, k1);
        }
        if (nt < 8) TR(1 + warp, 3 * nt + 2);
      }
      // merge this thread's two chains: smaller score wins, equal scores keep the lower centre index
      const bool a_wins = (b1a < b1b) || (b1a == b1b && i1a < i1b);
      const uint32_t ab = it & 1;
      mbar_wait_ns(&sm->a_empty[ab], ((it >> 1) & 1) ^ 1, 64);
      if (H16 && TB_GRID) {
        // the three smallest A-class minima (keys: value * 16 + class); a warp without a centre tile (k = 128)
        // hands over INT_MAX everywhere
        int a1 = 0x7fffffff, a2 = 0x7fffffff, a3 = 0x7fffffff;
        if (got) {
#pragma unroll
          for (int p = 0; p < 16; ++p) A[p] = (int)((uint32_t)A[p] * 16u + ((uint32_t)p - (uint32_t)TB_QBASE * 16u));
          int y1, y2, y3;
          tb_top3_of8(A, a1, a2, a3);
          tb_top3_of8(A + 8, y1, y2, y3);
          tb_merge3(a1, a2, a3, y1, y2, y3);
        }
        const uint32_t hs = SKEW ? 0u : half;  // (TB_SKEW: the only hand-off of this row tile)
        sm->p_b1[ab][hs][row_in_tile] = __int_as_float(B1);
        sm->p_b2[ab][hs][row_in_tile] = __int_as_float(B2);
        sm->p_b3[ab][hs][row_in_tile] = B3;
        // first centres of the best and the second best M-class (k <= 65536)
        sm->p_i1[ab][hs][row_in_tile] =
            (int)((uint32_t)(T1 * TN + (B1 & 7) * 16) | ((uint32_t)(T2 * TN + (B2 & 7) * 16) << 16));
        sm->p_i2[ab][hs][row_in_tile] = a2;
        sm->p_a3[ab][hs][row_in_tile] = (a3 & ~15) | (a1 & 15);
      } else if (H16) {
        sm->p_b1[ab][half][row_in_tile] = __int_as_float(B1);
        sm->p_b2[ab][half][row_in_tile] = __int_as_float(B2);
        sm->p_i1[ab][half][row_in_tile] = T1 * TN + tb_key_col(B1);
        sm->p_b3[ab][half][row_in_tile] = B3;
        sm->p_i2[ab][half][row_in_tile] = T2 * TN + tb_key_col(B2);
      } else {
        sm->p_b1[ab][half][row_in_tile] = a_wins ? b1a : b1b;
        sm->p_i1[ab][half][row_in_tile] = a_wins ? i1a : i1b;
        sm->p_b2[ab][half][row_in_tile] = fminf(fmaxf(b1a, b1b), fminf(b2a, b2b));  // second smallest of both chains
      }
      mbar_arrive(&sm->a_full[ab]);
      TR(1 + warp, 31);
    }
  } else if (warp < W_FIN0) {
    // ===================== split: fp32 row -> (x1, x2) bf16 pairs, or S * x as fp16 pairs, in tensor memory;
    // row norms
    const int w4 = warp - W_SPLIT0;
    const int r = w4 * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(w4 * 32) << 16;
    const float S = H16 ? misc[KM_MISC_SCALE] : 1.f;
    const float xlim = 65504.f / S;  // entries beyond this overflow fp16 after scaling
    // global key scale: a row longer than the norm the planes were built for may leave the key window
    const float xn_lim = (H16 && misc[KM_MISC_SC] > 0.f) ? misc[KM_MISC_XMAX] * misc[KM_MISC_XMAX] * 1.0001f : INFINITY;
    uint32_t it = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      float xn = 0.f, amax = 0.f;
      if (w4 == 0) TR(9, 0);
      for (int a = 0; a < KA; ++a) {
        uint32_t p1[32], p2[32];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int kc = 2 * a + h;
          const uint32_t xslot = RFS ? (it % TB_NX) * KC + kc : kc;
          mbar_wait_ns(&sm->xs_full[xslot], RFS ? ((it / TB_NX) & 1) : (it & 1), 64);
          const uint8_t* src = xs + (size_t)xslot * CHUNK_BYTES + (size_t)r * 128;
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            float4 v = *reinterpret_cast<const float4*>(src + ((p ^ (r & 7)) << 4));
            if (SCALED) {  // the row a materialised StandardScaler copy would hold: fl32(x * scale)
              const float4 sc = *reinterpret_cast<const float4*>(&sm->cscale[kc * 32 + p * 4]);
              v.x = __fmul_rn(v.x, sc.x); v.y = __fmul_rn(v.y, sc.y); v.z = __fmul_rn(v.z, sc.z); v.w = __fmul_rn(v.w, sc.w);
            }
            xn = fmaf(v.x, v.x, xn); xn = fmaf(v.y, v.y, xn); xn = fmaf(v.z, v.z, xn); xn = fmaf(v.w, v.w, xn);
            if (H16) {
              amax = fmaxf(fmaxf(amax, fabsf(v.x)), fmaxf(fabsf(v.y), fmaxf(fabsf(v.z), fabsf(v.w))));
              const __half2 g01 = __floats2half2_rn(v.x * S, v.y * S);  // .x -> low 16 bits = even k
              const __half2 g23 = __floats2half2_rn(v.z * S, v.w * S);
              p1[h * 16 + p * 2 + 0] = *reinterpret_cast<const uint32_t*>(&g01);
              p1[h * 16 + p * 2 + 1] = *reinterpret_cast<const uint32_t*>(&g23);
              continue;
            }
            const __nv_bfloat162 h01 = __floats2bfloat162_rn(v.x, v.y);  // .x -> low 16 bits = even k
            const __nv_bfloat162 h23 = __floats2bfloat162_rn(v.z, v.w);
            const float2 f01 = __bfloat1622float2(h01), f23 = __bfloat1622float2(h23);
            const __nv_bfloat162 l01 = __floats2bfloat162_rn(v.x - f01.x, v.y - f01.y);
            const __nv_bfloat162 l23 = __floats2bfloat162_rn(v.z - f23.x, v.w - f23.y);
            p1[h * 16 + p * 2 + 0] = *reinterpret_cast<const uint32_t*>(&h01);
            p1[h * 16 + p * 2 + 1] = *reinterpret_cast<const uint32_t*>(&h23);
            p2[h * 16 + p * 2 + 0] = *reinterpret_cast<const uint32_t*>(&l01);
            p2[h * 16 + p * 2 + 1] = *reinterpret_cast<const uint32_t*>(&l23);
          }
          // generic-proxy reads of the staging chunk before its refill through the async proxy (TMA): without the
          // proxy fence the refill may overtake reads still in flight (the arrive only clobbers memory, so the
          // arithmetic consuming them can be scheduled below it).  Found in the multinomial LR kernel, where it
          // showed; the same hand-off here never failed a repeatability test, and costs nothing to close.
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          if (!RFS) mbar_arrive(&sm->xs_empty[kc]);  // (RFS: the finish warps hand the whole tile back)
        }
        const uint32_t xb = it % XBUFS;
        if (w4 == 0) TR(9, 1 + 3 * a);
        // the MMAs that read this buffer last (tile it - XBUFS) are done with it
        if (a == 0) mbar_wait_ns(&sm->xa_empty[xb], ((it / XBUFS) & 1) ^ 1, 64);
        tc_fence_after();
        if (w4 == 0) TR(9, 2 + 3 * a);
        tmem_st32(tmem_base + lane_addr + TB_A1 + xb * TB_ABUF + a * 32, p1);
        if (!H16) tmem_st32(tmem_base + lane_addr + TB_A2 + xb * TB_ABUF + a * 32, p2);
        tmem_st_wait();
        tc_fence_before();
        if (a == KA - 1) mbar_arrive(&sm->xa_full[xb]);
        if (w4 == 0) TR(9, 3 + 3 * a);
      }
      // four slots: the split warps run up to two tiles ahead of the MMAs, which run ahead of the finish warps
      mbar_wait_ns(&sm->n_empty[it & 3], ((it >> 2) & 1) ^ 1, 128);
      sm->xnorm2[it & 3][r] = (H16 && !(amax <= xlim && xn <= xn_lim)) ? INFINITY : xn;
      mbar_arrive(&sm->n_full[it & 3]);
    }
  } else if (warp < W_XPROD) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(TB_REG_FIN));
    // ===================== finish: merge the two epilogue warps' results, certainty test, labels, re-check list,
    // accumulation of certain rows (16-byte vector reductions), cost, counts.  Thread == row.
    const int w4 = warp - W_FIN0;
    const int row_in_tile = w4 * 32 + lane;
    const float cmax = misc[KM_MISC_CMAX];
    // fp16 underflow term of the bound: sqrt(d) eta (||x|| / S_b + ||-2c|| / S_a)
    const float tau0a = H16 ? kappa0 / misc[KM_MISC_SCALE] : 0.f, tau0b = H16 ? kappa0 / misc[KM_MISC_SCALE_B] : 0.f;
    const float inv_s2 = H16 ? misc[KM_MISC_INV_SCALE2] : 0.f;     // (read once: there is no L1 behind these loads)
    const float sc_glob = H16 ? misc[KM_MISC_SC] : 0.f;            // > 0: one key scale for every row
    double cost = 0.0;
    unsigned long long nrows = 0;
    float* myrep = rep ? rep + (size_t)(blockIdx.x % KM_REPLICAS) * k * d : nullptr;
    uint32_t it = 0, fxc = 0;  // fxc: batches this warp has pushed through its row buffers
    // list appends of the previous tile whose slots are reserved but whose entries are not written yet
    unsigned pend_pm = 0, pend_lm = 0;       // lanes with a pair / re-check entry
    int pend_pbase = 0, pend_lbase = 0;      // (leader lane) first reserved slot
    bool pend_pair = false, pend_list = false;
    int4 pend_ent = make_int4(0, 0, 0, 0);
    int32_t pend_lrow = 0;
    auto flush_pending = [&]() {
      const unsigned lt = (1u << lane) - 1;
      bool over = false;  // pair list full: the row goes to the re-check list after all (rare: own atomic, waited for)
      if (pend_pm) {
        const int base = __shfl_sync(0xffffffffu, pend_pbase, __ffs(pend_pm) - 1);
        if (pend_pair) {
          const int slot = base + __popc(pend_pm & lt);
          if (slot < pcap) pairs[slot] = pend_ent;
          else over = true;
        }
        const unsigned om = __ballot_sync(0xffffffffu, over);
        if (om) {
          int ob = 0;
          if (lane == (__ffs(om) - 1)) ob = atomicAdd(&ws->count, __popc(om));
          ob = __shfl_sync(0xffffffffu, ob, __ffs(om) - 1);
          if (over) {
            const int slot = ob + __popc(om & lt);
            if (slot < cap) amb_idx[slot] = pend_ent.x;
          }
        }
      }
      if (pend_lm) {
        const int base = __shfl_sync(0xffffffffu, pend_lbase, __ffs(pend_lm) - 1);
        if (pend_list) {
          const int slot = base + __popc(pend_lm & lt);
          if (slot < cap) amb_idx[slot] = pend_lrow;
        }
      }
      pend_pm = 0;
      pend_lm = 0;
    };
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ab = it & 1;
      if (w4 == 0) TR(10, 0);
      mbar_wait_ns(&sm->a_full[ab], (it >> 1) & 1, 128);
      if (w4 == 0) TR(10, 1);
      float b1 = sm->p_b1[ab][0][row_in_tile], b2 = sm->p_b2[ab][0][row_in_tile];
      int i1 = sm->p_i1[ab][0][row_in_tile];
      float gap_units = 0.f;   // fp16 filter: second best - best in key units
      float gap3_units = 0.f;  // (lower bound of the) third best - best
      int i2 = -1;             // index of the second best
      // TB_GRID, per epilogue warp h: values of the three smallest M-class minima (gv1 = the warp's best score) and
      // of the second / third smallest A-class minima; first centres of the two best M-classes; the two best A-classes
      int gv1[2], gm2[2], gm3[2], ga2[2], ga3[2], gb1[2], gb2[2], gc1[2], gc2[2];
      if (H16 && TB_GRID) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          int B1 = 0x7fffffff, B2 = 0x7fffffff, B3 = 0x7fffffff, a2 = 0x7fffffff, a3 = 0x7fffffff;
          uint32_t i12 = 0;
          if (!SKEW || h == 0) {  // (TB_SKEW: the row tile has ONE hand-off; the other stays at "no centre tile")
            B1 = __float_as_int(sm->p_b1[ab][h][row_in_tile]);
            B2 = __float_as_int(sm->p_b2[ab][h][row_in_tile]);
            B3 = sm->p_b3[ab][h][row_in_tile];
            a2 = sm->p_i2[ab][h][row_in_tile];
            a3 = sm->p_a3[ab][h][row_in_tile];
            i12 = (uint32_t)sm->p_i1[ab][h][row_in_tile];
          }
          gv1[h] = B1 >> 3; gm2[h] = B2 >> 3; gm3[h] = B3 >> 3;  // (INT_MAX stays far above every real value)
          ga2[h] = a2 >> 4; ga3[h] = a3 >> 4;
          gb1[h] = (int)(i12 & 0xffffu); gb2[h] = (int)(i12 >> 16);
          gc1[h] = a3 & 15; gc2[h] = a2 & 15;
        }
        mbar_arrive(&sm->a_empty[ab]);
      } else {
        const float ob1 = sm->p_b1[ab][1][row_in_tile], ob2 = sm->p_b2[ab][1][row_in_tile];
        const int oi1 = sm->p_i1[ab][1][row_in_tile];
        int B3 = 0x7fffffff, O3 = 0x7fffffff, mi2 = -1, oi2 = -1;
        if (H16) {
          B3 = sm->p_b3[ab][0][row_in_tile];
          O3 = sm->p_b3[ab][1][row_in_tile];
          mi2 = sm->p_i2[ab][0][row_in_tile];
          oi2 = sm->p_i2[ab][1][row_in_tile];
        }
        mbar_arrive(&sm->a_empty[ab]);
        if (H16) {
          const int B1 = __float_as_int(b1), B2 = __float_as_int(b2);
          const int O1 = __float_as_int(ob1), O2 = __float_as_int(ob2);
          const bool mine = (B1 < O1) || (B1 == O1 && i1 < oi1);
          const int lose = max(B1, O1);
          const int S2 = min(lose, min(B2, O2)), S1 = min(B1, O1);
          const int S3 = min(min(B3, O3), min(max(B2, O1), max(B1, O2)));
          // (a warp without a centre tile hands over INT_MAX keys: at least one warp has real ones)
          gap_units = (S2 == 0x7fffffff) ? 0.f : (float)(tb_key_value(S2) - tb_key_value(S1));
          gap3_units = (S3 == 0x7fffffff) ? 0.f : (float)(tb_key_value(S3) - tb_key_value(S1));
          i2 = (S2 == lose) ? (mine ? oi1 : i1) : ((S2 == B2) ? mi2 : oi2);
          i1 = mine ? i1 : oi1;
        } else {
          const bool mine = (b1 < ob1) || (b1 == ob1 && i1 < oi1);
          b2 = fminf(fmaxf(b1, ob1), fminf(b2, ob2));
          i1 = mine ? i1 : oi1;
          b1 = mine ? b1 : ob1;
        }
      }
      mbar_wait_ns(&sm->n_full[it & 3], (it >> 2) & 1, 64);
      const float xn = sm->xnorm2[it & 3][row_in_tile];
      mbar_arrive(&sm->n_empty[it & 3]);
      const float xl = sqrtf(xn);  // xn = +inf marks a row whose scaled entries overflow fp16: never certain
      // 2 * tau for this row: operand rounding (~ ||x|| max||c||), fp16 underflow, and the fp32 roundings of ||c||^2
      // itself and of the final add (~ max||c||^2: matters only for rows much closer to the origin than the centres)
      // (fp16 filter: ||c||^2 enters as three fp16 pieces times 2^15; their underflow adds 3 * 2^-10 / S^2)
      const float tau2 = kappa2 * xl * cmax + tau0b * xl + tau0a * 2.f * cmax + 3.6e-7f * cmax * cmax +
                         (H16 ? 5.9e-3f * inv_s2 : 0.f);
      const int64_t row = tile * TM + row_in_tile;
      const bool live = row < n;
      bool certain = (b2 - b1) > tau2;  // false for NaN as well
      bool pair_ok = false;  // uncertain, but every centre other than the best two is out of reach
      int quad_lo = 0, quad_hi = 0;  // TB_GRID: the candidates of such a row
      if (H16) {
        // key units per score unit, and what the keys themselves may be off by: per-row scale: one rounding each
        // (2.5 units for a difference); global scale: the tensor core TRUNCATES when it adds -- 8 accumulations at up
        // to 1/4 unit, then the k-block with the magic constant, whose five addends are aligned to the constant --
        // <= 7 units per key, 20 for a difference with room to spare (tau2 * sc is in the thousands)
        const float sc = sc_glob > 0.f ? sc_glob : tb_row_scale(xn, cmax);
        const float arow = inv_s2 * sc;
        const float thr = fmaf(tau2, sc, sc_glob > 0.f ? 20.f : 2.5f);
        const bool usable = arow > 0.f && arow < INFINITY && xn < INFINITY;
        certain = gap_units > thr && usable;
        pair_ok = !certain && usable && pairs != nullptr && gap3_units > thr && i2 >= 0 && i2 < k && i2 != i1;
        if (TB_GRID) {
          // candidates = centres that may be within thr of the best: (M-classes within thr) x (A-classes within
          // thr), per epilogue warp; a third class of either kind within thr = too many to enumerate
          const int S1 = min(gv1[0], gv1[1]);
          const int m = (gv1[1] < gv1[0]) ? 1 : 0;  // the warp that holds the best
          int nM[2], nA[2];
          bool many = false;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int n1 = !((float)(gv1[h] - S1) > thr);
            const int m2 = !((float)(gm2[h] - S1) > thr), m3 = !((float)(gm3[h] - S1) > thr);
            const int a2 = !((float)(ga2[h] - S1) > thr), a3 = !((float)(ga3[h] - S1) > thr);
            nM[h] = n1 + m2;
            nA[h] = n1 + a2;
            many |= (m3 | a3) != 0;
          }
          const int tot = nM[0] * nA[0] + nM[1] * nA[1];
          // (selects, not indexing by m: a dynamically indexed array would live in local memory)
          i1 = m ? gb1[1] + gc1[1] : gb1[0] + gc1[0];
          certain = usable && !many && tot == 1;
          pair_ok = !certain && usable && pairs != nullptr && !many && tot <= 4;
          if (pair_ok) {  // up to four candidate centres, the best first, 16 bits each
            unsigned long long pk = 0;
            int nc = 0;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const bool o = (hh ^ m) != 0;  // warp whose candidates come now: the one holding the best first
              const int cM = o ? nM[1] : nM[0], cA = o ? nA[1] : nA[0];
              const int b1c = o ? gb1[1] : gb1[0], b2c = o ? gb2[1] : gb2[0];
              const int c1c = o ? gc1[1] : gc1[0], c2c = o ? gc2[1] : gc2[0];
#pragma unroll
              for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                for (int ai = 0; ai < 2; ++ai)
                  if (mi < cM && ai < cA) {
                    const int idx = (mi ? b2c : b1c) + (ai ? c2c : c1c);
                    pk |= (unsigned long long)(uint32_t)idx << (16 * nc);
                    ++nc;
                  }
            }
            quad_lo = (int)(uint32_t)pk;
            quad_hi = (int)(uint32_t)(pk >> 32);
            i2 = tot;  // number of candidates
          }
        }
      }
      if (live) labels[row] = i1;             // provisional for uncertain rows (rewritten by the re-check)
      const int my_lab = (live && certain) ? i1 : -1;
      const int64_t r0 = tile * TM + w4 * 32;
      // The copies of the first two batches of certain rows go out NOW: the list appends below are atomics whose
      // return value is an L2 round trip (~94 % of the warp-tiles have a pair row at cfg 2), and that wait then
      // overlaps the copies instead of preceding them.
      if (RFS) {
        // The row tile is still in its staging slot: two tensor reductions per certain row, straight from the
        // (swizzled) chunks into the replica row of the label.  The slot goes back to the producer after the list
        // appends below, whose atomics' round trip covers the engine's read of the source.
        if (accumulate && my_lab >= 0) {
          const uint32_t slot = it % TB_NX;
          const int rrow = (int)(blockIdx.x % KM_REPLICAS) * k + my_lab;
          for (int kc = 0; kc < KC; ++kc)
            tma_reduce_add_2d(&map_rep, kc * 32, rrow, xs + (size_t)(slot * KC + kc) * CHUNK_BYTES + (size_t)row_in_tile * 128);
        }
        bulk_commit();
      }
      TbAcc acc_rows;
      if (H16 && accumulate && !RFS) {
        acc_rows.X = X; acc_rows.n = n; acc_rows.ld = ld; acc_rows.r0 = r0; acc_rows.d = d; acc_rows.lane = lane;
        acc_rows.myrep = myrep;
        acc_rows.mybuf = fxbuf + (size_t)w4 * 2 * TB_FX_ROWS * d * 4;
        acc_rows.fx_full = sm->fx_full[w4];
        acc_rows.begin(fxc, my_lab);
      }
      // Uncertain rows with at most four (two) candidates: pair list, decided by fp64 distances; a full pair list sends
      // the row to the re-check list instead.  The other uncertain rows: re-check list.  One atomic per warp and list
      // reserves the slots -- and its return value is an L2 round trip, which used to sit in this loop for ~94 % of
      // the warp-tiles (cfg 2).  So the entries are WRITTEN one tile later (flush_pending), when the reservation has
      // long arrived, and the atomics of this tile are only issued here.
      flush_pending();
      bool to_list = live && !certain && !pair_ok;
      if (H16) {
        pend_pair = live && pair_ok;
        pend_pm = __ballot_sync(0xffffffffu, pend_pair);
        pend_ent = TB_GRID ? make_int4((int)(row & 0x7fffffff), quad_lo, quad_hi, i2)
                           : make_int4((int)(row & 0x7fffffff), i1, i2, 0);
        if (pend_pm && lane == (__ffs(pend_pm) - 1)) pend_pbase = atomicAdd(&ws->pcount, __popc(pend_pm));
      }
      pend_list = to_list;
      pend_lrow = (int32_t)(row & 0x7fffffff);
      pend_lm = __ballot_sync(0xffffffffu, to_list);
      if (pend_lm && lane == (__ffs(pend_lm) - 1)) pend_lbase = atomicAdd(&ws->count, __popc(pend_lm));
      if (w4 == 0) TR(10, 2);
      if (RFS) {
        bulk_wait_read<0>();
        mbar_arrive(&sm->xs_empty[it % TB_NX]);
      }
      if (!accumulate) continue;
      // bf16x3: squared distance to the chosen centre.  fp16: ||x||^2 only -- the remaining part of the cost,
      // sum_j (n_j ||c_j||^2 - 2 c_j . S_j), is added exactly from the sums (tcb_cost_terms_kernel)
      const float my_sc = H16 ? xn : fmaxf(xn + b1, 0.f);
      if (H16 && !RFS) acc_rows.finish(fxc);
      // TB_FIN_ROWS rows in flight per warp: the re-reads come from L2 (~800 cycles each way under load) and this
      // loop, not the epilogue, set the pace of the whole kernel at 8 rows in flight (hand-off trace)
      for (int g = 0; g < (H16 ? 0 : 32); g += TB_FIN_ROWS) {
        if (w4 == 0) TR(10, 4 + 3 * (g / TB_FIN_ROWS));
        float4 xv[TB_FIN_ROWS];
#pragma unroll
        for (int u = 0; u < TB_FIN_ROWS; ++u) {
          const int lbl = __shfl_sync(0xffffffffu, my_lab, g + u);
          xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (lbl >= 0 && lane * 4 < d) xv[u] = *reinterpret_cast<const float4*>(X + (r0 + g + u) * ld + lane * 4);
        }
        if (w4 == 0) TR(10, 5 + 3 * (g / TB_FIN_ROWS));
#pragma unroll
        for (int u = 0; u < TB_FIN_ROWS; ++u) {
          const int lbl = __shfl_sync(0xffffffffu, my_lab, g + u);
          if (lbl >= 0 && lane * 4 < d) red_add_v4(myrep + (size_t)lbl * d + lane * 4, xv[u]);
        }
        if (w4 == 0) TR(10, 6 + 3 * (g / TB_FIN_ROWS));
      }
      if (my_lab >= 0) {
        cost += (double)my_sc;
        // (measured at k = 256, 40M rows: one global count atomic per row cost 2.7 ms of a 10.2 ms step -- 256 hot
        // addresses -- against 1.6 ms for the row sums themselves)
        if (cnt_smem) atomicAdd(&cnt_s[my_lab], 1);
        else atomicAdd(cnt + my_lab, 1ull);
        ++nrows;
      }
      if (w4 == 0) TR(10, 3);
    }
    flush_pending();
    if (H16) bulk_wait_all();  // the reductions read this CTA's shared memory
    if (accumulate) {
      cost = warp_sum(cost);
      unsigned long long tot = nrows;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
      if (lane == 0) {
        atomicAdd(payload + (size_t)k * d + k, cost);
        atomicAdd(payload + (size_t)k * d + k + 1, (double)tot);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (accumulate && cnt_smem)
    for (int i = threadIdx.x; i < k; i += TB_THREADS) {
      const int c = cnt_s[i];
      if (c) atomicAdd(cnt + i, (unsigned long long)c);
    }
  if (warp == W_ALLOC) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TB_TMEM_COLS)
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// re-check plumbing: gather listed rows into a compact matrix, scatter their labels back
// ---------------------------------------------------------------------------------------------
__global__ void tcb_gather_kernel(const float* __restrict__ X, int64_t ld, int d, const float* __restrict__ colscale,
                                  TcbWs* __restrict__ ws, const int32_t* __restrict__ amb_idx,
                                  float* __restrict__ Xamb, int cap) {
  const int used = min(ws->count, cap);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ws->used = used;
    ws->cap = cap;
  }
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int i = warp; i < used; i += nwarps) {
    const float* src = X + (int64_t)amb_idx[i] * ld;
    float* dst = Xamb + (int64_t)i * d;
    for (int c = lane * 4; c < d; c += 128) {
      float4 v = *reinterpret_cast<const float4*>(src + c);
      if (colscale) {  // scaled pass: the compact matrix holds fl32(x * scale)
        const float4 sc = *reinterpret_cast<const float4*>(colscale + c);
        v.x = __fmul_rn(v.x, sc.x); v.y = __fmul_rn(v.y, sc.y); v.z = __fmul_rn(v.z, sc.z); v.w = __fmul_rn(v.w, sc.w);
      }
      *reinterpret_cast<float4*>(dst + c) = v;
    }
  }
}

__global__ void tcb_scatter_kernel(const TcbWs* __restrict__ ws, const int32_t* __restrict__ amb_idx,
                                   const int32_t* __restrict__ lab_amb, int32_t* __restrict__ labels) {
  const int used = ws->used;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < used; i += gridDim.x * blockDim.x)
    labels[amb_idx[i]] = lab_amb[i];
}

__global__ void tcb_set_cap_kernel(TcbWs* __restrict__ ws, int cap) { ws->cap = cap; }

// ---------------------------------------------------------------------------------------------
// Pair check: rows of the fp16 filter whose label is uncertain but can only be their best or second-best centre.
// fp64 on the fp64 master centres, the same form as km_exact_kernel (||x||^2 + ||c||^2 - 2 x.c, clamped at 0), equal
// distances go to the lower index (findClosest order).  Warp per row, lane = 4 columns (d <= 128); the centre table
// (k * d * 8 bytes) stays in L2.  Accumulates like the tensor-core kernels: 16-byte vector reductions into the fp32
// replicas, integer counts, fp64 cost.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) km_pair_exact_kernel(const float* __restrict__ X, int64_t ld, int d,
                                                            const float* __restrict__ colscale,
                                                            const double* __restrict__ C64,
                                                            const double* __restrict__ CN64,
                                                            const float* __restrict__ C32, int k,
                                                            double* __restrict__ payload, float* __restrict__ rep,
                                                            unsigned long long* __restrict__ cnt,
                                                            int32_t* __restrict__ labels,
                                                            const TcbWs* __restrict__ ws,
                                                            const int4* __restrict__ pairs, int pcap, int accumulate) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const int used = min(ws->pcount, pcap);
  const int c = lane * 4;
  const bool on = c < d;
  float* myrep = rep ? rep + (size_t)(blockIdx.x % KM_REPLICAS) * k * d : nullptr;
  double cost = 0.0;
  unsigned long long rows = 0;
  // One warp per entry, ~110 entries per warp, and each is a chain of dependent round trips (entry -> row from HBM
  // -> centres from L2): the entry after next and the next entry's row are fetched while the current one is worked on.
  const int4 none = make_int4(0, 0, 0, 0);
  auto load_row = [&](int64_t r) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (on) v = ldg_stream(reinterpret_cast<const float4*>(X + r * ld + c));
    return v;
  };
  int4 pr = warp < used ? pairs[warp] : none;
  int4 pr1 = warp + nwarps < used ? pairs[warp + nwarps] : none;
  float4 xraw = load_row(pr.x), xraw1 = make_float4(0.f, 0.f, 0.f, 0.f);
  int4 pr2 = none;
  for (int e = warp; e < used; e += nwarps, pr = pr1, pr1 = pr2, xraw = xraw1) {
    pr2 = (int64_t)e + 2 * (int64_t)nwarps < used ? pairs[e + 2 * nwarps] : none;
    xraw1 = load_row(pr1.x);  // (row 0 when there is no next entry: a harmless read)
    const int64_t row = pr.x;
    // candidates of the row: TB_GRID: up to four, 16 bits each in (y, z), their number in w; else the best two
    int cand[4];
    int nc;
    if (TB_GRID) {
      cand[0] = pr.y & 0xffff; cand[1] = (int)((uint32_t)pr.y >> 16);
      cand[2] = pr.z & 0xffff; cand[3] = (int)((uint32_t)pr.z >> 16);
      nc = min(max(pr.w, 1), 4);
    } else {
      cand[0] = pr.y; cand[1] = pr.z; cand[2] = pr.y; cand[3] = pr.y;
      nc = 2;
    }
    float4 x = xraw;
    if (on) {
      if (colscale) {
        const float4 sc = *reinterpret_cast<const float4*>(colscale + c);
        x.x = __fmul_rn(x.x, sc.x); x.y = __fmul_rn(x.y, sc.y); x.z = __fmul_rn(x.z, sc.z); x.w = __fmul_rn(x.w, sc.w);
      }
    }
    // First in fp32, sum (x - c32)^2 on the fp32 centres (half the bytes of the fp64 table, and this kernel is bound
    // by those L2 reads): relative error <= 12 * 2^-24 from the arithmetic (positive terms) and, from c32 = fl32(c),
    // 2 sqrt(dist2) 2^-24 ||c||; twice that is allowed for.  If the smallest distance beats every other by more than
    // both bounds the fp64 order is the same; else (about one entry in a few hundred: these rows were uncertain at
    // fp16 precision, not fp32) the fp64 evaluation below decides.
    if (C32 != nullptr) {
      float dq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q < nc) {
          float a = 0.f;
          if (on) {
            const float4 cv = *reinterpret_cast<const float4*>(C32 + (size_t)cand[q] * d + c);
            const float t0 = x.x - cv.x, t1 = x.y - cv.y, t2 = x.z - cv.z, t3 = x.w - cv.w;
            a = fmaf(t0, t0, fmaf(t1, t1, fmaf(t2, t2, t3 * t3)));
          }
          dq[q] = warp_sum(a);
        }
      }
      int qb = 0;
      float dbest = dq[0];
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (q < nc && dq[q] < dbest) {
          qb = q;
          dbest = dq[q];
        }
      bool sure = true;
      float hi_best = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q < nc) {
          const float cnq = (float)CN64[cand[q]];
          const float err = fmaf(2.0e-6f, dq[q], 3.0e-7f * sqrtf(dq[q] * cnq)) + 1e-30f;
          if (q == qb) hi_best = dq[q] + err;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (q < nc && q != qb) {
          const float cnq = (float)CN64[cand[q]];
          const float err = fmaf(2.0e-6f, dq[q], 3.0e-7f * sqrtf(dq[q] * cnq)) + 1e-30f;
          sure = sure && (dq[q] - err > hi_best);
        }
      }
      if (sure) {
        int pick32 = cand[0];
#pragma unroll
        for (int q = 1; q < 4; ++q) pick32 = (qb == q) ? cand[q] : pick32;
        if (lane == 0) labels[row] = pick32;
        if (accumulate) {
          if (on) red_add_v4(myrep + (size_t)pick32 * d + c, x);
          if (lane == 0) {
            atomicAdd(cnt + pick32, 1ull);
            cost += (double)dbest;
            ++rows;
          }
        }
        continue;
      }
    }
    const double xd[4] = {(double)x.x, (double)x.y, (double)x.z, (double)x.w};
    double xn = 0.0, sdot[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < 4; ++j) xn = fma(xd[j], xd[j], xn);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (q < nc && on) {  // (nc is the same for the whole warp)
        const double2* pc = reinterpret_cast<const double2*>(C64 + (size_t)cand[q] * d + c);
        const double2 u0 = pc[0], u1 = pc[1];
        sdot[q] = fma(xd[0], u0.x, fma(xd[1], u0.y, fma(xd[2], u1.x, xd[3] * u1.y)));
      }
    }
    xn = warp_sum(xn);
    int pick = cand[0];
    double vbest = 0.0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (q < nc) {
        const double sq = warp_sum(sdot[q]);
        const double vq = fmax(xn + CN64[cand[q]] - 2.0 * sq, 0.0);
        if (q == 0 || vq < vbest || (vq == vbest && cand[q] < pick)) {
          vbest = vq;
          pick = cand[q];
        }
      }
    }
    if (lane == 0) labels[row] = pick;
    if (accumulate) {
      if (on) red_add_v4(myrep + (size_t)pick * d + c, x);
      if (lane == 0) {
        atomicAdd(cnt + pick, 1ull);
        cost += vbest;
        ++rows;
      }
    }
  }
  if (accumulate && lane == 0 && rows) {
    atomicAdd(payload + (size_t)k * d + k, cost);
    atomicAdd(payload + (size_t)k * d + k + 1, (double)rows);
  }
}

// The same check with HALF a warp per entry (TB_GRID entries, d = 64 * NV): one warp per entry is a chain of dependent
// round trips (entry -> row from HBM -> centres from L2) with 32 lanes waiting on it, and at d = 64 half of them had no
// column at all.  Two entries per warp double the loads in flight and shorten the reductions to four shuffles.
// (ncu launch list at cfg 2: 2.9 ms per iteration for 5-8M entries with the warp-per-entry kernel.)
template <int NV>
__global__ void __launch_bounds__(256) km_cand_kernel(const float* __restrict__ X, int64_t ld, int d,
                                                      const float* __restrict__ colscale,
                                                      const double* __restrict__ C64, const double* __restrict__ CN64,
                                                      const float* __restrict__ C32, int k,
                                                      double* __restrict__ payload, float* __restrict__ rep,
                                                      unsigned long long* __restrict__ cnt, int32_t* __restrict__ labels,
                                                      const TcbWs* __restrict__ ws, const int4* __restrict__ pairs,
                                                      int pcap, int accumulate) {
  const int lane = threadIdx.x & 31, hl = lane & 15;
  const int hw = (blockIdx.x * blockDim.x + threadIdx.x) >> 4, nhw = (gridDim.x * blockDim.x) >> 4;
  const int used = min(ws->pcount, pcap);
  const int c0 = hl * 4;  // this lane's columns: c0 .. c0 + 3 (and 64 further on when NV == 2)
  float* myrep = rep ? rep + (size_t)(blockIdx.x % KM_REPLICAS) * k * d : nullptr;
  double cost = 0.0;
  unsigned long long rows = 0;
  const int4 none = make_int4(0, 0, 0, 0);
  float4 xr[NV], xr1[NV];
  int4 pr = hw < used ? pairs[hw] : none;
  int4 pr1 = hw + nhw < used ? pairs[hw + nhw] : none;
#pragma unroll
  for (int v = 0; v < NV; ++v) xr[v] = ldg_stream(reinterpret_cast<const float4*>(X + (int64_t)pr.x * ld + c0 + 64 * v));
  for (int e = hw; __any_sync(0xffffffffu, e < used); e += nhw) {
    const bool valid = e < used;
    const int4 pr2 = (int64_t)e + 2 * (int64_t)nhw < used ? pairs[e + 2 * nhw] : none;
#pragma unroll
    for (int v = 0; v < NV; ++v)  // (row 0 when there is no next entry: a harmless read)
      xr1[v] = ldg_stream(reinterpret_cast<const float4*>(X + (int64_t)pr1.x * ld + c0 + 64 * v));
    const int64_t row = pr.x;
    const int cand[4] = {pr.y & 0xffff, (int)((uint32_t)pr.y >> 16), pr.z & 0xffff, (int)((uint32_t)pr.z >> 16)};
    const int nc = valid ? min(max(pr.w, 1), 4) : 0;
    float4 x[NV];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      x[v] = xr[v];
      if (colscale) {
        const float4 sc = *reinterpret_cast<const float4*>(colscale + c0 + 64 * v);
        x[v].x = __fmul_rn(x[v].x, sc.x); x[v].y = __fmul_rn(x[v].y, sc.y);
        x[v].z = __fmul_rn(x[v].z, sc.z); x[v].w = __fmul_rn(x[v].w, sc.w);
      }
    }
    // fp32 first (see km_pair_exact_kernel for the bounds)
    float dq[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float a = 0.f;
      if (q < nc) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          const float4 cv = *reinterpret_cast<const float4*>(C32 + (size_t)cand[q] * d + c0 + 64 * v);
          const float t0 = x[v].x - cv.x, t1 = x[v].y - cv.y, t2 = x[v].z - cv.z, t3 = x[v].w - cv.w;
          a = fmaf(t0, t0, fmaf(t1, t1, fmaf(t2, t2, fmaf(t3, t3, a))));
        }
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      dq[q] = a;
    }
    int pick = cand[0];
    float dbest = dq[0], ebest = 0.f;
    bool sure = true;
#pragma unroll
    for (int q = 1; q < 4; ++q)
      if (q < nc && dq[q] < dbest) {
        dbest = dq[q];
        pick = cand[q];
      }
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (q < nc && cand[q] == pick)
        ebest = fmaf(2.0e-6f, dq[q], 3.0e-7f * sqrtf(dq[q] * (float)CN64[cand[q]])) + 1e-30f;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (q < nc && cand[q] != pick) {
        const float err = fmaf(2.0e-6f, dq[q], 3.0e-7f * sqrtf(dq[q] * (float)CN64[cand[q]])) + 1e-30f;
        sure = sure && (dq[q] - err > dbest + ebest);
      }
    double vbest = (double)dbest;
    const bool need64 = valid && !sure;
    if (__any_sync(0xffffffffu, need64)) {  // fp64 on the master centres, same form as km_exact_kernel
      double xn = 0.0;
#pragma unroll
      for (int v = 0; v < NV; ++v)
        xn = fma((double)x[v].x, (double)x[v].x, fma((double)x[v].y, (double)x[v].y,
                 fma((double)x[v].z, (double)x[v].z, fma((double)x[v].w, (double)x[v].w, xn))));
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) xn += __shfl_xor_sync(0xffffffffu, xn, o);
      int pick64 = cand[0];
      double v64 = 0.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        double sdot = 0.0;
        if (need64 && q < nc) {
#pragma unroll
          for (int v = 0; v < NV; ++v) {
            const double2* pc = reinterpret_cast<const double2*>(C64 + (size_t)cand[q] * d + c0 + 64 * v);
            const double2 u0 = pc[0], u1 = pc[1];
            sdot = fma((double)x[v].x, u0.x, fma((double)x[v].y, u0.y, fma((double)x[v].z, u1.x, fma((double)x[v].w, u1.y, sdot))));
          }
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) sdot += __shfl_xor_sync(0xffffffffu, sdot, o);
        if (need64 && q < nc) {
          const double vq = fmax(xn + CN64[cand[q]] - 2.0 * sdot, 0.0);
          if (q == 0 || vq < v64 || (vq == v64 && cand[q] < pick64)) {  // lower index wins ties
            v64 = vq;
            pick64 = cand[q];
          }
        }
      }
      if (need64) {
        pick = pick64;
        vbest = v64;
      }
    }
    if (valid) {
      if (hl == 0) labels[row] = pick;
      if (accumulate) {
#pragma unroll
        for (int v = 0; v < NV; ++v) red_add_v4(myrep + (size_t)pick * d + c0 + 64 * v, x[v]);
        if (hl == 0) {
          atomicAdd(cnt + pick, 1ull);
          cost += vbest;
          ++rows;
        }
      }
    }
    pr = pr1;
    pr1 = pr2;
#pragma unroll
    for (int v = 0; v < NV; ++v) xr[v] = xr1[v];
  }
  if (accumulate && hl == 0 && rows) {
    atomicAdd(payload + (size_t)k * d + k, cost);
    atomicAdd(payload + (size_t)k * d + k + 1, (double)rows);
  }
}

// Centre-dependent part of the cost of everything accumulated so far:
//   F = sum_j ( n_j ||c_j||^2 - 2 c_j . S_j ),   S_j / n_j = fp64 payload + fp32 replicas / integer counts.
// The fp16 filter kernel adds only sum ||x||^2 of its certain rows to the cost; the exact remainder of their cost is
// F(after the kernel) - F(before it), whatever the accumulator held on entry.  One warp per centre, fp64.
__global__ void tcb_cost_terms_kernel(double* __restrict__ payload, const float* __restrict__ rep,
                                      const unsigned long long* __restrict__ cnt, const double* __restrict__ c64,
                                      int k, int d, double sign) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const size_t kd = (size_t)k * d;
  double f = 0.0;
  for (int j = warp; j < k; j += nwarps) {
    double dot = 0.0, nn = 0.0;
    for (int c = lane; c < d; c += 32) {
      const size_t i = (size_t)j * d + c;
      double sum = payload[i];
#pragma unroll
      for (int rp = 0; rp < KM_REPLICAS; ++rp) sum += (double)rep[rp * kd + i];
      const double cv = c64[i];
      dot += cv * sum;
      nn += cv * cv;
    }
    dot = warp_sum(dot);
    nn = warp_sum(nn);
    f += (payload[kd + j] + (double)cnt[j]) * nn - 2.0 * dot;
  }
  if (lane == 0 && f != 0.0) atomicAdd(payload + kd + k, sign * f);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int km_fold_scaled(void* acc, int k, int d, const float* colscale, cudaStream_t stream);  // kmeans_generic.cu

struct TcbWsLayout {
  size_t off_idx, off_lab, off_idx2, off_pairs, off_x, bytes;
  int cap;
};
static TcbWsLayout tcb_ws_layout(int cap, int d) {
  TcbWsLayout L;
  L.cap = cap;
  size_t o = sizeof(TcbWs);
  L.off_idx = o; o = b2f_align_up(o + (size_t)cap * 4, 1024);
  L.off_lab = o; o = b2f_align_up(o + (size_t)cap * 4, 1024);
  L.off_idx2 = o; o = b2f_align_up(o + (size_t)cap * 4, 1024);
  L.off_pairs = o; o = b2f_align_up(o + (size_t)cap * 16, 1024);
  L.off_x = o;   o = b2f_align_up(o + (size_t)cap * d * 4, 1024);
  L.bytes = o;
  return L;
}
static int tcb_cap_for_rows(int64_t n) {
  int64_t c = (n / 5 + TM - 1) / TM * TM;  // room for 1 row in 5 (cfg 2 lists 5-9.5 % with random initial centres)
  if (c < 4096) c = 4096;
  if (c > (1 << 30)) c = 1 << 30;
  return (int)c;
}
static int tcb_cap_from_bytes(size_t bytes, int d) {
  if (bytes < sizeof(TcbWs) + 5 * 1024) return 0;
  int64_t c = (int64_t)((bytes - sizeof(TcbWs) - 5 * 1024) / (28 + (size_t)d * 4));
  c = c / TM * TM;
  if (c > (1 << 30)) c = 1 << 30;
  while (c > 0 && tcb_ws_layout((int)c, d).bytes > bytes) c -= TM;
  return (int)c;
}

static int tcb_tile_bytes(int d, int filter) {
  return filter == KM_FILTER_FP16 ? (d / 64) * CHUNK_BYTES + TB_CN_BYTES : (d / 64) * tb_stage_bytes<KM_FILTER_BF16X3>();
}
static bool tcb_rfs(int d, int filter) {
#ifdef B2F_DEV
  if (getenv("B2F_TC_NO_RFS")) return false;
#endif
  return filter == KM_FILTER_FP16 && d == 64;
}
static int km_tcb_stages(int k, int d, int filter) {
  const bool rfs = tcb_rfs(d, filter);
  // row slots of the finish warps (fp16 filter; none when they reduce from the staging ring)
  const int fx_bytes = rfs ? 0 : TB_FX_WARPS * 2 * TB_FX_ROWS * d * 4;
  int budget_kb = 227, cap = 4;
#ifdef B2F_DEV  // dev knobs of the timing experiments; not in the product build
  if (filter == KM_FILTER_FP16 && getenv("B2F_TC_SMEM_KB")) budget_kb = atoi(getenv("B2F_TC_SMEM_KB"));
  if (getenv("B2F_TC_STAGES")) cap = atoi(getenv("B2F_TC_STAGES"));
#endif
  const int left = budget_kb * 1024 - (int)sizeof(SmemB) - tcb_cnt_smem_bytes(k) - 64 - (rfs ? TB_NX : 1) * (d / 32) * CHUNK_BYTES -
                   (filter == KM_FILTER_FP16 ? TB_CN_BYTES + fx_bytes : 0);
  // a stage holds one centre tile: d / 64 atoms, fp16: + the ||c||^2 block
  int s = left / tcb_tile_bytes(d, filter);
  return s > cap ? cap : s;
}

bool km_tcb_supported(int k, int d, int64_t ld) {
  if (d != 64 && d != 128) return false;  // whole 64-k atoms, X hi/lo pairs fit the tensor memory budget
  if (k % TN != 0 || k < TN || k > 65536) return false;  // (the filter hands centre indices over in 16 bits)
  if (ld % 4 != 0) return false;
  return km_tcb_stages(k, d, KM_FILTER_BF16X3) >= 2 && km_tcb_stages(k, d, KM_FILTER_FP16) >= 2 &&
         km_tc_supported(k, d, ld);
}

size_t km_tcb_ws_bytes(int64_t n, int k, int d) {
  (void)k;
  return tcb_ws_layout(tcb_cap_for_rows(n), d).bytes;
}

template <int MODE, bool SCALED, bool RFS>
static int tcb_launch_filter(int grid, size_t smem, cudaStream_t stream, const CUtensorMap& mx, const CUtensorMap& m1,
                             const CUtensorMap& m2, const CUtensorMap& mr, const float* X, int64_t n, int d, int64_t ld, const float* colscale,
                             const float* cn, const float* misc, float kappa2, float kappa0, int k, int nstages,
                             double* payload, float* rep, unsigned long long* cnt, int32_t* labels, int accumulate,
                             TcbWs* hdr, int32_t* amb_idx, int cap, int4* pairs, int pcap) {
  unsigned long long* trace = nullptr;
#ifdef B2F_DEV  // hand-off tracer of the timing experiments (allocates; not in the product build)
  const char* trace_path = getenv("B2F_TC_TRACE");
  const size_t trace_bytes = (size_t)TR_ROLES * TR_TILES * TR_SLOTS * sizeof(unsigned long long);
  if (trace_path) {
    B2F_CUDA(cudaMalloc(&trace, trace_bytes));
    B2F_CUDA(cudaMemsetAsync(trace, 0, trace_bytes, stream));
    auto kern = km_tcb_kernel<MODE, true, SCALED, RFS>;
    B2F_CUDA(b2f_smem_attr(kern, smem));
    kern<<<grid, TB_THREADS, smem, stream>>>(mx, m1, m2, mr, X, n, d, ld, colscale, cn, misc, kappa2, kappa0, k, nstages,
                                             payload, rep, cnt, labels, accumulate, hdr, amb_idx, cap, pairs, pcap, trace);
    B2F_LAUNCH_CHECK();
    std::vector<unsigned long long> host(trace_bytes / sizeof(unsigned long long));
    B2F_CUDA(cudaStreamSynchronize(stream));
    B2F_CUDA(cudaMemcpy(host.data(), trace, trace_bytes, cudaMemcpyDeviceToHost));
    B2F_CUDA(cudaFree(trace));
    if (FILE* f = fopen(trace_path, "wb")) {
      fwrite(host.data(), 1, trace_bytes, f);
      fclose(f);
    }
    return B2F_OK;
  }
#endif
  auto kern = km_tcb_kernel<MODE, false, SCALED, RFS>;
  B2F_CUDA(b2f_smem_attr(kern, smem));
  kern<<<grid, TB_THREADS, smem, stream>>>(mx, m1, m2, mr, X, n, d, ld, colscale, cn, misc, kappa2, kappa0, k, nstages,
                                           payload, rep, cnt, labels, accumulate, hdr, amb_idx, cap, pairs, pcap, trace);
  B2F_LAUNCH_CHECK();
  return B2F_OK;
}

int km_tcb_step(const float* X, int64_t n, int d, int64_t ld, const float* colscale, const void* prep, int k, void* acc,
                int32_t* labels, void* ws, size_t ws_bytes, bool accumulate, int filter, cudaStream_t stream) {
  B2F_REQUIRE(labels, "b2f_kmeans (filter path): a labels buffer is required");
  B2F_REQUIRE(ws, "b2f_kmeans (filter path): a workspace is required (b2f_kmeans_ws_bytes)");
  B2F_REQUIRE(n < (1ll << 31), "b2f_kmeans (filter path): at most 2^31-1 rows per call");
  B2F_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "b2f_kmeans (filter path): workspace must be 256-byte aligned");
  B2F_REQUIRE(!colscale || filter == KM_FILTER_FP16, "b2f_kmeans (filter path): a column scale needs the fp16 filter");
  B2F_REQUIRE(!colscale || (reinterpret_cast<uintptr_t>(colscale) & 15) == 0,
              "b2f_kmeans (filter path): the column scale must be 16-byte aligned");
  char* wp = static_cast<char*>(ws);
  TcbWs* hdr = reinterpret_cast<TcbWs*>(wp);
  KmPrepLayout P = km_prep_layout(k, d);
  const char* pp = static_cast<const char*>(prep);
  const float* cn = reinterpret_cast<const float*>(pp + P.off_cn);
  const float* misc = reinterpret_cast<const float*>(pp + P.off_misc);
  const float* cmax = misc + KM_MISC_CMAX;
  // 3xTF32 bound: dropped products 3 u^2 with u = 2^-11 on c' = -2c, plus 48 fp32 accumulation roundings
  const float kappa2_tc3 = 2.0f * (6.0f * 1.02f / 4194304.0f + 2.0f * (3.0f * d / 8.0f + 16.0f) / 16777216.0f);
  B2F_CUDA(cudaMemsetAsync(hdr, 0, sizeof(TcbWs), stream));

  if (filter == KM_FILTER_NONE) {
    // No filter pass: 3xTF32 over every row, rows inside its own (narrow) uncertainty band to the fp64 kernel.
    // The whole workspace behind the header is the list: room for every row when the workspace has its standard size.
    B2F_REQUIRE(ws_bytes >= sizeof(TcbWs) + 4096, "b2f_kmeans (3xTF32 + re-check): workspace of %zu bytes is too small",
                ws_bytes);
    int64_t cap3 = (int64_t)((ws_bytes - sizeof(TcbWs)) / 4);
    if (cap3 > (1 << 30)) cap3 = 1 << 30;
    int32_t* idx3 = reinterpret_cast<int32_t*>(wp + sizeof(TcbWs));
    tcb_set_cap_kernel<<<1, 1, 0, stream>>>(hdr, (int)cap3);
    B2F_LAUNCH_CHECK();
    KmRecheck rck;
    rck.count = &hdr->count2;
    rck.idx = idx3;
    rck.cap = (int)cap3;
    rck.src_idx = nullptr;
    rck.kappa2 = kappa2_tc3;
    rck.cmax = cmax;
    int rc = km_tc_step(X, n, d, ld, prep, k, acc, labels, 3, accumulate, nullptr, &rck, stream);
    if (rc != B2F_OK) return rc;
    return km_exact_rows(X, ld, d, nullptr, prep, k, acc, labels, &hdr->count2, idx3, (int)cap3, accumulate, stream);
  }

  const int cap = tcb_cap_from_bytes(ws_bytes, d);
  B2F_REQUIRE(cap >= TM, "b2f_kmeans (filter path): workspace of %zu bytes is too small", ws_bytes);
  TcbWsLayout W = tcb_ws_layout(cap, d);
  int32_t* amb_idx = reinterpret_cast<int32_t*>(wp + W.off_idx);
  int32_t* lab_amb = reinterpret_cast<int32_t*>(wp + W.off_lab);
  int32_t* idx2 = reinterpret_cast<int32_t*>(wp + W.off_idx2);
  int4* pairs = reinterpret_cast<int4*>(wp + W.off_pairs);
  float* Xamb = reinterpret_cast<float*>(wp + W.off_x);
  const bool h16 = filter == KM_FILTER_FP16;
  const bool rfs = tcb_rfs(d, filter);
  const int fx_bytes = rfs ? 0 : TB_FX_WARPS * 2 * TB_FX_ROWS * d * 4;
  const int nstages = km_tcb_stages(k, d, filter);
  B2F_REQUIRE(nstages >= 2, "b2f_kmeans (filter path): k = %d, d = %d does not fit in shared memory", k, d);
  CUtensorMap mx, m1, m2;
  int rc;
  if ((rc = make_map_f32(&mx, X, (uint64_t)n, (uint64_t)d, (uint64_t)ld)) != B2F_OK) return rc;
  // (the fp16 plane goes through a 16-bit map as well: the copy does not interpret the elements)
  if ((rc = make_map_bf16(&m1, pp + (h16 ? P.off_ch16 : P.off_cb1), (uint64_t)k, (uint64_t)d, (uint64_t)d)) != B2F_OK)
    return rc;
  // second map: bf16x3 -> the low bf16 plane; fp16 -> the ||c||^2 plane (64 halves per centre)
  if (h16) rc = make_map_b16_sw32(&m2, pp + P.off_cnh, (uint64_t)k, (uint64_t)KM_CNH_COLS, (uint64_t)KM_CNH_COLS);
  else rc = make_map_bf16(&m2, pp + P.off_cb2, (uint64_t)k, (uint64_t)d, (uint64_t)d);
  if (rc != B2F_OK) return rc;
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
  // 2 * tau = kappa2 * ||x|| * max||c||:  dropped products 3 u^2 (u = 2^-8) on c' = -2c, i.e. 6 * 2^-16,
  // plus fp32 accumulation / rounding slack (3 d + 16) * 2^-23 on |x||c'| = 2 |x||c|
  // fp16 filter (one product, x and -2c each rounded once, u = 2^-11):  |error| <= (2u + u^2) |x||c'| = 4.01 u |x||c|
  // plus the accumulation slack for d products; operands below the fp16 normal range carry an absolute error of
  // eta = 2^-25 / S per entry instead:  + sqrt(d) eta (|c'| + |x|)  ->  kappa0 (divided by S in the kernel)
  const float kappa = h16 ? 4.0f * 1.01f / 2048.0f + 2.0f * (d + 20.0f) / 8388608.0f
                          : 6.0f * 1.02f / 65536.0f + 2.0f * (3.0f * d + 16.0f) / 8388608.0f;
  const float kappa2 = 2.0f * kappa;
  const float kappa0 = h16 ? 2.0f * 1.01f * sqrtf((float)d) / 33554432.0f : 0.f;
  bool acc_main = accumulate;
#ifdef B2F_DEV
  if (getenv("B2F_TC_DBG") && (atoi(getenv("B2F_TC_DBG")) & 4)) acc_main = false;  // results invalid when set
#endif
  // RFS: the replicas as ONE [KM_REPLICAS * k, d] tensor, boxes of one row x 32 floats in the staging swizzle
  CUtensorMap mr = mx;
  if (rfs && rep &&
      (rc = make_map_2d(&mr, rep, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)KM_REPLICAS * k, (uint64_t)d, (uint64_t)d,
                        1)) != B2F_OK)
    return rc;
  const size_t smem = (size_t)(rfs ? TB_NX : 1) * (d / 32) * CHUNK_BYTES + (size_t)nstages * tcb_tile_bytes(d, filter) +
                      (h16 ? TB_CN_BYTES + fx_bytes : 0) + sizeof(SmemB) + tcb_cnt_smem_bytes(k) + 64;
  const int64_t ntiles = (n + TM - 1) / TM;
  int grid = b2f_sm_count();
  if ((int64_t)grid > ntiles) grid = (int)ntiles;
  if (h16) {
    const double* c64 = reinterpret_cast<const double*>(pp + P.off_c64);
    const double* cn64 = reinterpret_cast<const double*>(pp + P.off_cn64);
    const float* c32 = TB_GRID ? reinterpret_cast<const float*>(pp + P.off_c32) : nullptr;
    const int cgrid = (k + 7) / 8;
    // scaled pass: this kernel adds RAW rows to the replicas; whatever they hold on entry is folded away first and
    // the raw sums are folded through the scale right after the kernel, so everything downstream sees scaled sums
    if (acc_main && colscale && (rc = km_fold_scaled(acc, k, d, nullptr, stream)) != B2F_OK) return rc;
    if (acc_main) tcb_cost_terms_kernel<<<cgrid, 256, 0, stream>>>(payload, rep, cnt, c64, k, d, -1.0);
#define TCB_H16(SC, RF)                                                                                              \
  tcb_launch_filter<KM_FILTER_FP16, SC, RF>(grid, smem, stream, mx, m1, m2, mr, X, n, d, ld, colscale, cn, misc, kappa2, \
                                            kappa0, k, nstages, payload, rep, cnt, labels, acc_main ? 1 : 0, hdr,     \
                                            amb_idx, cap, pairs, cap)
    if (colscale) rc = rfs ? TCB_H16(true, true) : TCB_H16(true, false);
    else rc = rfs ? TCB_H16(false, true) : TCB_H16(false, false);
#undef TCB_H16
    if (rc != B2F_OK) return rc;
    if (acc_main && colscale && (rc = km_fold_scaled(acc, k, d, colscale, stream)) != B2F_OK) return rc;
    if (acc_main) tcb_cost_terms_kernel<<<cgrid, 256, 0, stream>>>(payload, rep, cnt, c64, k, d, 1.0);
    // rows with exactly two candidate centres: two fp64 distances each
    if (TB_GRID && d == 64)
      km_cand_kernel<1><<<b2f_sm_count() * 8, 256, 0, stream>>>(X, ld, d, colscale, c64, cn64, c32, k, payload, rep, cnt,
                                                                labels, hdr, pairs, cap, accumulate ? 1 : 0);
    else if (TB_GRID && d == 128)
      km_cand_kernel<2><<<b2f_sm_count() * 8, 256, 0, stream>>>(X, ld, d, colscale, c64, cn64, c32, k, payload, rep, cnt,
                                                                labels, hdr, pairs, cap, accumulate ? 1 : 0);
    else
      km_pair_exact_kernel<<<b2f_sm_count() * 8, 256, 0, stream>>>(X, ld, d, colscale, c64, cn64, c32, k, payload, rep,
                                                                   cnt, labels, hdr, pairs, cap, accumulate ? 1 : 0);
    B2F_LAUNCH_CHECK();
  } else {
    rc = tcb_launch_filter<KM_FILTER_BF16X3, false, false>(grid, smem, stream, mx, m1, m2, mr, X, n, d, ld, nullptr, cn, misc, kappa2,
                                                    kappa0, k, nstages, payload, rep, cnt, labels, acc_main ? 1 : 0, hdr,
                                                    amb_idx, cap, nullptr, 0);
    if (rc != B2F_OK) return rc;
  }
  tcb_gather_kernel<<<b2f_sm_count() * 4, 256, 0, stream>>>(X, ld, d, colscale, hdr, amb_idx, Xamb, cap);
  B2F_LAUNCH_CHECK();
  KmRecheck rck;
  rck.count = &hdr->count2;
  rck.idx = idx2;
  rck.cap = cap;
  rck.src_idx = amb_idx;
  rck.kappa2 = kappa2_tc3;
  rck.cmax = cmax;
  rc = km_tc_step(Xamb, (int64_t)cap, d, (int64_t)d, prep, k, acc, lab_amb, 3, accumulate, &hdr->used, &rck, stream);
  if (rc != B2F_OK) return rc;
  tcb_scatter_kernel<<<b2f_sm_count() * 2, 256, 0, stream>>>(hdr, amb_idx, lab_amb, labels);
  B2F_LAUNCH_CHECK();
  // rows still uncertain after 3xTF32: exact fp64 assignment + accumulation (writes labels after the scatter)
  return km_exact_rows(X, ld, d, colscale, prep, k, acc, labels, &hdr->count2, idx2, cap, accumulate, stream);
}
