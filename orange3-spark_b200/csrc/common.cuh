// Shared helpers for libb200fit.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "b200fit.h"

#ifndef __CUDA_ARCH__
#define B2F_HOST 1
#endif

void b2f_set_error(const char* fmt, ...);

#define B2F_REQUIRE(cond, ...)        \
  do {                                \
    if (!(cond)) {                    \
      b2f_set_error(__VA_ARGS__);     \
      return B2F_EINVAL;              \
    }                                 \
  } while (0)

#define B2F_CUDA(call)                                                                  \
  do {                                                                                  \
    cudaError_t _e = (call);                                                            \
    if (_e != cudaSuccess) {                                                            \
      b2f_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
      return B2F_ECUDA;                                                                 \
    }                                                                                   \
  } while (0)

#define B2F_LAUNCH_CHECK() B2F_CUDA(cudaGetLastError())

static inline cudaStream_t b2f_cs(b2f_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// number of SMs of the current device (cached per device)
int b2f_sm_count();

// cudaFuncAttributeMaxDynamicSharedMemorySize, set once per (kernel, device) and again only when a launch needs
// more (api.cu keeps the table; the attribute call is not free and sat on every step before)
cudaError_t b2f_smem_attr_impl(const void* kernel, size_t bytes);
template <typename F>
static inline cudaError_t b2f_smem_attr(F kernel, size_t bytes) {
  return b2f_smem_attr_impl(reinterpret_cast<const void*>(kernel), bytes);
}

static inline size_t b2f_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- kmeans buffer layouts (shared by kmeans_generic.cu / kmeans_tc.cu) ----
// prep:  [ C32  float[k*d]   fp32 centres (rounded from the fp64 master)
//          CH   float[k*d]   -2*c rounded to TF32 (low 13 mantissa bits zero)
//          CL   float[k*d]   -2*c - CH, rounded to TF32
//          CN   float[k]     ||c||^2 (fp32, from fp64)
//          C64  double[k*d]  copy of the fp64 master (refine / exact paths)
//          CB1  bf16[k*d]    -2*c rounded to bf16          (operands of the bf16x3 kernel)
//          CB2  bf16[k*d]    -2*c - CB1, rounded to bf16
//          CH16 fp16[k*d]    -2*c * S rounded to fp16, S a power of two (operand of the fp16 filter kernel)
//          CN64 double[k]    ||c||^2 in fp64 (exact re-check)
//          CNH  fp16[k*16]   ||c||^2 * S^2 / 2^15 as three fp16 pieces in columns 0..2, zeros elsewhere: one more
//                            k-block of the fp16 filter's centre operand (its x side is the constant 2^15), so the
//                            tensor core adds ||c||^2 and the epilogue does not have to
//          MISC float[64]    [0] = max_j ||c_j|| rounded up   [1] = max |c_jl|   (float bits, atomicMax)
//                            [2] = S_a, the scale of the x operand   [3] = 1 / (S_a S_b)   [4] = S_b, the scale of
//                            the centre operand   [5] = max ||x|| the planes were built for (0: unknown)
//                            [6] = global key scale sc = S_a S_b (0: per-row key scales)
//                            (written by the second prepare kernel)                                      ]
// fp16 filter, two modes.  Per-row (max ||x|| unknown): S_a = S_b = S, the accumulator holds S^2 * score and the
// epilogue scales every score into its row's 22-bit key window (one multiply-add per score).  Global (max ||x||
// given, b2f_kmeans_prepare_x): S_a S_b = sc is chosen so that sc * |score| < 2^21 for EVERY row, and the same
// k-block that adds ||c||^2 also adds the magic constant 1.5 * 2^23 (A side 3072, centre side 4096 in column 3 of
// CNH): the accumulator then holds the key's integer directly and the multiply-add disappears from the epilogue.
#define KM_CNH_COLS 16
#define KM_CNH_E1 32768.f
#define KM_CNH_MA 3072.f   // 1.5 * 2^11: x-side half of the magic constant (always present in the constant block)
#define KM_CNH_MB 4096.f   // centre-side half (column 3 of CNH; written in global mode only)
#define KM_MISC_CMAX 0
#define KM_MISC_CABS 1
#define KM_MISC_SCALE 2
#define KM_MISC_INV_SCALE2 3
#define KM_MISC_SCALE_B 4
#define KM_MISC_XMAX 5
#define KM_MISC_SC 6
struct KmPrepLayout {
  size_t off_c32, off_ch, off_cl, off_cn, off_c64, off_cb1, off_cb2, off_ch16, off_cn64, off_cnh, off_misc, bytes;
};
static inline KmPrepLayout km_prep_layout(int k, int d) {
  KmPrepLayout L;
  size_t kd = (size_t)k * d;
  size_t o = 0;
  L.off_c32 = o; o = b2f_align_up(o + kd * 4, 1024);
  L.off_ch = o;  o = b2f_align_up(o + kd * 4, 1024);
  L.off_cl = o;  o = b2f_align_up(o + kd * 4, 1024);
  L.off_cn = o;  o = b2f_align_up(o + (size_t)k * 4, 1024);
  L.off_c64 = o; o = b2f_align_up(o + kd * 8, 1024);
  L.off_cb1 = o; o = b2f_align_up(o + kd * 2, 1024);
  L.off_cb2 = o; o = b2f_align_up(o + kd * 2, 1024);
  L.off_ch16 = o; o = b2f_align_up(o + kd * 2, 1024);
  L.off_cn64 = o; o = b2f_align_up(o + (size_t)k * 8, 1024);
  L.off_cnh = o;  o = b2f_align_up(o + (size_t)k * KM_CNH_COLS * 2, 1024);
  L.off_misc = o; o = b2f_align_up(o + 256, 1024);
  L.bytes = o;
  return L;
}

// acc:   [ payload double[k*d + k + 2] = sums | counts | cost | nrows
//          rep     float[KM_REPLICAS * k * d]   fp32 replicas written with red.global.add
//          cnt32   unsigned long long[k]        integer counts (unweighted fast path)        ]
#ifndef KM_REPLICAS
#define KM_REPLICAS 4
#endif
struct KmAccLayout {
  size_t off_payload, off_rep, off_cnt, bytes;
  size_t payload_len;
};
static inline KmAccLayout km_acc_layout(int k, int d) {
  KmAccLayout L;
  size_t kd = (size_t)k * d;
  L.payload_len = kd + k + 2;
  size_t o = 0;
  L.off_payload = o; o = b2f_align_up(o + L.payload_len * 8, 1024);
  L.off_rep = o;     o = b2f_align_up(o + (size_t)KM_REPLICAS * kd * 4, 1024);
  L.off_cnt = o;     o = b2f_align_up(o + (size_t)k * 8, 1024);
  L.bytes = o;
  return L;
}

#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// round-to-nearest fp32 -> tf32 (result is an fp32 bit pattern with the low 13 bits zero)
__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
#endif

// Optional certainty test of the 3xTF32 kernel: rows whose best / second-best scores are within
// kappa2 * ||x|| * max||c|| are not accumulated but appended (as ORIGINAL row indices, through src_idx when
// the kernel runs on a gathered matrix) to idx[0..cap) for the exact fp64 kernel.  count == NULL: off.
struct KmRecheck {
  int* count;
  int32_t* idx;
  int cap;
  const int32_t* src_idx;
  float kappa2;
  const float* cmax;
};

// implemented in kmeans_generic.cu: exact (fp64) assignment + accumulation of listed rows.
// colscale (optional, fp32[d]): the rows are taken as fl32(x * colscale) -- the StandardScaler fold.
int km_exact_rows(const float* X, int64_t ld, int d, const float* colscale, const void* prep, int k, void* acc,
                  int32_t* labels, const int* count, const int32_t* idx, int cap, bool accumulate,
                  cudaStream_t stream);
// squared distance (fp32, direct form on the fp32 centres) of every row to the centre its label names
int km_dist_to_label(const float* X, int64_t n, int d, int64_t ld, const float* colscale, const void* prep, int k,
                     const int32_t* labels, float* dist2, cudaStream_t stream);

// implemented in kmeans_tc.cu
// n_dev (optional, device): the kernel processes min(n, *n_dev) rows (row count known only on the device)
int km_tc_step(const float* X, int64_t n, int d, int64_t ld, const void* prep, int k, void* acc, int32_t* labels,
               int passes, bool accumulate, const int* n_dev, const KmRecheck* recheck, cudaStream_t stream);
bool km_tc_supported(int k, int d, int64_t ld);

// implemented in kmeans_tcb.cu (filter pass + pair check + 3xTF32 re-check of uncertain rows + exact fp64 tail)
// filter: KM_FILTER_BF16X3 = three bf16 MMAs per product, KM_FILTER_FP16 = one fp16 MMA (scaled operands),
//         KM_FILTER_NONE = no filter pass: 3xTF32 over every row with its certainty test, then the fp64 tail
#define KM_FILTER_BF16X3 0
#define KM_FILTER_FP16 1
#define KM_FILTER_NONE 2
bool km_tcb_supported(int k, int d, int64_t ld);
size_t km_tcb_ws_bytes(int64_t n, int k, int d);
// colscale (optional, fp32[d], FP16 filter only): the pass runs on the rows fl32(x * colscale) without a scaled copy
int km_tcb_step(const float* X, int64_t n, int d, int64_t ld, const float* colscale, const void* prep, int k, void* acc,
                int32_t* labels, void* ws, size_t ws_bytes, bool accumulate, int filter, cudaStream_t stream);

// implemented in lr_tc.cu (tcgen05 multinomial loss + gradient)
bool lr_mn_tc_supported(int d, int64_t ld, int K);
size_t lr_mn_tc_extra_scratch(int d);
int lr_mn_tc_launch(const float* X, const float* y, const float* w, int64_t n, int d, int K, const double* beta,
                    const double* icpt, double* partials, int grid, void* wbuf, cudaStream_t st);
