// Exact arithmetic of the I-ViT / I-BERT integer path, as device functions.
//
// The reference simulates integers in float32: every module first
// recovers x/s *without rounding*, then floors / truncates / rounds in float32, and the
// dyadic requantisation runs in float64.  Bit-exact parity therefore means replaying the
// same IEEE operations, one rounding at a time.  Everything here uses explicit
// round-to-nearest intrinsics (__fmul_rn, __fdiv_rn, ...) so that nvcc can neither
// contract a*b+c into an FMA nor substitute approximate division.
//
// Citations: /root/reference/models/quantization_utils/<file>:<line>.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ivit {

// x/s as the next module sees it: the carrier holds fl32(q*s); the consumer divides by s
// again and does NOT round (quant_modules.py:223, ivit_modules.py:36,106,165,
// ibert_modules.py:129,227,304).  Differs from q by one ulp for ~11% of (q,s).
__device__ __forceinline__ float x_eff(float q, float s) { return __fdiv_rn(__fmul_rn(q, s), s); }

// 2^k for small integer k (the reference's `2 ** tensor`, exact for integer exponents).
__device__ __forceinline__ float pow2i(int k) { return __int_as_float((127 + k) << 23); }

// ---------------------------------------------------------------- fixedpoint_mul
// quant_utils.py:220-230: out = round_he( fp64(z) * fp64(m) / 2^e ).  `mult` = m * 2^-e is
// exact in fp64 (m has <= 32 significant bits), and scaling by a power of two commutes with
// rounding, so rint(fl64(z * mult)) replays both fp64 roundings of the reference.
__device__ __forceinline__ double requant_d(double z, double mult) { return rint(__dmul_rn(z, mult)); }

// Fast form for |z| < 2^22 (then z*m < 2^53 is exact and one fused rounding is enough):
// int32 -> fp64 through the 2^52 exponent trick, FMA against 1.5*2^52 leaves the
// round-half-even integer in the low word.  |z*mult| must stay below 2^31.
__device__ __forceinline__ int requant_i32_fast(int z, double mult) {
  const double zd = __hiloint2double(0x43300000, static_cast<int>(static_cast<uint32_t>(z) ^ 0x80000000u)) -
                    4503601774854144.0;  // 2^52 + 2^31
  return __double2loint(__fma_rn(zd, mult, 6755399441055744.0));  // 1.5 * 2^52
}
// The same result with the int32 -> fp64 conversion done by I2F.F64 (one issue slot on the otherwise idle
// conversion pipe instead of XOR + MOV + DADD): for kernels bound by issue slots, not by that pipe.
__device__ __forceinline__ int requant_i32_cvt(int z, double mult) {
  return __double2loint(__fma_rn(__int2double_rn(z), mult, 6755399441055744.0));  // 1.5 * 2^52
}
__device__ __forceinline__ bool fast_requant_ok(int z) {
  return static_cast<uint32_t>(z + (1 << 22)) < (1u << 23);
}
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// General entry: z is the integer the consumer's round(x / s_in) recovers (quant_utils.py:220).
// For |z| >= 2^22 the recovery itself is replayed in float32 (s_in = per-channel scale of z).
__device__ __forceinline__ double requant_any(int z, float s_in, double mult) {
  double zd;
  if (fast_requant_ok(z)) {
    zd = static_cast<double>(z);
  } else {
    const float zf = static_cast<float>(z);  // the fp32 GEMM result the reference carries
    zd = static_cast<double>(rintf(__fdiv_rn(__fmul_rn(zf, s_in), s_in)));
  }
  return requant_d(zd, mult);
}

// ------------------------------------------------------------------- int_exp_shift
// ivit_modules.py:89-103 (ShiftGELU, n = 23) and :150-162 (Shiftmax, n = 15).
// x: x_int - max (float32, integer valued for GELU, un-rounded for softmax).
struct ExpShiftParams {
  float x0;    // floor(-1 / scaling_factor)          (:94 / :154)
  float nx0;   // n * x0                              (:95 / :155)
  int n;
};
__device__ __forceinline__ float exp_shift(float x, const ExpShiftParams& p) {
  float t = __fadd_rn(x, floorf(__fmul_rn(x, 0.5f)));            // x + floor(x/2)
  t = __fsub_rn(t, floorf(__fmul_rn(x, 0.0625f)));               //   - floor(x/2^4)
  t = fmaxf(t, p.nx0);                                           // max(x, n*x0)
  const float q = floorf(__fdiv_rn(t, p.x0));                    // floor(x / x0)
  const float r = __fsub_rn(t, __fmul_rn(p.x0, q));              // x - x0*q
  float e = __fsub_rn(__fmul_rn(r, 0.5f), p.x0);                 // r/2 - x0
  e = floorf(__fmul_rn(e, pow2i(p.n - static_cast<int>(q))));    // floor(e * 2^(n-q))
  return fmaxf(e, 0.0f);                                         // clamp(min=0)
}

// --------------------------------------------------------------------- ShiftGELU
// ivit_modules.py:105-126.  q, qmax: int8 codes out of qact_gelu (qmax = row max of q; the
// truncated values are monotone in q so max(trunc) = trunc(max)).  Returns pre_x_int*sigmoid_int.
struct GeluIvitParams {
  float s_in;            // scaling_factor of the input
  ExpShiftParams ex;     // built from s_in * 1.702 (:108)
  float two31m1;         // float(2^31 - 1) = 2^31
  float inv_shift;       // 1 / 2^(31 - output_bit + 1)
};
__device__ __forceinline__ float gelu_ivit(int q, int qmax, const GeluIvitParams& p) {
  const int pre = static_cast<int>(x_eff(static_cast<float>(q), p.s_in));       // :106-107 trunc
  const int pmx = static_cast<int>(x_eff(static_cast<float>(qmax), p.s_in));    // :110
  const float e = exp_shift(static_cast<float>(pre - pmx), p.ex);               // :113
  const float em = exp_shift(static_cast<float>(-pmx), p.ex);                   // :115
  const float sum = fminf(__fadd_rn(e, em), p.two31m1);                         // :116-118
  const float factor = floorf(__fdiv_rn(p.two31m1, sum));                       // :119
  const float sig = floorf(__fmul_rn(__fmul_rn(e, factor), p.inv_shift));       // :120
  return __fmul_rn(static_cast<float>(pre), sig);                               // :123
}

// ---------------------------------------------------------------- I-BERT GELU
// ibert_modules.py:203-235.  Works on the un-rounded x_int; result is a (non-integer) float.
struct GeluIbertParams {
  float s_in;     // input scaling factor
  float b_int;    // floor(-1.769 / (s/1.4142))                 (:205)
  float c_int;    // floor((1/-0.2888) / (s/1.4142)^2)          (:206)
  float shift;    // floor(1 / sigmoid_scaling_factor)          (:230)
  float s_out;    // s * sigmoid_scaling_factor / 2             (:233)
};
__device__ __forceinline__ float gelu_ibert(int q, const GeluIbertParams& p) {
  const float x = x_eff(static_cast<float>(q), p.s_in);                         // :227
  const float sgn = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);                 // :209
  float a = fminf(fabsf(x), -p.b_int);                                          // :210-211
  a = __fadd_rn(a, p.b_int);
  float y = __fadd_rn(__fmul_rn(a, a), p.c_int);                                // :212
  y = __fmul_rn(sgn, y);                                                        // :213
  y = floorf(__fmul_rn(y, 0.015625f));                                          // :215  / 2^6
  return __fmul_rn(x, __fadd_rn(y, p.shift));                                   // :232
}

// ------------------------------------------------- torch CPU float32 row-sum order
// The reference's float32 row reductions are order dependent once partial sums stop being
// exact (hazard H6).  ATen's vectorized_inner_sum (SumKernel.cpp) adds element e of the
// first 32*floor(n/32) to accumulator e%32 sequentially, cascading into a higher level
// every 16 steps; see oracle/ivit_oracle.py::torch_sum_lastdim (verified against torch).
// A 32-lane warp with lane = e%32 is exactly that accumulator file.
struct TorchSum32 {          // per-lane state for rows with n <= 32*16*16 elements
  float a0 = 0.f, a1 = 0.f;  // cascade levels 0 and 1
  int steps = 0;
  __device__ __forceinline__ void add(float v) {   // call with element e = 32*i + lane, i ascending
    a0 = __fadd_rn(a0, v);
    if (++steps == 16) { a1 = __fadd_rn(a1, a0); a0 = 0.f; steps = 0; }
  }
  __device__ __forceinline__ float level_sum() const { return __fadd_rn(a0, a1); }
};
// Combine: `acc` = this lane's accumulator (lane = k*8 + l), `left` = left-over whole vectors
// (elements 32*floor(n/32) + 8*v + l, added to row 0 lanes, lanes >= 8 pass 0), `tail` = the
// n%8 scalar tail elements summed in order (same value in every lane).  Result in all lanes.
__device__ __forceinline__ float torch_sum_finish(float acc, float tail_sum, const float* left, int nleft,
                                                  int lane) {
  // rows 1..3 -> row 0, in order (lanes 0..7 hold row 0)
  float p = acc;
  // left-over vectors are added to row 0 BEFORE the rows are merged (ps[0] += vecs[v])
  if (lane < 8) {
    for (int v = 0; v < nleft; ++v) p = __fadd_rn(p, left[v]);
  }
  for (int k = 1; k < 4; ++k) {
    const float o = __shfl_sync(0xffffffffu, acc, (lane & 7) + 8 * k);
    if (lane < 8) p = __fadd_rn(p, o);
  }
  float fin = tail_sum;
  for (int l = 0; l < 8; ++l) fin = __fadd_rn(fin, __shfl_sync(0xffffffffu, p, l));
  return fin;
}

}  // namespace ivit
