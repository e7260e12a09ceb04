// Row-wise kernels for the 16-bit activation locations (vit_quant.py:180-187 `*_bw = 16`, the
// `--bitwidth 16` sweep of scripts/inference.py:297): the residual stream then carries int16 codes.
//
//   * IntLayerNorm (I-ViT ivit_modules.py:30-65 / I-BERT ibert_modules.py:112-158) on int16 rows,
//     fused with the 8-bit QuantAct that follows it (Block.qact1 / qact3, final qact2).
//   * class-token row fill and the stand-alone QuantAct with 16-bit inputs / identities / outputs.
//
// With 16-bit codes the shortcuts of the int8 kernels (rowwise.cu) no longer hold: x/s is off an
// integer by up to 2^-9 per element, so the float32 row sum is replayed in torch's CPU order for
// EVERY row (not just exact ties), and |y| exceeds 46340, so the int32 square wraps (hazard H9)
// exactly as `y_int ** 2` does in the reference.  One warp per row; lane t owns elements 32 i + t
// (torch's accumulator layout, intmath.cuh TorchSum32).
#include <cmath>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "intmath.cuh"
#include "ivit_b200.h"

namespace ivit {
int fail(int code, const char* msg);
int check_launch(const char* what);
int ln_variant();          // ivit_layernorm_set_variant (rowwise.cu)

// exact replay of the QuantAct behind the LayerNorm for one element (quant_utils.py:220-251)
__device__ __noinline__ int ln16_requant(float v, float sc, double mult, int qmin, int qmax) {
  const float xo = __fmul_rn(v, sc);
  const float z = rintf(__fdiv_rn(xo, sc));
  const double rd = requant_d(static_cast<double>(z), mult);
  return static_cast<int>(fmin(fmax(rd, static_cast<double>(qmin)), static_cast<double>(qmax)));
}

// torch.sum over a channel-STRIDED row (ATen outer reduction, oracle/ivit_oracle.py::torch_outer_sum_lastdim),
// evaluated by one lane over the row's x/s values in shared memory.  tail = false: the token lies in a
// whole block of 32 tokens of its image -> sequential with a cascade every 2^lp elements
// (multi_row_sum); tail = true: row_sum -> 4 interleaved accumulators cascaded the same way over
// C/4 steps, left-overs into accumulator 0, then 1..3 added in order.
__device__ __noinline__ float outer_order_sum(const float* xe, int C, bool tail) {
  if (!tail) {
    int cl = 0;
    while ((1 << cl) < C) ++cl;
    const int lp = cl / 4 > 4 ? cl / 4 : 4, step = 1 << lp, mask = step - 1;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int i = 0;
    while (i + step <= C) {
      for (int u = 0; u < step; ++u, ++i) acc[0] = __fadd_rn(acc[0], xe[i]);
      for (int j = 1; j < 4; ++j) {
        acc[j] = __fadd_rn(acc[j], acc[j - 1]);
        acc[j - 1] = 0.f;
        if ((i & (mask << (j * lp))) != 0) break;
      }
    }
    for (; i < C; ++i) acc[0] = __fadd_rn(acc[0], xe[i]);
    for (int j = 1; j < 4; ++j) acc[0] = __fadd_rn(acc[0], acc[j]);
    return acc[0];
  }
  const int n = C / 4;
  int cl = 0;
  while ((1 << cl) < n) ++cl;
  const int lp = cl / 4 > 4 ? cl / 4 : 4, step = 1 << lp, mask = step - 1;
  float acc[4][4];
  for (int j = 0; j < 4; ++j)
    for (int k = 0; k < 4; ++k) acc[j][k] = 0.f;
  int i = 0;
  while (i + step <= n) {
    for (int u = 0; u < step; ++u, ++i)
      for (int k = 0; k < 4; ++k) acc[0][k] = __fadd_rn(acc[0][k], xe[4 * i + k]);
    for (int j = 1; j < 4; ++j) {
      for (int k = 0; k < 4; ++k) {
        acc[j][k] = __fadd_rn(acc[j][k], acc[j - 1][k]);
        acc[j - 1][k] = 0.f;
      }
      if ((i & (mask << (j * lp))) != 0) break;
    }
  }
  for (; i < n; ++i)
    for (int k = 0; k < 4; ++k) acc[0][k] = __fadd_rn(acc[0][k], xe[4 * i + k]);
  for (int j = 1; j < 4; ++j)
    for (int k = 0; k < 4; ++k) acc[0][k] = __fadd_rn(acc[0][k], acc[j][k]);
  float p0 = acc[0][0];
  for (int e = 4 * n; e < C; ++e) p0 = __fadd_rn(p0, xe[e]);
  for (int k = 1; k < 4; ++k) p0 = __fadd_rn(p0, acc[0][k]);
  return p0;
}

// kind 0: I-ViT, kind 1: I-BERT.  x int16 (or int8) [rows, C] with row stride ldx ELEMENTS; out int8
// [rows, C].  OUTER: the reference's tensor is channel-strided with `tokens` tokens per image (Swin
// stage 0: the memory layout of the patch-embedding conv survives every elementwise op up to the
// first PatchMerging), so torch sums the mean in its outer-reduction order (row % tokens decides
// between its two forms).
template <int NPL, typename XT, bool OUTER>
__global__ void __launch_bounds__(256)
layernorm_x16_requant_kernel(int kind, const XT* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows,
                             float s_in, const float* __restrict__ bias_int, const float* __restrict__ sf,
                             const double* __restrict__ mult, int qmin, int qmax, float shift_pow, float inv_shift_pow,
                             int use_int_sqrt, int tokens) {
  constexpr int C = NPL * 32;
  __shared__ __align__(16) int8_t stage[8][C];
  __shared__ float srow[OUTER ? 8 : 1][OUTER ? C : 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= rows) return;
  const XT* xr = x + static_cast<long long>(row) * ldx;
  float xe[NPL];
  TorchSum32 acc;
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    xe[i] = x_eff(static_cast<float>(__ldg(xr + 32 * i + lane)), s_in);       // ivit :36 / ibert :129
    acc.add(xe[i]);
    if (OUTER) srow[warp][32 * i + lane] = xe[i];
  }
  float sum_f;
  if (OUTER) {
    __syncwarp();
    if ((row % tokens) >= (tokens / 32) * 32) {          // warp-uniform; tail tokens of an image only (row_sum form)
      float sv = 0.f;
      if (lane == 0) sv = outer_order_sum(srow[warp], C, true);
      sum_f = __shfl_sync(0xffffffffu, sv, 0);
    } else {
      // multi_row_sum form, in parallel: the cascade adds whole groups of 16 elements (each summed
      // sequentially from zero) into the level-1 accumulator, and 16 of those into level 2 -- lane g sums
      // group g, then the groups are combined in order (C <= 384: at most 24 groups, two levels)
      constexpr int G = C / 16;
      float sg = 0.f;
      if (lane < G) {
#pragma unroll
        for (int u = 0; u < 16; ++u) sg = __fadd_rn(sg, srow[warp][16 * lane + u]);
      }
      float a1 = 0.f, a2 = 0.f;
#pragma unroll
      for (int g = 0; g < G; ++g) {
        a1 = __fadd_rn(a1, __shfl_sync(0xffffffffu, sg, g));
        if (((g + 1) & 15) == 0) { a2 = __fadd_rn(a2, a1); a1 = 0.f; }
      }
      sum_f = __fadd_rn(__fadd_rn(0.f, a1), a2);           // acc[0] (empty: C % 16 == 0) += acc[1]; += acc[2]
    }
  } else {
    sum_f = torch_sum_finish(acc.level_sum(), 0.f, nullptr, 0, lane);
  }
  const float mean = rintf(__fdiv_rn(sum_f, static_cast<float>(C)));          // ivit :37 / ibert :130
  float y[NPL];
  float factor;
  if (kind == 0) {
    const int mean_i = static_cast<int>(mean);
    long long var = 0;
#pragma unroll
    for (int i = 0; i < NPL; ++i) {
      const int yi = static_cast<int>(xe[i]) - mean_i;                        // :38 trunc, :40
      const int sq = static_cast<int>(static_cast<uint32_t>(yi) * static_cast<uint32_t>(yi));   // :41 int32 wrap (H9)
      var += static_cast<long long>(sq);                                      // :42 int64 sum
      y[i] = static_cast<float>(yi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
    const float var_f = static_cast<float>(var);
    float k = 65536.f;
#pragma unroll
    for (int it = 0; it < 10; ++it) k = floorf(__fmul_rn(__fadd_rn(k, floorf(__fdiv_rn(var_f, k))), 0.5f));   // :45-49
    factor = floorf(__fdiv_rn(2147483648.f, k));                              // :51 (2^31-1 -> 2^31 in float32)
  } else {
    TorchSum32 vacc;
#pragma unroll
    for (int i = 0; i < NPL; ++i) {
      y[i] = __fsub_rn(xe[i], mean);                                          // :131
      const float ys = floorf(__fmul_rn(y[i], inv_shift_pow));                // :132
      vacc.add(__fmul_rn(ys, ys));                                            // :133
    }
    const float var_f = torch_sum_finish(vacc.level_sum(), 0.f, nullptr, 0, lane);   // :134
    float std_i;
    if (use_int_sqrt) {                                                       // :85-109
      const float n = fmaxf(var_f, 0.f);
      const float bits = floorf(log2f(fmaxf(n, 1.f))) + 1.f;
      float xk = pow2i(static_cast<int>(ceilf(__fmul_rn(bits, 0.5f))));
      for (int it = 0; it < 4; ++it) {
        const float inv = floorf(__fdiv_rn(n, fmaxf(xk, 1.f)));
        xk = floorf(__fmul_rn(__fadd_rn(xk, inv), 0.5f));
      }
      std_i = (var_f > 0.f) ? xk : 0.f;
    } else {
      std_i = floorf(__fsqrt_rn(var_f));                                      // :145
    }
    std_i = __fmul_rn(std_i, shift_pow);
    factor = floorf(__fdiv_rn(2147483648.f, std_i));                          // :146 (2^31)
  }
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    const int c = 32 * i + lane;
    float v = floorf(__fmul_rn(__fmul_rn(y[i], factor), 0.5f));               // ivit :52 / ibert :147
    v = __fadd_rn(v, __ldg(bias_int + c));                                    // :61 / :154
    // QuantAct: the exact value is rint(z * mult), z = rint(fl32(fl32(v sf) / sf)), |z - v| <= |v| 2^-22.  Inside the
    // clamp range the float32 product v * fl32(mult) is within 5e-5 of z * mult, so unless it lies within 1e-4
    // of a rounding tie (or is a NaN from a zero-variance row) its rounding IS the exact answer (same argument
    // as rowwise.cu); the replay runs for those elements only -- it costs four XU-pipe operations.
    const double md = __ldg(mult + c);
    const float ra = fminf(fmaxf(__fmul_rn(v, static_cast<float>(md)), -200.f), 200.f);
    const float tm = __fadd_rn(ra, 12582912.f);                                 // rint(ra) in the low mantissa bits
    const float rr = __fsub_rn(tm, 12582912.f);
    int code = clampi(__float_as_int(tm) - 0x4B400000, qmin, qmax);
    if (!(fabsf(__fsub_rn(ra, rr)) < 0.4999f)) code = ln16_requant(v, __ldg(sf + c), md, qmin, qmax);
    stage[warp][c] = static_cast<int8_t>(code);
  }
  __syncwarp();
  int8_t* orow = out + static_cast<long long>(row) * C;
  for (int v = lane; v < C / 16; v += 32)
    *reinterpret_cast<int4*>(orow + v * 16) = *reinterpret_cast<const int4*>(&stage[warp][v * 16]);
}

// Swin stage 0 (C = 96, channel-strided order, every token inside a whole block of 32 of its image): eight
// lanes per row, four rows per warp.  Lane g < 6 of a row owns channels [16 g, 16 g + 16) -- exactly one group
// of torch's 16-element cascade -- so the ordered mean needs 16 sequential adds per lane plus a 6-step
// in-order combine, the per-row scalar work (ten Newton steps) is replicated on 8 lanes instead of 32,
// rows come in as 32-byte and leave as 16-byte vectors without shared-memory staging, and the
// per-channel constants live in registers across the grid-stride loop.  I-ViT kind only.
template <typename XT>
__global__ void __launch_bounds__(256, 4)
layernorm96_strided_kernel(const XT* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows, float s_in,
                           const float* __restrict__ bias_int, const float* __restrict__ sf,
                           const double* __restrict__ mult, int qmin, int qmax) {
  constexpr int C = 96;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane & 7, sub = lane >> 3;
  const bool owner = g < 6;
  const int c0 = 16 * (owner ? g : 0);
  float bia[16], mf[16];
#pragma unroll
  for (int u = 0; u < 16; ++u) {
    bia[u] = __ldg(bias_int + c0 + u);
    mf[u] = static_cast<float>(__ldg(mult + c0 + u));
  }
  const int quads = (rows + 3) / 4;
  for (int q = blockIdx.x * 8 + warp; q < quads; q += gridDim.x * 8) {
    const int row = 4 * q + sub;
    const bool live = owner && row < rows;
    int qv[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) qv[u] = 0;
    if (live) {
      const XT* xr = x + static_cast<long long>(row) * ldx + c0;
      if (sizeof(XT) == 2) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int4 w = __ldg(reinterpret_cast<const int4*>(xr) + h);
          const int ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
          for (int u = 0; u < 8; ++u) qv[8 * h + u] = static_cast<int16_t>(ww[u >> 1] >> (16 * (u & 1)));
        }
      } else {
        const int4 w = __ldg(reinterpret_cast<const int4*>(xr));
        const int ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int u = 0; u < 16; ++u) qv[u] = static_cast<int8_t>(ww[u >> 2] >> (8 * (u & 3)));
      }
    }
    float xe[16];
    float sg = 0.f;
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      xe[u] = x_eff(static_cast<float>(qv[u]), s_in);                           // ivit_modules.py:36
      sg = __fadd_rn(sg, xe[u]);                                               // one 16-element cascade group
    }
    float a1 = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) a1 = __fadd_rn(a1, __shfl_sync(0xffffffffu, sg, k, 8));   // groups in order
    const int mean_i = static_cast<int>(rintf(__fdiv_rn(a1, static_cast<float>(C))));    // :37
    long long var = 0;
    int yi[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      yi[u] = static_cast<int>(xe[u]) - mean_i;                                // :38 trunc, :40
      const int sq = static_cast<int>(static_cast<uint32_t>(yi[u]) * static_cast<uint32_t>(yi[u]));   // :41 wraps (H9)
      var += owner ? static_cast<long long>(sq) : 0ll;                         // :42
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o, 8);
    const float var_f = static_cast<float>(var);
    float k = 65536.f;
#pragma unroll
    for (int it = 0; it < 10; ++it) k = floorf(__fmul_rn(__fadd_rn(k, floorf(__fdiv_rn(var_f, k))), 0.5f));   // :45-49
    const float factor = floorf(__fdiv_rn(2147483648.f, k));                   // :51
    int code[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      float v = floorf(__fmul_rn(__fmul_rn(static_cast<float>(yi[u]), factor), 0.5f));      // :52
      v = __fadd_rn(v, bia[u]);                                                // :61
      const float ra = fminf(fmaxf(__fmul_rn(v, mf[u]), -200.f), 200.f);
      const float tm = __fadd_rn(ra, 12582912.f);
      const float rr = __fsub_rn(tm, 12582912.f);
      code[u] = clampi(__float_as_int(tm) - 0x4B400000, qmin, qmax);
      if (!(fabsf(__fsub_rn(ra, rr)) < 0.4999f))                               // near a tie / NaN: exact replay
        code[u] = ln16_requant(v, __ldg(sf + c0 + u), __ldg(mult + c0 + u), qmin, qmax);
    }
    if (live) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        w[j] = (code[4 * j] & 0xff) | ((code[4 * j + 1] & 0xff) << 8) | ((code[4 * j + 2] & 0xff) << 16) |
               (static_cast<uint32_t>(code[4 * j + 3]) << 24);
      *reinterpret_cast<int4*>(out + static_cast<long long>(row) * C + c0) = make_int4(w[0], w[1], w[2], w[3]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Round 2: the I-ViT kind again, restructured like rowwise.cu's int8 kernel.  Same arithmetic, fewer
// instructions per element:
//   * x/s as c = fl(r inv_s), c' = fma(fma(-c, s, r), inv_s, c) with r = fl(q s), inv_s = fl(1/s): three FMA-pipe
//     operations instead of the IEEE division sequence.  It is used only when the HOST has checked, for this s and
//     EVERY code of the input width, that c' equals the IEEE quotient (xeff_fma_ok below: same IEEE arithmetic on
//     the host cores, cached per scale); otherwise the kernels above run.
//   * R rows in flight per warp; the per-row scalar work (mean division, the ten Newton steps of :45-49, the
//     2^31 / std division) runs once with row r on lane r, and the Newton loop stops as soon as every lane has
//     reached a fixed point (further steps would return the same value).
//   * the QuantAct tail of rowwise.cu: fl(y factor) / 2 == fl(y (factor / 2)); clamp before rounding; the near-tie
//     test as two FMAs whose roundings must agree; the code is the low byte of the magic-number sum; an element
//     that fails the test is replayed exactly inside its own group of four (ln_tail_group).
__device__ __forceinline__ float x_eff_fma(float q, float s, float inv_s) {
  const float r = __fmul_rn(q, s);
  const float c = __fmul_rn(r, inv_s);
  return __fmaf_rn(__fmaf_rn(-c, s, r), inv_s, c);
}

// The QuantAct behind the LayerNorm for G values of one lane (channels ch[0..G)): float32 product clamped before
// rounding; rint(ra (1 -+ 2^-18) + 1.5 * 2^23) must agree (the exact value lies between them: float32(mult), z vs v and
// the product rounding are each < 2^-21.5 relative), the code is then the low byte of the sum.  Where they differ
// (~7e-4 of the elements) the exact replay runs for that element, inline and on its own group only -- a second pass
// over the whole row would run for most warps.
template <int G>
__device__ __forceinline__ void ln_tail_group(const float (&v)[G], const float (&mf)[G], const int (&ch)[G],
                                              const float* __restrict__ sf, const double* __restrict__ mult, bool fast_ok,
                                              int qmin, int qmax, uint32_t (&lob)[G]) {
  uint32_t x[G], any = 0;
#pragma unroll
  for (int e = 0; e < G; ++e) {
    const float ra = fminf(fmaxf(__fmul_rn(v[e], mf[e]), -128.4f), 127.4f);
    const float lo = __fmaf_rn(ra, 0.999996185302734375f, 12582912.f);        // 1 - 2^-18
    const float hi = __fmaf_rn(ra, 1.000003814697265625f, 12582912.f);        // 1 + 2^-18
    x[e] = __float_as_uint(lo) ^ __float_as_uint(hi);
    any |= x[e];
    lob[e] = __float_as_uint(lo);
  }
  if ((any & 1u) || !fast_ok) {
#pragma unroll
    for (int e = 0; e < G; ++e)
      if (!fast_ok || (x[e] & 1u))
        lob[e] = static_cast<uint32_t>(ln16_requant(v[e], __ldg(sf + ch[e]), __ldg(mult + ch[e]), qmin, qmax)) & 0xffu;
  }
}

// ten steps of k <- floor((k + floor(v / k)) / 2) from 2^16 (ivit_modules.py:45-49), then floor(2^31 / k) (:51)
__device__ __forceinline__ float ivit_std_factor(float var_f) {
  float k = 65536.f;
#pragma unroll 1
  for (int it = 0; it < 10; ++it) {
    const float kn = floorf(__fmul_rn(__fadd_rn(k, floorf(__fdiv_rn(var_f, k))), 0.5f));
    const bool same = kn == k;                  // fixed point: every later step returns k again (NaN never matches)
    k = kn;
    if (__all_sync(0xffffffffu, same)) break;
  }
  return floorf(__fdiv_rn(2147483648.f, k));
}

template <int NPL, int R, typename XT, int MINB>
__global__ void __launch_bounds__(256, MINB)
layernorm16_ivit_kernel(const XT* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows, float s_in,
                        float inv_s, const float* __restrict__ bias_int, const float* __restrict__ sf,
                        const double* __restrict__ mult, int qmin, int qmax) {
  constexpr int C = NPL * 32;
  constexpr int VPR = C / 16;
  __shared__ __align__(16) int8_t stage[8][R][C];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float bia[NPL], mf[NPL];
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    bia[i] = __ldg(bias_int + 32 * i + lane);
    mf[i] = static_cast<float>(__ldg(mult + 32 * i + lane));
  }
  const bool fast_ok = qmin == -128 && qmax == 127;
  const int ngroups = (rows + R - 1) / R;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8) {
    const int row0 = g * R;
    const int nr = min(R, rows - row0);
    int y[R][NPL];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const XT* xr = x + static_cast<long long>(row0 + min(r, nr - 1)) * ldx + lane;
#pragma unroll
      for (int i = 0; i < NPL; ++i) y[r][i] = static_cast<int>(__ldg(xr + 32 * i));
    }
    float sum_l = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      TorchSum32 acc;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        const float xe = x_eff_fma(static_cast<float>(y[r][i]), s_in, inv_s);     // ivit_modules.py:36
        acc.add(xe);
        y[r][i] = static_cast<int>(xe);                                          // :38 trunc
      }
      const float sr = torch_sum_finish(acc.level_sum(), 0.f, nullptr, 0, lane);
      sum_l = (lane == r) ? sr : sum_l;
    }
    const int mean_l = static_cast<int>(rintf(__fdiv_rn(sum_l, static_cast<float>(C))));   // :37, row r on lane r
    long long var_l = 1ll << 32;              // lanes without a row: 2^16 is a fixed point at once (early exit)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int mean = __shfl_sync(0xffffffffu, mean_l, r);
      long long var = 0;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] -= mean;                                                          // :40
        const int sq = static_cast<int>(static_cast<uint32_t>(y[r][i]) * static_cast<uint32_t>(y[r][i]));   // :41 wraps (H9)
        var += static_cast<long long>(sq);                                        // :42
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
      var_l = (lane == r) ? var : var_l;
    }
    const float fct_l = ivit_std_factor(static_cast<float>(var_l));
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float fh = __fmul_rn(__shfl_sync(0xffffffffu, fct_l, r), 0.5f);
      constexpr int G = NPL % 4 == 0 ? 4 : 2;
#pragma unroll
      for (int i0 = 0; i0 < NPL; i0 += G) {
        float v[G], mg[G];
        int ch[G];
        uint32_t lob[G];
#pragma unroll
        for (int e = 0; e < G; ++e) {
          const float yf = __fsub_rn(__int_as_float(0x4B400000 + y[r][i0 + e]), 12582912.f);   // |y| < 2^17: exact
          v[e] = __fadd_rn(floorf(__fmul_rn(yf, fh)), bia[i0 + e]);               // :52, :61
          mg[e] = mf[i0 + e];
          ch[e] = 32 * (i0 + e) + lane;
        }
        ln_tail_group<G>(v, mg, ch, sf, mult, fast_ok, qmin, qmax, lob);
#pragma unroll
        for (int e = 0; e < G; ++e) stage[warp][r][ch[e]] = static_cast<int8_t>(lob[e]);
      }
    }
    __syncwarp();
    for (int v = lane; v < R * VPR; v += 32) {
      const int r = v / VPR, vv = v - r * VPR;
      if (r < nr)
        *reinterpret_cast<int4*>(out + static_cast<long long>(row0 + r) * C + vv * 16) =
            *reinterpret_cast<const int4*>(&stage[warp][r][vv * 16]);
    }
    __syncwarp();
  }
}

// The I-BERT kind (ibert_modules.py:112-158) in the same shape: y stays a float (x/s - mean, :131), the variance is a
// second float32 row sum in torch's order over floor(y / 2^shift)^2 (:132-134), the standard deviation one square
// root (or the four integer Newton steps of :85-109) -- evaluated with row r on lane r.
template <int NPL, int R, typename XT, int MINB>
__global__ void __launch_bounds__(256, MINB)
layernorm16_ibert_kernel(const XT* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows, float s_in,
                         float inv_s, const float* __restrict__ bias_int, const float* __restrict__ sf,
                         const double* __restrict__ mult, int qmin, int qmax, float shift_pow, float inv_shift_pow,
                         int use_int_sqrt) {
  constexpr int C = NPL * 32;
  constexpr int VPR = C / 16;
  __shared__ __align__(16) int8_t stage[8][R][C];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float bia[NPL], mf[NPL];
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    bia[i] = __ldg(bias_int + 32 * i + lane);
    mf[i] = static_cast<float>(__ldg(mult + 32 * i + lane));
  }
  const bool fast_ok = qmin == -128 && qmax == 127;
  const int ngroups = (rows + R - 1) / R;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8) {
    const int row0 = g * R;
    const int nr = min(R, rows - row0);
    float y[R][NPL];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const XT* xr = x + static_cast<long long>(row0 + min(r, nr - 1)) * ldx + lane;
#pragma unroll
      for (int i = 0; i < NPL; ++i) y[r][i] = static_cast<float>(static_cast<int>(__ldg(xr + 32 * i)));
    }
    float sum_l = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      TorchSum32 acc;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] = x_eff_fma(y[r][i], s_in, inv_s);                               // ibert_modules.py:129
        acc.add(y[r][i]);
      }
      const float sr = torch_sum_finish(acc.level_sum(), 0.f, nullptr, 0, lane);
      sum_l = (lane == r) ? sr : sum_l;
    }
    const float mean_l = rintf(__fdiv_rn(sum_l, static_cast<float>(C)));          // :130, row r on lane r
    float var_l = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float mean = __shfl_sync(0xffffffffu, mean_l, r);
      TorchSum32 vacc;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] = __fsub_rn(y[r][i], mean);                                       // :131
        const float ys = floorf(__fmul_rn(y[r][i], inv_shift_pow));               // :132
        vacc.add(__fmul_rn(ys, ys));                                              // :133
      }
      const float vr = torch_sum_finish(vacc.level_sum(), 0.f, nullptr, 0, lane); // :134
      var_l = (lane == r) ? vr : var_l;
    }
    float std_l;
    if (use_int_sqrt) {                                                           // :85-109
      const float n = fmaxf(var_l, 0.f);
      const float bits = floorf(log2f(fmaxf(n, 1.f))) + 1.f;
      float xk = pow2i(static_cast<int>(ceilf(__fmul_rn(bits, 0.5f))));
      for (int it = 0; it < 4; ++it) {
        const float inv = floorf(__fdiv_rn(n, fmaxf(xk, 1.f)));
        xk = floorf(__fmul_rn(__fadd_rn(xk, inv), 0.5f));
      }
      std_l = (var_l > 0.f) ? xk : 0.f;
    } else {
      std_l = floorf(__fsqrt_rn(var_l));                                          // :145
    }
    std_l = __fmul_rn(std_l, shift_pow);
    const float fct_l = floorf(__fdiv_rn(2147483648.f, std_l));                   // :146 (2^31)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float fh = __fmul_rn(__shfl_sync(0xffffffffu, fct_l, r), 0.5f);
      constexpr int G = NPL % 4 == 0 ? 4 : 2;
#pragma unroll
      for (int i0 = 0; i0 < NPL; i0 += G) {
        float v[G], mg[G];
        int ch[G];
        uint32_t lob[G];
#pragma unroll
        for (int e = 0; e < G; ++e) {
          v[e] = __fadd_rn(floorf(__fmul_rn(y[r][i0 + e], fh)), bia[i0 + e]);     // :147, :154
          mg[e] = mf[i0 + e];
          ch[e] = 32 * (i0 + e) + lane;
        }
        ln_tail_group<G>(v, mg, ch, sf, mult, fast_ok, qmin, qmax, lob);
#pragma unroll
        for (int e = 0; e < G; ++e) stage[warp][r][ch[e]] = static_cast<int8_t>(lob[e]);
      }
    }
    __syncwarp();
    for (int v = lane; v < R * VPR; v += 32) {
      const int r = v / VPR, vv = v - r * VPR;
      if (r < nr)
        *reinterpret_cast<int4*>(out + static_cast<long long>(row0 + r) * C + vv * 16) =
            *reinterpret_cast<const int4*>(&stage[warp][r][vv * 16]);
    }
    __syncwarp();
  }
}

// Channel-strided order (Swin stage 0: every token inside a whole block of 32 of its image, so the row sum is ATen's
// multi_row_sum: groups of 16 channels summed sequentially from zero, the group sums added in order).  Two lanes per
// row, GPL groups each (C = 32 GPL: 96 for Swin-T/S, 128 for Swin-B), sixteen rows per warp: all lanes carry
// elements, the scalar work of a row runs on its own two lanes, rows come in and leave as 16-byte vectors.
template <int GPL, typename XT>
__global__ void __launch_bounds__(256, 2)
layernorm16_outer2_kernel(const XT* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows, float s_in,
                          float inv_s, const float* __restrict__ bias_int, const float* __restrict__ sf,
                          const double* __restrict__ mult, int qmin, int qmax) {
  constexpr int C = 32 * GPL, E = 16 * GPL;
  __shared__ __align__(16) float sb[C];
  __shared__ __align__(16) float sm[C];
  if (threadIdx.x < C) {
    sb[threadIdx.x] = __ldg(bias_int + threadIdx.x);
    sm[threadIdx.x] = static_cast<float>(__ldg(mult + threadIdx.x));
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = lane & 1, sub = lane >> 1;
  const int c0 = half * E;
  const bool fast_ok = qmin == -128 && qmax == 127;
  const int ngroups = (rows + 15) / 16;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8) {
    const int row = 16 * g + sub;
    const bool live = row < rows;
    const XT* xr = x + static_cast<long long>(live ? row : rows - 1) * ldx + c0;
    int t[E];
    if (sizeof(XT) == 2) {
#pragma unroll
      for (int h = 0; h < E / 8; ++h) {
        const int4 w = __ldg(reinterpret_cast<const int4*>(xr) + h);
        const int ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int u = 0; u < 8; ++u) t[8 * h + u] = static_cast<int16_t>(ww[u >> 1] >> (16 * (u & 1)));
      }
    } else {
#pragma unroll
      for (int h = 0; h < E / 16; ++h) {
        const int4 w = __ldg(reinterpret_cast<const int4*>(xr) + h);
        const int ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int u = 0; u < 16; ++u) t[16 * h + u] = static_cast<int8_t>(ww[u >> 2] >> (8 * (u & 3)));
      }
    }
    float gs[GPL];
#pragma unroll
    for (int k = 0; k < GPL; ++k) {
      float sg = 0.f;
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const float xe = x_eff_fma(static_cast<float>(t[16 * k + u]), s_in, inv_s);   // ivit_modules.py:36
        sg = __fadd_rn(sg, xe);                                                      // one 16-element cascade group
        t[16 * k + u] = static_cast<int>(xe);                                        // :38 trunc
      }
      gs[k] = sg;
    }
    // group sums in order: ((((0 + g0) + g1) + ...) + g_{2 GPL - 1}); the even lane holds the first GPL groups
    float pre = 0.f;
#pragma unroll
    for (int k = 0; k < GPL; ++k) pre = __fadd_rn(pre, gs[k]);
    float tot = __shfl_sync(0xffffffffu, pre, lane & ~1);
#pragma unroll
    for (int k = 0; k < GPL; ++k) tot = __fadd_rn(tot, gs[k]);
    tot = __shfl_sync(0xffffffffu, tot, lane | 1);
    const int mean = static_cast<int>(rintf(__fdiv_rn(tot, static_cast<float>(C))));  // :37
    long long var = 0;
#pragma unroll
    for (int u = 0; u < E; ++u) {
      t[u] -= mean;                                                                  // :40
      const int sq = static_cast<int>(static_cast<uint32_t>(t[u]) * static_cast<uint32_t>(t[u]));   // :41 wraps (H9)
      var += static_cast<long long>(sq);                                             // :42
    }
    var += __shfl_xor_sync(0xffffffffu, var, 1);
    const float fh = __fmul_rn(ivit_std_factor(static_cast<float>(var)), 0.5f);       // :45-51
    int8_t* orow = out + static_cast<long long>(row) * C + c0;
#pragma unroll
    for (int h = 0; h < E / 16; ++h) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float4 b4 = *reinterpret_cast<const float4*>(&sb[c0 + 16 * h + 4 * j]);
        const float4 m4 = *reinterpret_cast<const float4*>(&sm[c0 + 16 * h + 4 * j]);
        const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w};
        float v[4];
        int ch[4];
        uint32_t lob[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float yf = __fsub_rn(__int_as_float(0x4B400000 + t[16 * h + 4 * j + e]), 12582912.f);
          v[e] = __fadd_rn(floorf(__fmul_rn(yf, fh)), bb[e]);                        // :52, :61
          ch[e] = c0 + 16 * h + 4 * j + e;
        }
        ln_tail_group<4>(v, mm, ch, sf, mult, fast_ok, qmin, qmax, lob);
        w[j] = __byte_perm(__byte_perm(lob[0], lob[1], 0x0040), __byte_perm(lob[2], lob[3], 0x0040), 0x5410);
      }
      if (live) *(reinterpret_cast<int4*>(orow) + h) = make_int4(w[0], w[1], w[2], w[3]);
    }
  }
}

__global__ void fill_cls_rows16_kernel(int16_t* __restrict__ out, const int16_t* __restrict__ cls_row, int B, int N,
                                       int C) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B * C) return;
  const int b = t / C, c = t % C;
  out[(static_cast<long long>(b) * N) * C + c] = cls_row[c];
}

// Stand-alone QuantAct, any of int32 / int16 / int8 codes in, int8 / int16 identity, int8 / int16 out.
template <typename ZT, typename IT, typename OT>
__global__ void requant_any_kernel(const ZT* __restrict__ z, const float* __restrict__ s_in,
                                   const double* __restrict__ mult, int per_channel, const IT* __restrict__ id,
                                   long long id_rows, double mult_id, OT* __restrict__ out, long long rows, int C,
                                   int qmin, int qmax) {
  const long long n = rows * C;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < n;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = per_channel ? static_cast<int>(t % C) : 0;
    double r = requant_any(static_cast<int>(z[t]), s_in[c], mult[c]);
    if (id != nullptr) {
      const long long it = (id_rows > 0 && id_rows < rows) ? ((t / C) % id_rows) * C + t % C : t;
      r += requant_d(static_cast<double>(id[it]), mult_id);
    }
    out[t] = static_cast<OT>(static_cast<int>(fmin(fmax(r, (double)qmin), (double)qmax)));
  }
}

// QuantAct on a float tensor at any width up to 16 bits: quant_utils.py:49 + :93 (reciprocal multiply)
template <typename OT>
__global__ void quantize_codes_kernel(const float* __restrict__ x, OT* __restrict__ out, float inv_s, long long n,
                                      float qmin, float qmax) {
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < n;
       t += static_cast<long long>(gridDim.x) * blockDim.x)
    out[t] = static_cast<OT>(static_cast<int>(fminf(fmaxf(rintf(__fmul_rn(inv_s, x[t])), qmin), qmax)));
}

template <typename ZT, typename IT, typename OT>
static int launch_requant(const void* z, const float* s_in, const double* mult, int per_channel, const void* id,
                          long long id_rows, double mult_id, void* out, long long rows, int C, int qmin, int qmax,
                          cudaStream_t st) {
  const long long n = rows * C;
  long long g = (n + 255) / 256;
  if (g > 148LL * 16) g = 148LL * 16;
  requant_any_kernel<ZT, IT, OT><<<static_cast<int>(g), 256, 0, st>>>(
      static_cast<const ZT*>(z), s_in, mult, per_channel, static_cast<const IT*>(id), id_rows, mult_id,
      static_cast<OT*>(out), rows, C, qmin, qmax);
  return check_launch("requant_any_kernel");
}

template <typename ZT>
static int dispatch_requant(int id_bits, int out_bits, const void* z, const float* s_in, const double* mult,
                            int per_channel, const void* id, long long id_rows, double mult_id, void* out,
                            long long rows, int C, int qmin, int qmax, cudaStream_t st) {
  if (out_bits == 16) {
    if (id_bits == 16)
      return launch_requant<ZT, int16_t, int16_t>(z, s_in, mult, per_channel, id, id_rows, mult_id, out, rows, C, qmin, qmax, st);
    return launch_requant<ZT, int8_t, int16_t>(z, s_in, mult, per_channel, id, id_rows, mult_id, out, rows, C, qmin, qmax, st);
  }
  if (id_bits == 16)
    return launch_requant<ZT, int16_t, int8_t>(z, s_in, mult, per_channel, id, id_rows, mult_id, out, rows, C, qmin, qmax, st);
  return launch_requant<ZT, int8_t, int8_t>(z, s_in, mult, per_channel, id, id_rows, mult_id, out, rows, C, qmin, qmax, st);
}

}  // namespace ivit

using namespace ivit;

// Host check behind x_eff_fma: for this scale and every code of the input width, c + fma(-c, s, r) * inv_s (two
// fused operations) must equal the IEEE quotient r / s.  The host cores do the same IEEE single-precision arithmetic
// as the device (divss, mulss, a correctly rounded fmaf), 65536 codes take ~0.2 ms, and the verdict is cached per
// (scale, width).  A scale that fails keeps the division kernels.
static bool xeff_fma_ok(float s, int bits) {
  static std::mutex mu;
  static std::unordered_map<uint64_t, bool> cache;
  uint32_t sb;
  std::memcpy(&sb, &s, 4);
  const uint64_t key = (static_cast<uint64_t>(bits) << 32) | sb;
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  bool ok = std::isfinite(s) && s > 0.f;
  if (ok) {
    const volatile float inv = 1.0f / s;
    const int lim = bits == 8 ? 128 : 32768;
    for (int q = -lim; q < lim && ok; ++q) {
      const volatile float r = static_cast<float>(q) * s;
      const volatile float ref = r / s;
      const volatile float c = r * inv;
      const volatile float e = std::fmaf(-c, s, r);
      const volatile float c2 = std::fmaf(e, inv, c);
      ok = c2 == ref;
    }
  }
  if (cache.size() > 4096) cache.clear();
  cache[key] = ok;
  return ok;
}

template <typename K>
static int resident_ctas(K kernel, int fallback) {
  int nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, 256, 0) != cudaSuccess || nb < 1) nb = fallback;
  return nb;
}


template <int NPL, int R, typename XT, int MINB>
static int launch_ln16_rows(const XT* x, int64_t ldx, int8_t* out, int rows, float s_in, const float* bias_int,
                            const float* sf, const double* mult, int qmin, int qmax, cudaStream_t st) {
  static int resident[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev = dev < 0 || dev >= 16 ? 0 : dev;
  if (resident[dev] == 0) resident[dev] = resident_ctas(layernorm16_ivit_kernel<NPL, R, XT, MINB>, MINB);
  const int ngroups = (rows + R - 1) / R;
  int g = (ngroups + 7) / 8;
  if (g > 148 * resident[dev]) g = 148 * resident[dev];
  layernorm16_ivit_kernel<NPL, R, XT, MINB><<<g, 256, 0, st>>>(x, ldx, out, rows, s_in, 1.0f / s_in, bias_int, sf, mult,
                                                               qmin, qmax);
  return check_launch("layernorm16_ivit_kernel");
}

template <int NPL, int R, typename XT, int MINB>
static int launch_ln16_rows_ibert(const XT* x, int64_t ldx, int8_t* out, int rows, float s_in, const float* bias_int,
                                  const float* sf, const double* mult, int qmin, int qmax, float sp, int use_int_sqrt,
                                  cudaStream_t st) {
  static int resident[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev = dev < 0 || dev >= 16 ? 0 : dev;
  if (resident[dev] == 0) resident[dev] = resident_ctas(layernorm16_ibert_kernel<NPL, R, XT, MINB>, MINB);
  const int ngroups = (rows + R - 1) / R;
  int g = (ngroups + 7) / 8;
  if (g > 148 * resident[dev]) g = 148 * resident[dev];
  layernorm16_ibert_kernel<NPL, R, XT, MINB><<<g, 256, 0, st>>>(x, ldx, out, rows, s_in, 1.0f / s_in, bias_int, sf, mult,
                                                                qmin, qmax, sp, 1.f / sp, use_int_sqrt);
  return check_launch("layernorm16_ibert_kernel");
}

template <int GPL, typename XT>
static int launch_ln16_outer2(const XT* x, int64_t ldx, int8_t* out, int rows, float s_in, const float* bias_int,
                              const float* sf, const double* mult, int qmin, int qmax, cudaStream_t st) {
  static int resident[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev = dev < 0 || dev >= 16 ? 0 : dev;
  if (resident[dev] == 0) resident[dev] = resident_ctas(layernorm16_outer2_kernel<GPL, XT>, 2);
  const int ngroups = (rows + 15) / 16;
  int g = (ngroups + 7) / 8;
  if (g > 148 * resident[dev]) g = 148 * resident[dev];
  layernorm16_outer2_kernel<GPL, XT><<<g, 256, 0, st>>>(x, ldx, out, rows, s_in, 1.0f / s_in, bias_int, sf, mult, qmin,
                                                        qmax);
  return check_launch("layernorm16_outer2_kernel");
}

template <typename XT>
static int launch_ln16(int kind, const XT* x, int64_t ldx, int8_t* out, int rows, int C, float s_in,
                       const float* bias_int, const float* sf, const double* mult, int qmin, int qmax, float shift,
                       int use_int_sqrt, int strided_tokens, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (kind != 0 && kind != 1) return fail(IVIT_ERR_BAD_ARG, "layernorm16: kind must be 0 (ivit) or 1 (ibert)");
  if (ldx < C) return fail(IVIT_ERR_BAD_ARG, "layernorm16: row stride (elements) must be >= C");
  if (qmin < -128 || qmax > 127) return fail(IVIT_ERR_UNSUPPORTED, "layernorm16: int8 outputs only");
  const float sp = ldexpf(1.f, static_cast<int>(shift));
  const int grid = (rows + 7) / 8;
  const bool outer = strided_tokens > 0;
  // I-ViT kind, scale validated for the two-FMA quotient: the restructured kernels
  if (kind == 0 && ln_variant() == 0 && xeff_fma_ok(s_in, static_cast<int>(8 * sizeof(XT)))) {
    const bool vec_ok = (ldx * sizeof(XT)) % 16 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
                        (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    if (outer && strided_tokens % 32 == 0 && vec_ok) {
      if (C == 96) return launch_ln16_outer2<3, XT>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
      if (C == 128) return launch_ln16_outer2<4, XT>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
    }
    if (!outer && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
      switch (C) {
        case 192: return launch_ln16_rows<6, 8, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
        case 256: return launch_ln16_rows<8, 4, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
        case 384: return launch_ln16_rows<12, 4, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
        case 512: return launch_ln16_rows<16, 2, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
        case 768: return launch_ln16_rows<24, 2, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, st);
        default: break;
      }
    }
  }
  if (kind == 1 && ln_variant() == 0 && !outer && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
      xeff_fma_ok(s_in, static_cast<int>(8 * sizeof(XT)))) {
    switch (C) {
      case 192: return launch_ln16_rows_ibert<6, 8, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
      case 384: return launch_ln16_rows_ibert<12, 4, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
      case 768: return launch_ln16_rows_ibert<24, 2, XT, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
      default: break;
    }
  }
  // Swin stage 0: C = 96, no tail tokens (L % 32 == 0), aligned rows -> the eight-lanes-per-row kernel
  if (outer && kind == 0 && C == 96 && strided_tokens % 32 == 0 && (ldx * sizeof(XT)) % 16 == 0 &&
      (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    const int quads = (rows + 3) / 4;
    static int resident = 0;          // persistent grid = the CTAs resident at once (one wave)
    if (resident == 0) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, layernorm96_strided_kernel<XT>, 256, 0) != cudaSuccess || nb < 1)
        nb = 3;
      resident = nb;
    }
    int g = (quads + 7) / 8;
    if (g > 148 * resident) g = 148 * resident;
    layernorm96_strided_kernel<XT><<<g, 256, 0, st>>>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax);
    return check_launch("layernorm96_strided_kernel");
  }
#define IVIT_LN16_CASE(c, ALLOW_OUTER)                                                                            \
  case c:                                                                                                         \
    if (outer) {                                                                                                  \
      if (!ALLOW_OUTER) return fail(IVIT_ERR_UNSUPPORTED, "layernorm16: channel-strided order for C <= 384 only");  \
      layernorm_x16_requant_kernel<c / 32, XT, (ALLOW_OUTER)><<<grid, 256, 0, st>>>(                               \
          kind, x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, 1.f / sp, use_int_sqrt, strided_tokens); \
    } else {                                                                                                      \
      layernorm_x16_requant_kernel<c / 32, XT, false><<<grid, 256, 0, st>>>(                                       \
          kind, x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, 1.f / sp, use_int_sqrt, 1);          \
    }                                                                                                             \
    break;
  switch (C) {
    IVIT_LN16_CASE(96, true)
    IVIT_LN16_CASE(128, true)
    IVIT_LN16_CASE(192, true)
    IVIT_LN16_CASE(256, true)
    IVIT_LN16_CASE(384, true)
    IVIT_LN16_CASE(512, false)
    IVIT_LN16_CASE(768, false)
    IVIT_LN16_CASE(1024, false)
    IVIT_LN16_CASE(1536, false)
    IVIT_LN16_CASE(2048, false)
    default:
      return fail(IVIT_ERR_UNSUPPORTED, "layernorm16: C must be one of 96,128,192,256,384,512,768,1024,1536,2048");
  }
#undef IVIT_LN16_CASE
  return check_launch("layernorm_x16_requant_kernel");
}

// int8 rows of the I-BERT kind through the restructured kernel (rowwise.cu's entry point asks here first);
// returns -1 when it does not apply (width, alignment, scale not accepted by the host check, variant 1)
namespace ivit {
int try_ln_ibert_rows_i8(const int8_t* x, int64_t ldx, int8_t* out, int rows, int C, float s_in, const float* bias_int,
                         const float* sf, const double* mult, int qmin, int qmax, float shift, int use_int_sqrt,
                         cudaStream_t st) {
  if (ln_variant() != 0 || qmin < -128 || qmax > 127 || (reinterpret_cast<uintptr_t>(out) & 15) != 0 ||
      !xeff_fma_ok(s_in, 8))
    return -1;
  const float sp = ldexpf(1.f, static_cast<int>(shift));
  switch (C) {
    case 192: return launch_ln16_rows_ibert<6, 8, int8_t, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
    case 384: return launch_ln16_rows_ibert<12, 4, int8_t, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
    case 768: return launch_ln16_rows_ibert<24, 2, int8_t, 2>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp, use_int_sqrt, st);
    default: return -1;
  }
}
}  // namespace ivit

extern "C" int ivit_xeff_fma_ok(float s, int bits) { return (bits == 8 || bits == 16) && xeff_fma_ok(s, bits) ? 1 : 0; }

extern "C" int ivit_layernorm_requant_i16_i8(int kind, const int16_t* x, int64_t ldx, int8_t* out, int rows, int C,
                                             float s_in, const float* bias_int, const float* sf, const double* mult,
                                             int qmin, int qmax, float shift, int use_int_sqrt, void* stream) {
  return launch_ln16<int16_t>(kind, x, ldx, out, rows, C, s_in, bias_int, sf, mult, qmin, qmax, shift, use_int_sqrt, 0,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int ivit_layernorm_requant_strided(int kind, const void* x, int x_bits, int64_t ldx, int8_t* out, int rows,
                                              int C, float s_in, const float* bias_int, const float* sf,
                                              const double* mult, int qmin, int qmax, float shift, int use_int_sqrt,
                                              int tokens_per_image, void* stream) {
  if (tokens_per_image <= 0) return fail(IVIT_ERR_BAD_ARG, "layernorm (strided order): tokens_per_image must be > 0");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (x_bits == 16)
    return launch_ln16<int16_t>(kind, static_cast<const int16_t*>(x), ldx, out, rows, C, s_in, bias_int, sf, mult, qmin,
                                qmax, shift, use_int_sqrt, tokens_per_image, st);
  if (x_bits == 8)
    return launch_ln16<int8_t>(kind, static_cast<const int8_t*>(x), ldx, out, rows, C, s_in, bias_int, sf, mult, qmin,
                               qmax, shift, use_int_sqrt, tokens_per_image, st);
  return fail(IVIT_ERR_BAD_ARG, "layernorm (strided order): x_bits must be 8 or 16");
}

extern "C" int ivit_fill_cls_rows_i16(int16_t* out, const int16_t* cls_row, int B, int N, int C, void* stream) {
  if (B * C <= 0) return 0;
  fill_cls_rows16_kernel<<<(B * C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(out, cls_row, B, N, C);
  return check_launch("fill_cls_rows16_kernel");
}

extern "C" int ivit_quantize_f32_codes(const float* x, void* out, int out_bits, float inv_scale, int64_t n, int qmin,
                                       int qmax, void* stream) {
  if (n <= 0) return 0;
  if (out_bits != 8 && out_bits != 16) return fail(IVIT_ERR_BAD_ARG, "quantize: out_bits must be 8 or 16");
  const int lim = out_bits == 16 ? 32768 : 128;
  if (qmin < -lim || qmax > lim - 1) return fail(IVIT_ERR_UNSUPPORTED, "quantize: clamp range exceeds the output type");
  long long g = (n + 255) / 256;
  if (g > 148LL * 16) g = 148LL * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_bits == 16)
    quantize_codes_kernel<int16_t><<<static_cast<int>(g), 256, 0, st>>>(x, static_cast<int16_t*>(out), inv_scale, n,
                                                                       static_cast<float>(qmin), static_cast<float>(qmax));
  else
    quantize_codes_kernel<int8_t><<<static_cast<int>(g), 256, 0, st>>>(x, static_cast<int8_t*>(out), inv_scale, n,
                                                                      static_cast<float>(qmin), static_cast<float>(qmax));
  return check_launch("quantize_codes_kernel");
}

extern "C" int ivit_requant_codes(const void* z, int z_bits, const float* s_in, const double* mult, int per_channel,
                                  const void* identity, int id_bits, int64_t id_rows, double mult_id, void* out,
                                  int out_bits, int64_t rows, int C, int qmin, int qmax, void* stream) {
  const long long n = rows * C;
  if (n <= 0) return 0;
  if (identity != nullptr && id_rows > 0 && rows % id_rows != 0)
    return fail(IVIT_ERR_BAD_ARG, "requant: id_rows must divide rows");
  if ((out_bits != 8 && out_bits != 16) || (identity != nullptr && id_bits != 8 && id_bits != 16))
    return fail(IVIT_ERR_BAD_ARG, "requant: out_bits / id_bits must be 8 or 16");
  const int lim = out_bits == 16 ? 32768 : 128;
  if (qmin < -lim || qmax > lim - 1) return fail(IVIT_ERR_UNSUPPORTED, "requant: clamp range exceeds the output type");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (z_bits) {
    case 32:
      return dispatch_requant<int32_t>(id_bits, out_bits, z, s_in, mult, per_channel, identity, id_rows, mult_id, out, rows,
                                       C, qmin, qmax, st);
    case 16:
      return dispatch_requant<int16_t>(id_bits, out_bits, z, s_in, mult, per_channel, identity, id_rows, mult_id, out, rows,
                                       C, qmin, qmax, st);
    case 8:
      return dispatch_requant<int8_t>(id_bits, out_bits, z, s_in, mult, per_channel, identity, id_rows, mult_id, out, rows,
                                      C, qmin, qmax, st);
    default:
      return fail(IVIT_ERR_BAD_ARG, "requant: z_bits must be 8, 16 or 32");
  }
}
