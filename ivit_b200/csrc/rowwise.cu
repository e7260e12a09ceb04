// Row-wise (HBM-bound) kernels of the I-ViT integer block: input quantiser (+im2col store
// layout), integer LayerNorm fused with the following QuantAct, ShiftGELU / I-BERT GELU as a
// per-layer (row-max, code) table, and small helpers.  One warp per row, 16-byte vector
// access, shuffle reductions.  All arithmetic through intmath.cuh (bit-exact replay).
#include "intmath.cuh"
#include "ivit_b200.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ivit {

static inline int cdiv(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// ------------------------------------------------------------------ input quantiser
// quant_utils.py:49 + :93: q = clamp(round(fl32(1/s) * x), -128, 127)   (reciprocal-multiply!)
__device__ __forceinline__ int quant8(float x, float inv_s) {
  const float q = rintf(__fmul_rn(inv_s, x));
  return static_cast<int>(fminf(fmaxf(q, -128.f), 127.f));
}

// fp32 image [B,Cin,H,W] -> int8 patch matrix [B*gh*gw, Cin*P*P] (k = c*P*P + i*P + j): the
// im2col of a kernel=stride=P convolution is a pure re-indexing, so it is folded into the
// quantiser's store layout.
// One CTA walks output rows (= patches); a thread owns 4 consecutive k of the row (one float4 of an
// image row).  32-bit index arithmetic, and shifts instead of divisions when P is a power of two
// (PLOG2 >= 0): the generic 64-bit div/mod version spent ~5x the memory time on index maths.
template <int PLOG2>
__global__ void __launch_bounds__(256)
quantize_patches_kernel(const float* __restrict__ x, int8_t* __restrict__ out, float inv_s, int B, int Cin, int H, int W,
                        int P) {
  const int gh = H / P, gw = W / P;
  const int K = Cin * P * P, K4 = K / 4;
  const int rows = B * gh * gw;
  // two rows per iteration: both float4 loads are in flight before the first store
  for (int row0 = blockIdx.x; row0 < rows; row0 += 2 * gridDim.x) {
    const float* img[2];
    uint32_t* orow[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int row = min(row0 + u * static_cast<int>(gridDim.x), rows - 1);
      const int pw = row % gw, t = row / gw;
      const int ph = t % gh, b = t / gh;
      img[u] = x + (static_cast<long long>(b) * Cin * H + ph * P) * W + pw * P;
      orow[u] = reinterpret_cast<uint32_t*>(out + static_cast<long long>(row) * K);
    }
    const bool two = row0 + static_cast<int>(gridDim.x) < rows;
    for (int k4 = threadIdx.x; k4 < K4; k4 += blockDim.x) {
      const int k = k4 * 4;
      int j, i, c;
      if (PLOG2 >= 0) {
        j = k & (P - 1); i = (k >> PLOG2) & (P - 1); c = k >> (2 * PLOG2);
      } else {
        j = k % P; i = (k / P) % P; c = k / (P * P);
      }
      const long long off = (static_cast<long long>(c) * H + i) * W + j;
      float4 v[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(img[u] + off));
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !two) break;
        const int q0 = quant8(v[u].x, inv_s), q1 = quant8(v[u].y, inv_s), q2 = quant8(v[u].z, inv_s),
                  q3 = quant8(v[u].w, inv_s);
        orow[u][k4] = (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | (static_cast<uint32_t>(q3) << 24);
      }
    }
  }
}

// Small patches (Swin's 4 x 4 stem: a 48-byte output row): one float4 of an image row per thread, fully
// coalesced reads; the float4 is row i of patch (y / P, x / P), i.e. 4 bytes at k = c P^2 + (y % P) P + x % P.
__global__ void __launch_bounds__(256)
quantize_patches_small_kernel(const float* __restrict__ x, int8_t* __restrict__ out, float inv_s, int B, int Cin, int H,
                              int W, int P) {
  const int W4 = W / 4, gw = W / P, gh = H / P, K = Cin * P * P;
  const long long total = static_cast<long long>(B) * Cin * H * W4;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x4 = static_cast<int>(t % W4);
    const int y = static_cast<int>((t / W4) % H);
    const int c = static_cast<int>((t / (static_cast<long long>(W4) * H)) % Cin);
    const int b = static_cast<int>(t / (static_cast<long long>(W4) * H * Cin));
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + t);
    const int xx = 4 * x4;
    const long long row = (static_cast<long long>(b) * gh + y / P) * gw + xx / P;
    const int k = c * P * P + (y % P) * P + xx % P;
    const int q0 = quant8(v.x, inv_s), q1 = quant8(v.y, inv_s), q2 = quant8(v.z, inv_s), q3 = quant8(v.w, inv_s);
    *reinterpret_cast<uint32_t*>(out + row * K + k) =
        (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | (static_cast<uint32_t>(q3) << 24);
  }
}

// uint8 images (what a data loader holds before torchvision's ToTensor + Normalize): a pixel has 256 values per
// channel, so ToTensor (/255), Normalize ((x - mean) / std) and the input QuantAct collapse into one [Cin][256]
// code table built by the caller in float32 (the reference's op order); the kernel is a byte gather into the
// im2col layout.  A quarter of the float32 path's input bytes: what the host->device link carries per step.
__global__ void __launch_bounds__(256)
quantize_patches_u8_kernel(const uint8_t* __restrict__ x, int8_t* __restrict__ out, const int8_t* __restrict__ lut, int B,
                           int Cin, int H, int W, int P) {
  extern __shared__ uint8_t slut[];                 // [Cin][256]
  for (int i = threadIdx.x; i < Cin * 256; i += blockDim.x) slut[i] = static_cast<uint8_t>(lut[i]);
  __syncthreads();
  const int W4 = W / 4, gw = W / P, gh = H / P, K = Cin * P * P;
  const long long total = static_cast<long long>(B) * Cin * H * W4;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x4 = static_cast<int>(t % W4);
    const int y = static_cast<int>((t / W4) % H);
    const int c = static_cast<int>((t / (static_cast<long long>(W4) * H)) % Cin);
    const int b = static_cast<int>(t / (static_cast<long long>(W4) * H * Cin));
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(x) + t);
    const uint8_t* l = slut + c * 256;
    const int xx = 4 * x4;
    const long long row = (static_cast<long long>(b) * gh + y / P) * gw + xx / P;
    const int k = c * P * P + (y % P) * P + xx % P;
    *reinterpret_cast<uint32_t*>(out + row * K + k) =
        l[v & 0xffu] | (l[(v >> 8) & 0xffu] << 8) | (l[(v >> 16) & 0xffu] << 16) | (static_cast<uint32_t>(l[v >> 24]) << 24);
  }
}

// Plain elementwise quantiser (QuantAct without a predecessor scale on an arbitrary tensor).
__global__ void quantize_kernel(const float* __restrict__ x, int8_t* __restrict__ out, float inv_s, long long n) {
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < n;
       t += static_cast<long long>(gridDim.x) * blockDim.x)
    out[t] = static_cast<int8_t>(quant8(x[t], inv_s));
}

// ------------------------------------------------------------------ LayerNorm + QuantAct
// Lane t owns elements {32*i + t} of a row: that is torch's CPU accumulator layout
// (intmath.cuh TorchSum32), so the tie-break slow path needs no data exchange.  Rows are
// staged through shared memory with 16-byte loads and read back byte-wise (conflict free).

// ivit_modules.py:30-65 followed by QuantAct (quant_modules.py:379-387, quant_utils.py:220-251).
// per-channel params: bias_int[c] = floor((b/w)/(sqrt(C)/2^30)), sf[c] = sqrt(C)/2^30*w[c],
// mult[c] = m*2^-e of sf[c]/s_out.
//
// Persistent grid; a warp owns R rows per iteration.  What keeps it off the ALU roof:
//  * x_eff(q) = fl32(fl32(q*s)/s) and its truncation have 256 possible values: per-block LUT.
//  * per-channel constants live in registers for the warp's whole lifetime.
//  * the ten float32 Newton steps run once per R rows (lane r iterates row r's variance).
//  * QuantAct: exact value is rint(z*mult) with z = rint(fl32(fl32(v*sf)/sf)), |z - v| <= |v| 2^-22.
//    Inside the clamp range |v*mult| < 130 the float32 product v*fl32(mult) is within 5e-5 of
//    z*mult, so unless it lies within 1e-4 of a rounding tie (p ~ 2e-4) rintf of it IS the
//    exact answer; outside the range both clamp.  Near ties take the replay path.
// Exact replay of the QuantAct that follows the LayerNorm for one element (out of line: it runs for
// ~2e-4 of the elements, and inlining it 48 times made the loop body miss the instruction cache).
__device__ __noinline__ int ln_requant_exact(float v, float sc, double mult, int qmin, int qmax) {
  const float xo = __fmul_rn(v, sc);                                                            // :63
  const float z = rintf(__fdiv_rn(xo, sc));                                                     // quant_utils.py:220
  const double rd = requant_d(static_cast<double>(z), mult);                                    // :229-230
  return static_cast<int>(fmin(fmax(rd, static_cast<double>(qmin)), static_cast<double>(qmax)));
}

// float32 sum of one row's un-rounded x/s in torch's CPU order (exact-tie rows only; out of line).
template <int NPL>
__device__ __noinline__ float ln_torch_row_sum(const float (&xe)[NPL], int lane) {
  TorchSum32 acc;
#pragma unroll
  for (int i = 0; i < NPL; ++i) acc.add(xe[i]);
  return torch_sum_finish(acc.level_sum(), 0.f, nullptr, 0, lane);
}

template <int NPL, int R, int MINB>
__global__ void __launch_bounds__(256, MINB)
layernorm_ivit_requant_kernel(const int8_t* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows,
                              float s_in, const float* __restrict__ bias_int, const float* __restrict__ sf,
                              const double* __restrict__ mult, int qmin, int qmax) {
  constexpr int C = NPL * 32;
  constexpr int VPR = C / 16;                       // 16-byte vectors per row
  __shared__ __align__(16) int8_t stage[8][2][R][C];     // double buffered: cp.async prefetch of the next group
  __shared__ float lut_f[256];
  __shared__ int lut_i[256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  {
    const float xe = x_eff(static_cast<float>(static_cast<int>(threadIdx.x) - 128), s_in);   // :36
    lut_f[threadIdx.x] = xe;
    lut_i[threadIdx.x] = static_cast<int>(xe);                                               // :38 trunc
  }
  float bia[NPL], mf[NPL];
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    bia[i] = __ldg(bias_int + 32 * i + lane);
    mf[i] = static_cast<float>(__ldg(mult + 32 * i + lane));
  }
  const bool fast_ok = qmin >= -128 && qmax <= 127;
  const int ngroups = (rows + R - 1) / R;
  auto prefetch = [&](int g, int buf) {
    if (g < ngroups) {
      const int row0 = g * R, nr = min(R, rows - row0);
      for (int v = lane; v < R * VPR; v += 32) {
        const int r = v / VPR, vv = v - r * VPR;
        if (r < nr) {
          const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(&stage[warp][buf][r][vv * 16]));
          const int8_t* src = x + static_cast<long long>(row0 + r) * ldx + vv * 16;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  prefetch(blockIdx.x * 8 + warp, 0);
  __syncthreads();
  int buf = 0;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8, buf ^= 1) {
    const int row0 = g * R;
    const int nr = min(R, rows - row0);
    prefetch(g + gridDim.x * 8, buf ^ 1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    // Phase-ordered (all R rows per phase, branch-free main path) so that independent rows and
    // elements interleave in the instruction stream; the rare paths are warp-level exceptions.
    int y[R][NPL];
    int S[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      S[r] = 0;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] = static_cast<int>(stage[warp][buf][r][32 * i + lane]) + 128;     // LUT index for now
        S[r] += y[r][i];
      }
      S[r] -= 128 * NPL;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] += __shfl_xor_sync(0xffffffffu, S[r], o);
    int mean[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      // :37 mean = round(mean(x_int)).  Only an exact tie S/C = k + 1/2 can be moved by the
      // float32 noise of the un-rounded inputs; then replay torch's summation order.
      float sum_f = static_cast<float>(S[r]);
      if (((S[r] % C) + C) % C * 2 == C) {                      // warp-uniform, rare
        float xe[NPL];
#pragma unroll
        for (int i = 0; i < NPL; ++i) xe[i] = lut_f[y[r][i]];
        sum_f = ln_torch_row_sum<NPL>(xe, lane);
      }
      mean[r] = static_cast<int>(rintf(__fdiv_rn(sum_f, static_cast<float>(C))));
    }
    long long var[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      long long vr = 0;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] = lut_i[y[r][i]] - mean[r];                         // :38-40 (trunc, int32)
        vr += static_cast<long long>(y[r][i] * y[r][i]);            // :41 int32 square (wraps, H9), :42 int64 sum
      }
      var[r] = vr;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) var[r] += __shfl_xor_sync(0xffffffffu, var[r], o);
    // :45-49 ten float32 Newton steps from k = 2^16, once per R rows: lane r iterates row r
    float factor[R];
    {
      long long mv = var[0];
#pragma unroll
      for (int r = 1; r < R; ++r) mv = (lane == r) ? var[r] : mv;
      const float var_f = static_cast<float>(mv);
      float k = 65536.f;
#pragma unroll
      for (int it = 0; it < 10; ++it) k = floorf(__fmul_rn(__fadd_rn(k, floorf(__fdiv_rn(var_f, k))), 0.5f));
      const float fct = floorf(__fdiv_rn(2147483648.f, k));        // :51 (2^31-1 -> 2^31 in float32)
#pragma unroll
      for (int r = 0; r < R; ++r) factor[r] = __shfl_sync(0xffffffffu, fct, r);
    }
    uint32_t slow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      slow[r] = 0;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        // int -> float and float -> nearest int through the 1.5*2^23 magic constant (exact for
        // |.| < 2^22): conversions share the quarter-rate XU pipe
        const float yf = __fsub_rn(__int_as_float(0x4B400000 + y[r][i]), 12582912.f);
        float v = floorf(__fmul_rn(__fmul_rn(yf, factor[r]), 0.5f));                            // :52
        v = __fadd_rn(v, bia[i]);                                                               // :61
        const float ra = fminf(fmaxf(__fmul_rn(v, mf[i]), -200.f), 200.f);   // beyond +-130 both paths clamp
        const float tm = __fadd_rn(ra, 12582912.f);             // rint(ra) in the low mantissa bits
        const float rr = __fsub_rn(tm, 12582912.f);
        // near a rounding tie (or a NaN from a zero scale): exact replay below
        slow[r] |= (fabsf(__fsub_rn(ra, rr)) < 0.4999f) ? 0u : (1u << i);
        stage[warp][buf][r][32 * i + lane] = static_cast<int8_t>(clampi(__float_as_int(tm) - 0x4B400000, qmin, qmax));
      }
      if (!fast_ok) slow[r] = 0xffffffffu;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (slow[r] != 0) {
#pragma unroll
        for (int i = 0; i < NPL; ++i) {
          if ((slow[r] >> i) & 1u) {
            const int c = 32 * i + lane;
            const float yf = static_cast<float>(y[r][i]);
            float v = floorf(__fmul_rn(__fmul_rn(yf, factor[r]), 0.5f));                        // :52
            v = __fadd_rn(v, bia[i]);                                                           // :61
            stage[warp][buf][r][c] = static_cast<int8_t>(ln_requant_exact(v, __ldg(sf + c), __ldg(mult + c), qmin, qmax));
          }
        }
      }
    }
    __syncwarp();
    for (int v = lane; v < R * VPR; v += 32) {
      const int r = v / VPR, vv = v - r * VPR;
      if (r < nr)
        *reinterpret_cast<int4*>(out + static_cast<long long>(row0 + r) * C + vv * 16) =
            *reinterpret_cast<const int4*>(&stage[warp][buf][r][vv * 16]);
    }
    __syncwarp();
  }
}

// Same computation for C % 128 == 0 with a packed lane mapping: a lane owns the four consecutive
// channels 128 j + 4 lane + {0..3} of every 128-channel group, so a row comes in as NJ 32-bit
// shared-memory loads per lane and leaves as NJ coalesced 4-byte global stores per lane (no output
// staging), codes are unpacked with one sign-replicating PRMT and packed with cvt.pack.sat.
// The near-tie test is two FMAs: rint(ra (1 -+ 2^-18) + 1.5*2^23) must agree (the exact value lies
// between them: float32(mult), z vs v and the product rounding are each < 2^-21.5 relative);
// where they differ -- ~1e-3 of the elements -- the exact replay runs for that element.
__device__ __forceinline__ int sext_byte(uint32_t w, int k) {
  uint32_t d;
  const uint32_t sel = 0x8880u + 0x1111u * static_cast<uint32_t>(k);
  asm("prmt.b32 %0, %1, 0, %2;" : "=r"(d) : "r"(w), "r"(sel));
  return static_cast<int>(d);
}
__device__ __forceinline__ uint32_t pack_sat_s8_ln(int a, int b, uint32_t hi16) {
  uint32_t d;
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(hi16));
  return d;
}

template <int NJ, int R, int MINB>
__global__ void __launch_bounds__(256, MINB)
layernorm_ivit_requant4_kernel(const int8_t* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows,
                               float s_in, const float* __restrict__ bias_int, const float* __restrict__ sf,
                               const double* __restrict__ mult, int qmin, int qmax) {
  constexpr int NPL = NJ * 4;
  constexpr int C = NJ * 128;
  constexpr int VPR = C / 16;
  __shared__ __align__(16) int8_t stage[8][2][R][C];     // input rows, cp.async double buffer
  __shared__ float lut_f[256];
  // trunc(x_eff(code)) (:36-38) as int8, indexed by the code's UNSIGNED byte and 256-byte aligned: one PRMT builds the
  // shared address of an element's entry from the packed word (line base | byte), one LDS.S8 fetches it
  __shared__ __align__(256) int8_t lut_t[256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  {
    const float xe = x_eff(static_cast<float>(static_cast<int>(threadIdx.x) - 128), s_in);   // :36
    lut_f[threadIdx.x] = xe;
    lut_t[threadIdx.x ^ 0x80] = static_cast<int8_t>(static_cast<int>(xe));                   // :38 trunc; |.| <= 128
  }
  const uint32_t tb = smem_u32(lut_t);
  float bia[NPL], mf[NPL];
#pragma unroll
  for (int j = 0; j < NJ; ++j)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int c = 128 * j + 4 * lane + b;
      bia[4 * j + b] = __ldg(bias_int + c);
      mf[4 * j + b] = static_cast<float>(__ldg(mult + c));
    }
  const bool fast_ok = qmin == -128 && qmax == 127;       // cvt.pack.sat clamps to exactly this range
  const int ngroups = (rows + R - 1) / R;
  auto prefetch = [&](int g, int buf) {
    if (g < ngroups) {
      const int row0 = g * R, nr = min(R, rows - row0);
      for (int v = lane; v < R * VPR; v += 32) {
        const int r = v / VPR, vv = v - r * VPR;
        if (r < nr) {
          const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(&stage[warp][buf][r][vv * 16]));
          const int8_t* src = x + static_cast<long long>(row0 + r) * ldx + vv * 16;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  pdl_wait();                   // programmatic dependent launch: the rows are the predecessor's output
  prefetch(blockIdx.x * 8 + warp, 0);
  __syncthreads();
  int buf = 0;
  for (int g = blockIdx.x * 8 + warp; g < ngroups; g += gridDim.x * 8, buf ^= 1) {
    const int row0 = g * R;
    const int nr = min(R, rows - row0);
    prefetch(g + gridDim.x * 8, buf ^ 1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    int y[R][NPL];
    int S[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      S[r] = 0;
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(&stage[warp][buf][r][128 * j + 4 * lane]);
        S[r] = __dp4a(static_cast<int>(w), 0x01010101, S[r]);              // sum of the four codes
        y[r][4 * j + 0] = lds_s8(__byte_perm(w, tb, 0x7650));               // trunc(x_eff(code)) of each
        y[r][4 * j + 1] = lds_s8(__byte_perm(w, tb, 0x7651));
        y[r][4 * j + 2] = lds_s8(__byte_perm(w, tb, 0x7652));
        y[r][4 * j + 3] = lds_s8(__byte_perm(w, tb, 0x7653));
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) S[r] += __shfl_xor_sync(0xffffffffu, S[r], o);
    int mean[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      // :37 mean = round(mean(x_int)); an exact tie S/C = k + 1/2 replays torch's float32 summation
      // order over the un-rounded inputs (channel 32 i + lane, read back from the staged row)
      float sum_f = static_cast<float>(S[r]);
      if (((S[r] % C) + C) % C * 2 == C) {                      // warp-uniform, rare
        float xe[NPL];
#pragma unroll
        for (int i = 0; i < NPL; ++i) xe[i] = lut_f[static_cast<int>(stage[warp][buf][r][32 * i + lane]) + 128];
        sum_f = ln_torch_row_sum<NPL>(xe, lane);
      }
      mean[r] = static_cast<int>(rintf(__fdiv_rn(sum_f, static_cast<float>(C))));
    }
    int var[R];     // :41-42: |y| <= 258, so neither the int32 squares nor their sum over C <= 1024 channels wrap
#pragma unroll
    for (int r = 0; r < R; ++r) {
      int vr = 0;
#pragma unroll
      for (int i = 0; i < NPL; ++i) {
        y[r][i] -= mean[r];                                         // :38-40 (trunc done by the table, int32)
        vr += y[r][i] * y[r][i];
      }
      var[r] = vr;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) var[r] += __shfl_xor_sync(0xffffffffu, var[r], o);
    float factor[R];
    {
      int mv = var[0];
#pragma unroll
      for (int r = 1; r < R; ++r) mv = (lane == r) ? var[r] : mv;
      const float var_f = static_cast<float>(mv);
      float k = 65536.f;
#pragma unroll
      for (int it = 0; it < 10; ++it) k = floorf(__fmul_rn(__fadd_rn(k, floorf(__fdiv_rn(var_f, k))), 0.5f));   // :45-49
      const float fct = floorf(__fdiv_rn(2147483648.f, k));        // :51 (2^31-1 -> 2^31 in float32)
#pragma unroll
      for (int r = 0; r < R; ++r) factor[r] = __shfl_sync(0xffffffffu, fct, r);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      uint32_t lob[NPL];                  // 1.5 * 2^23 + code as float bits: the code is the low byte
      const float fh = __fmul_rn(factor[r], 0.5f);                  // fl(y * factor) / 2 == fl(y * (factor / 2)): exact scaling
#pragma unroll
      for (int j = 0; j < NJ; ++j) {       // the four channels of one output word; a near-tie is replayed inside its group
        float v[4];
        uint32_t x[4], any = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = 4 * j + e;
          const float yf = __fsub_rn(__int_as_float(0x4B400000 + y[r][i]), 12582912.f);
          v[e] = __fadd_rn(floorf(__fmul_rn(yf, fh)), bia[i]);                                  // :52, :61
          // clamp(rint(ra), -128, 127) == rint(clamp(ra, -128.4, 127.4)): the code then fits the low byte
          const float ra = fminf(fmaxf(__fmul_rn(v[e], mf[i]), -128.4f), 127.4f);
          const float lo = __fmaf_rn(ra, 0.999996185302734375f, 12582912.f);   // 1 - 2^-18
          const float hi = __fmaf_rn(ra, 1.000003814697265625f, 12582912.f);   // 1 + 2^-18
          x[e] = __float_as_uint(lo) ^ __float_as_uint(hi);                    // consecutive integers differ in bit 0
          any |= x[e];
          lob[i] = __float_as_uint(lo);
        }
        if ((any & 1u) || !fast_ok) {      // within 2^-18 of a rounding tie (~1e-3 of the elements): exact replay
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (!fast_ok || (x[e] & 1u)) {
              const int c = 128 * j + 4 * lane + e;
              lob[4 * j + e] = static_cast<uint32_t>(ln_requant_exact(v[e], __ldg(sf + c), __ldg(mult + c), qmin, qmax)) & 0xffu;
            }
        }
      }
      if (r < nr) {
        int8_t* orow = out + static_cast<long long>(row0 + r) * C + 4 * lane;
#pragma unroll
        for (int j = 0; j < NJ; ++j)
          *reinterpret_cast<uint32_t*>(orow + 128 * j) =
              __byte_perm(__byte_perm(lob[4 * j], lob[4 * j + 1], 0x0040), __byte_perm(lob[4 * j + 2], lob[4 * j + 3], 0x0040), 0x5410);
      }
    }
    __syncwarp();                 // the staged rows are overwritten by the prefetch two iterations on
  }
  pdl_launch_dependents();
}

// ibert_modules.py:112-158 followed by QuantAct.  shift_pow = 2^shift, inv_shift_pow = 2^-shift.
template <int NPL>
__global__ void __launch_bounds__(256)
layernorm_ibert_requant_kernel(const int8_t* __restrict__ x, long long ldx, int8_t* __restrict__ out, int rows, float s_in,
                               const float* __restrict__ bias_int, const float* __restrict__ sf,
                               const double* __restrict__ mult, int qmin, int qmax, float shift_pow,
                               float inv_shift_pow, int use_int_sqrt) {
  constexpr int C = NPL * 32;
  __shared__ __align__(16) int8_t stage[8][C];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= rows) return;
  const int8_t* xr = x + static_cast<long long>(row) * ldx;
  for (int v = lane; v < C / 16; v += 32)
    *reinterpret_cast<int4*>(&stage[warp][v * 16]) = *reinterpret_cast<const int4*>(xr + v * 16);
  __syncwarp();
  float xe[NPL];
  TorchSum32 acc;
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    xe[i] = x_eff(static_cast<float>(stage[warp][32 * i + lane]), s_in);   // :129
    acc.add(xe[i]);
  }
  const float sum_f = torch_sum_finish(acc.level_sum(), 0.f, nullptr, 0, lane);
  const float mean = rintf(__fdiv_rn(sum_f, static_cast<float>(C)));        // :130
  float y[NPL];
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
  const float factor = floorf(__fdiv_rn(2147483648.f, std_i));             // :146 (2^31)
  int8_t* orow = out + static_cast<long long>(row) * C;
#pragma unroll
  for (int i = 0; i < NPL; ++i) {
    const int c = 32 * i + lane;
    float v = floorf(__fmul_rn(__fmul_rn(y[i], factor), 0.5f));             // :147
    v = __fadd_rn(v, __ldg(bias_int + c));                                  // :154
    const float sc = __ldg(sf + c);
    const float xo = __fmul_rn(v, sc);                                      // :156
    const float z = rintf(__fdiv_rn(xo, sc));
    const double r = requant_d(static_cast<double>(z), __ldg(mult + c));
    const int code = static_cast<int>(fmin(fmax(r, static_cast<double>(qmin)), static_cast<double>(qmax)));
    stage[warp][c] = static_cast<int8_t>(code);
  }
  __syncwarp();
  for (int v = lane; v < C / 16; v += 32)
    *reinterpret_cast<int4*>(orow + v * 16) = *reinterpret_cast<const int4*>(&stage[warp][v * 16]);
}

// ------------------------------------------------------------------ GELU as a table
// ShiftGELU's output for one element depends only on (its int8 code, the row-max int8 code):
// ivit_modules.py:105-126, then mlp.qact1 (scalar requant + clamp).  I-BERT GELU
// (ibert_modules.py:220-235) depends on the code alone.  Both are tabulated once per layer at
// freeze time -- by this kernel, with the exact device arithmetic -- as lut[qmax+128][q+128]
// (I-BERT: every row identical), so the per-element work of the hot kernel is one byte lookup.
__global__ void gelu_lut_build_kernel(int8_t* __restrict__ lut, int mode, GeluIvitParams gi, GeluIbertParams gb,
                                      double mult_out, float s_gelu_out, int qmin, int qmax_out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 65536) return;
  const int q = (idx & 255) - 128, mx = (idx >> 8) - 128;
  float g;   // integer-valued (ivit) or float (ibert) GELU output "x_int"
  if (mode == 0) {
    g = (q <= mx) ? gelu_ivit(q, mx, gi) : 0.f;
  } else {
    g = gelu_ibert(q, gb);
  }
  // module returns x_int * scaling_factor; QuantAct recovers round((x_int*s)/s) (quant_utils.py:220)
  const float z = rintf(__fdiv_rn(__fmul_rn(g, s_gelu_out), s_gelu_out));
  const double r = requant_d(static_cast<double>(z), mult_out);
  lut[idx] = static_cast<int8_t>(static_cast<int>(fmin(fmax(r, (double)qmin), (double)qmax_out)));
}

// Raw (un-requantised) GELU integer output for tests / the generic path: int32 per table entry.
__global__ void gelu_ivit_raw_kernel(float* __restrict__ out, GeluIvitParams gi) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 65536) return;
  const int q = (idx & 255) - 128, mx = (idx >> 8) - 128;
  out[idx] = (q <= mx) ? gelu_ivit(q, mx, gi) : 0.f;
}

// One warp per row: row max (needed by ShiftGELU, ivit_modules.py:110), then table lookup.
// The row's 256-byte table line is staged per warp in shared memory, 256-byte aligned and rotated by 128, so
// that the shared address of a code's entry is (line base | code byte): ONE PRMT builds it from the packed
// word, one LDS.U8 fetches it, three PRMTs re-pack four results -- 2.75 instructions per byte (the pass runs
// in place on L2-resident rows, so instructions, not HBM, bound it).
template <int NV>                   // 16-byte vectors per lane: C <= NV * 512
__global__ void __launch_bounds__(256)
gelu_lut_apply_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ out, const int8_t* __restrict__ lut,
                      int rows, int C) {
  __shared__ __align__(256) uint8_t srow[8][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  pdl_wait();                   // programmatic dependent launch: x is the predecessor's output
  if (row >= rows) return;
  const int4* xr = reinterpret_cast<const int4*>(x + static_cast<long long>(row) * C);
  int4* orow = reinterpret_cast<int4*>(out + static_cast<long long>(row) * C);
  const int nv = C / 16;
  int4 v[NV];                      // NV = 6: up to C = 3072 (6 * 32 lanes * 16 B)
  uint32_t m = 0x80808080u;        // four int8 lanes of -128
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = i * 32 + lane;
    if (idx < nv) {
      v[i] = __ldg(xr + idx);
      m = __vmaxs4(m, v[i].x); m = __vmaxs4(m, v[i].y); m = __vmaxs4(m, v[i].z); m = __vmaxs4(m, v[i].w);
    }
  }
  int mx = max(max(static_cast<int8_t>(m), static_cast<int8_t>(m >> 8)),
               max(static_cast<int8_t>(m >> 16), static_cast<int8_t>(m >> 24)));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  // entry of code c at byte (c & 0xff): the line rotated by 128
  *reinterpret_cast<uint2*>(&srow[warp][(8 * lane + 128) & 255]) =
      __ldg(reinterpret_cast<const uint2*>(lut + (mx + 128) * 256) + lane);
  __syncwarp();
  const uint32_t tb = smem_u32(&srow[warp][0]);      // low byte zero
  auto look4 = [&](uint32_t w) -> uint32_t {
    uint32_t r0, r1, r2, r3;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r0) : "r"(__byte_perm(w, tb, 0x7650)));
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r1) : "r"(__byte_perm(w, tb, 0x7651)));
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r2) : "r"(__byte_perm(w, tb, 0x7652)));
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r3) : "r"(__byte_perm(w, tb, 0x7653)));
    return __byte_perm(__byte_perm(r0, r1, 0x0040), __byte_perm(r2, r3, 0x0040), 0x5410);
  };
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int idx = i * 32 + lane;
    if (idx < nv) {
      int4 o;
      o.x = look4(v[i].x); o.y = look4(v[i].y); o.z = look4(v[i].z); o.w = look4(v[i].w);
      orow[idx] = o;
    }
  }
}

// ------------------------------------------------------------------ small helpers
// Row 0 of every image = the (batch independent) class-token row, computed at freeze time.
__global__ void fill_cls_rows_kernel(int8_t* __restrict__ out, const int8_t* __restrict__ cls_row, int B, int N,
                                     int C) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B * C) return;
  const int b = t / C, c = t % C;
  out[(static_cast<long long>(b) * N) * C + c] = cls_row[c];
}

// Stand-alone QuantAct on integer codes (un-fused module path and tests):
// out = clamp(rint(z*mult[c]) [+ rint(id*mult_id)])   quant_utils.py:220-251
// z int32 (accumulators) or int8 (codes of a previous QuantAct); identity int8 [id_rows, C] with
// id_rows dividing rows (positional embedding broadcast over the batch: row % id_rows).
template <typename ZT>
__global__ void requant_kernel(const ZT* __restrict__ z, const float* __restrict__ s_in,
                               const double* __restrict__ mult, int per_channel, const int8_t* __restrict__ id,
                               long long id_rows, double mult_id, int8_t* __restrict__ out, long long rows, int C,
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
    out[t] = static_cast<int8_t>(static_cast<int>(fmin(fmax(r, (double)qmin), (double)qmax)));
  }
}

// codes * scale -> float32 (what the reference hands to un-quantised consumers)
__global__ void dequant_i8_kernel(const int8_t* __restrict__ q, float s, float* __restrict__ out, long long n) {
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < n;
       t += static_cast<long long>(gridDim.x) * blockDim.x)
    out[t] = __fmul_rn(static_cast<float>(q[t]), s);
}

}  // namespace ivit

// =============================================================================== C ABI
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ivit {
thread_local char g_err[256] = "";
int fail(int code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
  }
  return 0;
}
static int ew_grid(long long n, int block) {
  long long g = (n + block - 1) / block;
  const long long cap = 148LL * 16;   // grid-stride: a few waves of the 148 SMs
  return static_cast<int>(g < cap ? (g > 0 ? g : 1) : cap);
}
}  // namespace ivit

using namespace ivit;

extern "C" int ivit_abi_version(void) { return 2; }   // 2: round-2 fields appended to the attention / window-attention / GEMM epilogue structs
extern "C" const char* ivit_last_error(void) { return g_err; }
extern "C" int ivit_device_sm(int* major, int* minor) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
  cudaDeviceProp p;
  e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
  *major = p.major;
  *minor = p.minor;
  return 0;
}

extern "C" int ivit_quantize_f32_i8(const float* x, int8_t* out, float inv_scale, int64_t n, void* stream) {
  if (n <= 0) return 0;
  quantize_kernel<<<ew_grid(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, out, inv_scale, n);
  return check_launch("quantize_kernel");
}

extern "C" int ivit_quantize_patches_f32_i8(const float* x, int8_t* out, float inv_scale, int B, int Cin, int H,
                                            int W, int P, void* stream) {
  if (P % 4 != 0 || H % P != 0 || W % P != 0) return fail(IVIT_ERR_BAD_ARG, "patch size must divide H, W; P % 4 == 0");
  const long long rows = static_cast<long long>(B) * (H / P) * (W / P);
  if (rows <= 0) return 0;
  if (rows > 0x7fffffffLL || static_cast<long long>(Cin) * P * P > 0x7fffffffLL)
    return fail(IVIT_ERR_UNSUPPORTED, "quantize_patches: more than 2^31 patches");
  const int grid = static_cast<int>(rows < 148LL * 32 ? rows : 148LL * 32);
  const int k4 = Cin * P * P / 4;
  const int threads = k4 >= 256 ? 256 : (k4 + 31) / 32 * 32;      // one pass over a row where it fits
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (P <= 8) {          // short patch rows: the per-patch-row CTA walk leaves most lanes idle
    const long long total = static_cast<long long>(B) * Cin * H * (W / 4);
    const long long g = (total + 255) / 256;
    quantize_patches_small_kernel<<<static_cast<int>(g < 148LL * 32 ? g : 148LL * 32), 256, 0, st>>>(x, out, inv_scale, B,
                                                                                                 Cin, H, W, P);
    return check_launch("quantize_patches_small_kernel");
  }
  if (P == 16) quantize_patches_kernel<4><<<grid, threads, 0, st>>>(x, out, inv_scale, B, Cin, H, W, P);
  else if (P == 8) quantize_patches_kernel<3><<<grid, threads, 0, st>>>(x, out, inv_scale, B, Cin, H, W, P);
  else if (P == 4) quantize_patches_kernel<2><<<grid, threads, 0, st>>>(x, out, inv_scale, B, Cin, H, W, P);
  else quantize_patches_kernel<-1><<<grid, threads, 0, st>>>(x, out, inv_scale, B, Cin, H, W, P);
  return check_launch("quantize_patches_kernel");
}

extern "C" int ivit_quantize_patches_u8_i8(const uint8_t* x, int8_t* out, const int8_t* lut, int B, int Cin, int H, int W,
                                           int P, void* stream) {
  if (P % 4 != 0 || H % P != 0 || W % P != 0) return fail(IVIT_ERR_BAD_ARG, "patch size must divide H, W; P % 4 == 0");
  if (Cin < 1 || Cin > 16) return fail(IVIT_ERR_UNSUPPORTED, "quantize_patches_u8: 1..16 input channels");
  if ((reinterpret_cast<uintptr_t>(x) & 3) != 0) return fail(IVIT_ERR_BAD_ARG, "quantize_patches_u8: 4-byte aligned images");
  const long long total = static_cast<long long>(B) * Cin * H * (W / 4);
  if (total <= 0) return 0;
  const long long g = (total + 255) / 256;
  quantize_patches_u8_kernel<<<static_cast<int>(g < 148LL * 16 ? g : 148LL * 16), 256, Cin * 256,
                               static_cast<cudaStream_t>(stream)>>>(x, out, lut, B, Cin, H, W, P);
  return check_launch("quantize_patches_u8_kernel");
}

// 0: packed-lane kernel where C % 128 == 0 (default); 1: strided-lane kernel always (A/B, tests)
static int g_pdl = 0;
namespace ivit {
int pdl_enabled() { return g_pdl; }
}
extern "C" int ivit_set_pdl(int on) {
  const int prev = g_pdl;
  g_pdl = on != 0;
  return prev;
}

static int g_ln_variant = 0;
namespace ivit {
int ln_variant() { return g_ln_variant; }
int try_ln_ibert_rows_i8(const int8_t* x, int64_t ldx, int8_t* out, int rows, int C, float s_in, const float* bias_int,
                         const float* sf, const double* mult, int qmin, int qmax, float shift, int use_int_sqrt,
                         cudaStream_t st);        // rowwise16.cu
}
extern "C" int ivit_layernorm_set_variant(int variant) {
  const int prev = g_ln_variant;
  g_ln_variant = variant;
  return prev;
}

template <int NPL>
static int launch_ln(int kind, const int8_t* x, long long ldx, int8_t* out, int rows, float s_in, const float* bias_int,
                     const float* sf, const double* mult, int qmin, int qmax, float shift, int use_int_sqrt,
                     cudaStream_t st) {
  const int grid = (rows + 7) / 8;
  if (kind == 0) {
    constexpr int R = NPL <= 12 ? 4 : 2;
    const int groups = (rows + R - 1) / R;
    const int pgrid = min((groups + 7) / 8, 148 * 4);     // persistent: a few CTAs per SM
    if constexpr (NPL % 4 == 0) {
      if (g_ln_variant == 0) {
        // persistent grid = exactly the CTAs that are resident at once (122 registers: two per SM): with the
        // former 4 per SM the second wave started with cold pipelines (0.859 -> 0.829 ms per DeiT-S step)
        static int resident = 0;
        if (resident == 0) {
          int nb = 0;
          if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, layernorm_ivit_requant4_kernel<NPL / 4, R, 2>, 256, 0) !=
                  cudaSuccess || nb < 1)
            nb = 2;
          resident = nb;
        }
        const int pgrid4 = min((groups + 7) / 8, 148 * resident);
        cudaError_t le = launch_kernel(layernorm_ivit_requant4_kernel<NPL / 4, R, 2>, dim3(pgrid4), dim3(256), 0, st, x, ldx,
                                       out, rows, s_in, bias_int, sf, mult, qmin, qmax);
        if (le != cudaSuccess) return fail(static_cast<int>(le), cudaGetErrorString(le));
        return check_launch("layernorm_requant4_kernel");
      }
    }
    static int resident_s = 0;
    if (resident_s == 0) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, layernorm_ivit_requant_kernel<NPL, R, 2>, 256, 0) != cudaSuccess ||
          nb < 1)
        nb = 2;
      resident_s = nb;
    }
    layernorm_ivit_requant_kernel<NPL, R, 2><<<min(pgrid, 148 * resident_s), 256, 0, st>>>(x, ldx, out, rows, s_in, bias_int,
                                                                                          sf, mult, qmin, qmax);
  } else {
    const float sp = ldexpf(1.f, static_cast<int>(shift));
    layernorm_ibert_requant_kernel<NPL><<<grid, 256, 0, st>>>(x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, sp,
                                                             1.f / sp, use_int_sqrt);
  }
  return check_launch("layernorm_requant_kernel");
}

extern "C" int ivit_layernorm_requant_i8(int kind, const int8_t* x, int64_t ldx, int8_t* out, int rows, int C, float s_in,
                                         const float* bias_int, const float* sf, const double* mult, int qmin,
                                         int qmax, float shift, int use_int_sqrt, void* stream) {
  if (rows <= 0) return 0;
  if (ldx % 16 != 0 || ldx < C) return fail(IVIT_ERR_BAD_ARG, "layernorm: row stride must be >= C and a multiple of 16");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (kind == 1) {     // I-BERT kind: several rows per warp, x/s as two fused operations (rowwise16.cu) where that applies
    const int rc = try_ln_ibert_rows_i8(x, ldx, out, rows, C, s_in, bias_int, sf, mult, qmin, qmax, shift, use_int_sqrt, st);
    if (rc != -1) return rc;
  }
#define IVIT_LN_CASE(c)                                                                                          \
  case c:                                                                                                        \
    return launch_ln<c / 32>(kind, x, ldx, out, rows, s_in, bias_int, sf, mult, qmin, qmax, shift, use_int_sqrt, st);
  switch (C) {
    IVIT_LN_CASE(96)
    IVIT_LN_CASE(128)
    IVIT_LN_CASE(192)
    IVIT_LN_CASE(384)
    IVIT_LN_CASE(768)
    IVIT_LN_CASE(1024)
    default:
      return fail(IVIT_ERR_UNSUPPORTED, "layernorm: C must be one of 96,128,192,384,768,1024");
  }
#undef IVIT_LN_CASE
}

static GeluIvitParams make_gelu_ivit(float s_in, int output_bit) {
  GeluIvitParams g;
  g.s_in = s_in;
  const float s_sig = s_in * 1.702f;                 // ivit_modules.py:108 (float32 * python float)
  g.ex.x0 = floorf(-1.0f / s_sig);                   // :94
  g.ex.n = 23;
  g.ex.nx0 = 23.f * g.ex.x0;                         // :95
  g.two31m1 = 2147483648.f;                          // float32(2^31 - 1)
  g.inv_shift = 1.f / ldexpf(1.f, 31 - output_bit + 1);
  return g;
}

static GeluIbertParams make_gelu_ibert(float s_in) {
  GeluIbertParams g;
  g.s_in = s_in;
  const float k = 1.4142f;
  const double c0 = -0.2888, c1 = -1.769, c2 = 1.0 / c0;   // ibert_modules.py:194-195 (python floats)
  const float s = s_in / k;                                // :228
  g.b_int = floorf(static_cast<float>(c1) / s);            // :205
  g.c_int = floorf(static_cast<float>(c2) / (s * s));      // :206
  float s_erf = (s * s) * static_cast<float>(c0);          // :214
  s_erf = s_erf * 64.f;                                    // :216
  g.shift = floorf(1.0f / s_erf);                          // :230
  g.s_out = s_in * s_erf / 2.f;                            // :233
  return g;
}

extern "C" int ivit_gelu_lut_build(int kind, int8_t* lut, float s_in, int output_bit, double mult_out, int qmin,
                                   int qmax, float* s_gelu_out_host, void* stream) {
  GeluIvitParams gi = make_gelu_ivit(s_in, output_bit);
  GeluIbertParams gb = make_gelu_ibert(s_in);
  const float s_out = (kind == 0) ? s_in * (1.f / ldexpf(1.f, output_bit - 1)) : gb.s_out;   // ivit :121,:124
  if (s_gelu_out_host) *s_gelu_out_host = s_out;
  gelu_lut_build_kernel<<<256, 256, 0, static_cast<cudaStream_t>(stream)>>>(lut, kind, gi, gb, mult_out, s_out, qmin,
                                                                            qmax);
  return check_launch("gelu_lut_build_kernel");
}

extern "C" int ivit_gelu_ivit_raw_table(float* out65536, float s_in, int output_bit, void* stream) {
  gelu_ivit_raw_kernel<<<256, 256, 0, static_cast<cudaStream_t>(stream)>>>(out65536, make_gelu_ivit(s_in, output_bit));
  return check_launch("gelu_ivit_raw_kernel");
}

// Narrow rows (Swin stage 0 / 1: C = 384 / 768): R = 1536 / C rows per warp.  The R rows are one contiguous run of
// 1536 bytes -- three 16-byte vectors per lane, every lane busy, 1.5 KB in flight per warp (a 384-byte row per warp
// left the pass latency-bound at 3.2 TB/s) -- and the R row maxima travel through ONE shuffle reduction as the four
// bytes of a word (__vmaxs4).
template <int R>
__global__ void __launch_bounds__(256)
gelu_lut_apply_rows_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ out, const int8_t* __restrict__ lut,
                           int rows) {
  constexpr int NVR = 96 / R;                          // 16-byte vectors per row
  __shared__ __align__(256) uint8_t srow[8][R][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row0 = (static_cast<long long>(blockIdx.x) * 8 + warp) * R;
  pdl_wait();                   // programmatic dependent launch: x is the predecessor's output
  if (row0 >= rows) return;
  const int nr = static_cast<int>(min(static_cast<long long>(R), rows - row0));
  const int4* xr = reinterpret_cast<const int4*>(x + row0 * (NVR * 16));
  int4* orow = reinterpret_cast<int4*>(out + row0 * (NVR * 16));
  int4 v[3];
  int rid[3];
  uint32_t m = 0x80808080u;                            // byte r = running maximum of row r
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int t = i * 32 + lane;
    rid[i] = t / NVR;
    if (rid[i] < nr) {
      v[i] = __ldg(xr + t);
      uint32_t q = __vmaxs4(__vmaxs4(v[i].x, v[i].y), __vmaxs4(v[i].z, v[i].w));
      q = __vmaxs4(q, __byte_perm(q, 0, 0x1032));      // every byte = the maximum of the vector's 16 codes
      q = __vmaxs4(q, __byte_perm(q, 0, 0x2301));
      const uint32_t sel = 0xffu << (8 * rid[i]);
      m = __vmaxs4(m, (q & sel) | (0x80808080u & ~sel));
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = __vmaxs4(m, __shfl_xor_sync(0xffffffffu, m, o));
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int mx = static_cast<int8_t>(m >> (8 * r));
    // entry of code c at byte (c & 0xff): the line rotated by 128
    *reinterpret_cast<uint2*>(&srow[warp][r][(8 * lane + 128) & 255]) =
        __ldg(reinterpret_cast<const uint2*>(lut + (mx + 128) * 256) + lane);
  }
  __syncwarp();
  const uint32_t tb0 = smem_u32(&srow[warp][0][0]);    // low byte zero
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    if (rid[i] < nr) {
      const uint32_t tb = tb0 + 256u * static_cast<uint32_t>(rid[i]);
      auto look4 = [&](uint32_t w) -> uint32_t {
        uint32_t r0, r1, r2, r3;
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r0) : "r"(__byte_perm(w, tb, 0x7650)));
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r1) : "r"(__byte_perm(w, tb, 0x7651)));
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r2) : "r"(__byte_perm(w, tb, 0x7652)));
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r3) : "r"(__byte_perm(w, tb, 0x7653)));
        return __byte_perm(__byte_perm(r0, r1, 0x0040), __byte_perm(r2, r3, 0x0040), 0x5410);
      };
      int4 o;
      o.x = look4(v[i].x); o.y = look4(v[i].y); o.z = look4(v[i].z); o.w = look4(v[i].w);
      orow[i * 32 + lane] = o;
    }
  }
}

extern "C" int ivit_gelu_lut_apply_i8(const int8_t* x, int8_t* out, const int8_t* lut, int rows, int C,
                                      void* stream) {
  if (rows <= 0) return 0;
  if (C % 16 != 0 || C > 6144) return fail(IVIT_ERR_UNSUPPORTED, "gelu: C % 16 == 0 and C <= 6144 required");
  if ((C == 384 || C == 768) && g_ln_variant != 1 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    const int R = 1536 / C;
    const int grid = (rows + 8 * R - 1) / (8 * R);
    cudaError_t le = R == 4 ? launch_kernel(gelu_lut_apply_rows_kernel<4>, dim3(grid), dim3(256), 0,
                                            static_cast<cudaStream_t>(stream), x, out, lut, rows)
                            : launch_kernel(gelu_lut_apply_rows_kernel<2>, dim3(grid), dim3(256), 0,
                                            static_cast<cudaStream_t>(stream), x, out, lut, rows);
    if (le != cudaSuccess) return fail(static_cast<int>(le), cudaGetErrorString(le));
    return check_launch("gelu_lut_apply_rows_kernel");
  }
  if (C <= 3072)
  {
    cudaError_t le = launch_kernel(gelu_lut_apply_kernel<6>, dim3((rows + 7) / 8), dim3(256), 0,
                                   static_cast<cudaStream_t>(stream), x, out, lut, rows, C);
    if (le != cudaSuccess) return fail(static_cast<int>(le), cudaGetErrorString(le));
  }
  else      // ViT-L / Swin-B hidden widths (4096), Swin-L (6144)
    gelu_lut_apply_kernel<12><<<(rows + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, out, lut, rows, C);
  return check_launch("gelu_lut_apply_kernel");
}

extern "C" int ivit_requant_i32_i8(const int32_t* z, const float* s_in, const double* mult, int per_channel,
                                   const int8_t* identity, int64_t id_rows, double mult_id, int8_t* out, int64_t rows,
                                   int C, int qmin, int qmax, void* stream) {
  const long long n = rows * C;
  if (n <= 0) return 0;
  if (identity != nullptr && id_rows > 0 && rows % id_rows != 0) return fail(IVIT_ERR_BAD_ARG, "requant: id_rows must divide rows");
  requant_kernel<int32_t><<<ew_grid(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      z, s_in, mult, per_channel, identity, id_rows, mult_id, out, rows, C, qmin, qmax);
  return check_launch("requant_kernel<int32>");
}

extern "C" int ivit_requant_i8_i8(const int8_t* z, const float* s_in, const double* mult, int per_channel,
                                  const int8_t* identity, int64_t id_rows, double mult_id, int8_t* out, int64_t rows,
                                  int C, int qmin, int qmax, void* stream) {
  const long long n = rows * C;
  if (n <= 0) return 0;
  if (identity != nullptr && id_rows > 0 && rows % id_rows != 0) return fail(IVIT_ERR_BAD_ARG, "requant: id_rows must divide rows");
  requant_kernel<int8_t><<<ew_grid(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      z, s_in, mult, per_channel, identity, id_rows, mult_id, out, rows, C, qmin, qmax);
  return check_launch("requant_kernel<int8>");
}

extern "C" int ivit_dequant_i8_f32(const int8_t* q, float scale, float* out, int64_t n, void* stream) {
  if (n <= 0) return 0;
  dequant_i8_kernel<<<ew_grid(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(q, scale, out, n);
  return check_launch("dequant_i8_kernel");
}

extern "C" int ivit_fill_cls_rows_i8(int8_t* out, const int8_t* cls_row, int B, int N, int C, void* stream) {
  if (B * C <= 0) return 0;
  fill_cls_rows_kernel<<<(B * C + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(out, cls_row, B, N, C);
  return check_launch("fill_cls_rows_kernel");
}

// =============================================================================== no-allocation GELU
// ivit_gelu_i8_noalloc: the GELU table pass for callers that cannot allocate (CUDA-graph capture, tight serving
// loops).  No counterpart in the reference.  The scratch the two-call
// form (ivit_gelu_lut_build + ivit_gelu_lut_apply_i8) keeps in caller memory -- the layer's [256][256] table -- and
// the per-row maximum codes live in ONE file-scope device array, so nothing is allocated here, by torch or by
// the CUDA runtime.  Every GPU that loads the library holds the array (320 KB) until exit; every call shares it
// and finds the previous call's contents: each call rewrites all 65536 table entries and the `rows` maxima it
// reads before reading them (nothing has to start at zero), and two calls that overlap on the GPU (different
// streams, or concurrent host threads) race on it.
namespace ivit {
constexpr int kNoallocMaxRows = IVIT_GELU_NOALLOC_MAX_ROWS;
struct GeluNoallocScratch {
  int8_t lut[65536];
  int32_t rowmax[kNoallocMaxRows];
};
static __device__ __align__(256) unsigned char g_gelu_noalloc_scratch[sizeof(GeluNoallocScratch)];

// pass 1: one warp per row, row-max int8 code -> scratch
__global__ void __launch_bounds__(256)
gelu_noalloc_rowmax_kernel(const int8_t* __restrict__ x, int32_t* __restrict__ rowmax, int rows, int C) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = C / 16;
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const int4* xr = reinterpret_cast<const int4*>(x + static_cast<long long>(row) * C);
    uint32_t m = 0x80808080u;
    for (int idx = lane; idx < nv; idx += 32) {
      const int4 v = __ldg(xr + idx);
      m = __vmaxs4(m, v.x); m = __vmaxs4(m, v.y); m = __vmaxs4(m, v.z); m = __vmaxs4(m, v.w);
    }
    int mx = max(max(static_cast<int8_t>(m), static_cast<int8_t>(m >> 8)),
                 max(static_cast<int8_t>(m >> 16), static_cast<int8_t>(m >> 24)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) rowmax[row] = mx;
  }
}

// pass 2: the row's 256-byte table line staged per warp, byte lookups (as gelu_lut_apply_kernel)
__global__ void __launch_bounds__(256)
gelu_noalloc_apply_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ out, const int8_t* __restrict__ lut,
                          const int32_t* __restrict__ rowmax, int rows, int C) {
  __shared__ __align__(8) uint8_t srow[8][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = C / 16;
  for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
    const int mx = rowmax[row];
    __syncwarp();
    *reinterpret_cast<uint2*>(&srow[warp][8 * lane]) = *(reinterpret_cast<const uint2*>(lut + (mx + 128) * 256) + lane);
    __syncwarp();
    const uint8_t* trow = srow[warp];
    const int4* xr = reinterpret_cast<const int4*>(x + static_cast<long long>(row) * C);
    int4* orow = reinterpret_cast<int4*>(out + static_cast<long long>(row) * C);
    auto look4 = [&](uint32_t w) -> uint32_t {
      w ^= 0x80808080u;
      return trow[w & 0xffu] | (trow[(w >> 8) & 0xffu] << 8) | (trow[(w >> 16) & 0xffu] << 16) | (trow[w >> 24] << 24);
    };
    for (int idx = lane; idx < nv; idx += 32) {
      const int4 v = __ldg(xr + idx);
      int4 o;
      o.x = look4(v.x); o.y = look4(v.y); o.z = look4(v.z); o.w = look4(v.w);
      orow[idx] = o;
    }
  }
}

// Device address of the scratch array, fetched once per device with that device current.  In host code the
// array's name is only a handle for cudaGetSymbolAddress, never a pointer a kernel may be given.
static GeluNoallocScratch* gelu_noalloc_scratch(int* err) {
  constexpr int kMaxDev = 64;
  static GeluNoallocScratch* ptrs[kMaxDev] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess && (dev < 0 || dev >= kMaxDev)) e = cudaErrorInvalidDevice;
  if (e == cudaSuccess) {
    GeluNoallocScratch* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (p != nullptr) return p;
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_gelu_noalloc_scratch);
    if (e == cudaSuccess && sym != nullptr) {
      __atomic_store_n(&ptrs[dev], static_cast<GeluNoallocScratch*>(sym), __ATOMIC_RELEASE);
      return static_cast<GeluNoallocScratch*>(sym);
    }
    if (e == cudaSuccess) e = cudaErrorInvalidSymbol;
  }
  *err = fail(static_cast<int>(e), cudaGetErrorString(e));
  return nullptr;
}
}  // namespace ivit

extern "C" int ivit_gelu_i8_noalloc(int kind, const int8_t* x, int8_t* out, int rows, int C, float s_in, int output_bit,
                                    double mult_out, int qmin, int qmax, float* s_gelu_out_host, void* stream) {
  if (kind != 0 && kind != 1) return fail(IVIT_ERR_BAD_ARG, "gelu_noalloc: kind must be 0 (ivit) or 1 (ibert)");
  if (rows < 0) return fail(IVIT_ERR_BAD_ARG, "gelu_noalloc: rows < 0");
  if (rows > kNoallocMaxRows)
    return fail(IVIT_ERR_UNSUPPORTED, "gelu_noalloc: rows exceeds IVIT_GELU_NOALLOC_MAX_ROWS (65536), the size of the "
                                      "library's static scratch; use ivit_gelu_lut_build + ivit_gelu_lut_apply_i8");
  if (C % 16 != 0 || C <= 0 || C > 6144) return fail(IVIT_ERR_UNSUPPORTED, "gelu: C % 16 == 0 and C <= 6144 required");
  int err = 0;
  GeluNoallocScratch* sc = gelu_noalloc_scratch(&err);
  if (sc == nullptr) return err;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GeluIvitParams gi = make_gelu_ivit(s_in, output_bit);
  GeluIbertParams gb = make_gelu_ibert(s_in);
  const float s_out = (kind == 0) ? s_in * (1.f / ldexpf(1.f, output_bit - 1)) : gb.s_out;   // as ivit_gelu_lut_build
  if (s_gelu_out_host) *s_gelu_out_host = s_out;
  gelu_lut_build_kernel<<<256, 256, 0, st>>>(sc->lut, kind, gi, gb, mult_out, s_out, qmin, qmax);
  int rc = check_launch("gelu_lut_build_kernel (noalloc)");
  if (rc != 0 || rows == 0) return rc;
  const int grid = min((rows + 7) / 8, 148 * 8);
  gelu_noalloc_rowmax_kernel<<<grid, 256, 0, st>>>(x, sc->rowmax, rows, C);
  rc = check_launch("gelu_noalloc_rowmax_kernel");
  if (rc != 0) return rc;
  gelu_noalloc_apply_kernel<<<grid, 256, 0, st>>>(x, out, sc->lut, sc->rowmax, rows, C);
  return check_launch("gelu_noalloc_apply_kernel");
}
