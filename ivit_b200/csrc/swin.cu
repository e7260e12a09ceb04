// Swin-specific kernels of the I-ViT integer path (swin_quant.py): window attention with the
// quantised relative-position bias and the float shifted-window mask, the PatchMerging gather and
// the token average pool.  The window attention exists twice: `window_attention_tc_kernel` (head_dim
// 32, default) runs both QuantMatMul products on tcgen05 kind::i8 with hand-swizzled operand tiles,
// `window_attention_i8_kernel` (any head_dim % 4 == 0, and the cross-check) on dp4a.  Either way the
// kernel is bound by its per-score epilogue, not by the products.
//
// Roll and window partition / reverse (swin_quant.py:18-50, :258-289) are pure index maps: the
// kernel reads q / k / v of window token (iy, ix) of window (wy, wx) at image token
// ((wy ws + iy + shift) % res, (wx ws + ix + shift) % res) of the un-permuted [B, L, 3C] qkv
// buffer and writes its output row to the same token, so no permuted copy ever exists in HBM.
#include <cstdio>
#include <cstdlib>

#include "intmath.cuh"
#include "ivit_b200.h"
#include "ptx.cuh"

namespace ivit {
int fail(int code, const char* msg);
int check_launch(const char* what);

constexpr int WMAX = 64;      // tokens per window (7 x 7 = 49; up to 8 x 8)
constexpr int HDMAX = 64;

// torch.sum(x, -1) of a contiguous float32 row on CPU (ATen vectorized_inner_sum; see
// oracle/ivit_oracle.py::torch_sum_lastdim) for n < 512, evaluated by one thread.
__device__ float torch_row_sum_small(const float* e, int n) {
  const int nvec = n >> 3, size_ilp = nvec >> 2;
  float acc[32];
#pragma unroll
  for (int l = 0; l < 32; ++l) acc[l] = 0.f;
  for (int i = 0; i < size_ilp; ++i)          // size_ilp < 16: no cascade step is reached
#pragma unroll
    for (int l = 0; l < 32; ++l) acc[l] = __fadd_rn(acc[l], e[32 * i + l]);
  for (int v = size_ilp * 4; v < nvec; ++v)   // left-over whole vectors go to accumulator row 0
#pragma unroll
    for (int l = 0; l < 8; ++l) acc[l] = __fadd_rn(acc[l], e[8 * v + l]);
#pragma unroll
  for (int k = 1; k < 4; ++k)
#pragma unroll
    for (int l = 0; l < 8; ++l) acc[l] = __fadd_rn(acc[l], acc[8 * k + l]);
  float fin = 0.f;
  for (int t = nvec * 8; t < n; ++t) fin = __fadd_rn(fin, e[t]);
#pragma unroll
  for (int l = 0; l < 8; ++l) fin = __fadd_rn(fin, acc[l]);
  return fin;
}

struct WinAttnArgs {
  const int8_t* qkv;      // [B, L, 3, nH, hd]
  int8_t* out;            // [B, L, nH * hd]
  uint8_t* probs;         // optional tap [B * nW, nH, N, N]
  int res, ws, shift, nH, hd;
  ivit_window_attention_params_t p;
  float x0, nx0, out_shift;
};

// region label of a rolled coordinate for the shifted-window mask (swin_quant.py:226-236)
__device__ __forceinline__ int mask_region(int c, int res, int ws, int shift) {
  return c < res - ws ? 0 : (c < res - shift ? 1 : 2);
}

// a: four unsigned bytes, b: four signed bytes
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

constexpr int QK_LD = HDMAX + 4;      // row pitch of q / k in shared memory (words: odd multiple -> no bank conflicts)
constexpr int VT_LD = WMAX + 4;       // row pitch of v^T and of the probability rows

// TABLE: the Shiftmax exp term of an unmasked score depends only on (its qact2 code, the row-max code):
// gathered from the per-layer [256][256] table ivit_softmax_table_build makes with the very same device
// arithmetic, instead of ~45 instructions (two IEEE divisions) per score.  A masked score (carrier - 100)
// lies so far below the row maximum -- the diagonal is never masked -- that int_exp_shift clamps it to
// n x0 and returns exactly -x0; the host proves that bound for the layer's scale before choosing TABLE.
constexpr int kMasked = -1000;

template <bool TABLE>
__global__ void __launch_bounds__(256)
window_attention_i8_kernel(const WinAttnArgs a) {
  __shared__ __align__(16) int8_t sq[WMAX * QK_LD], sk[WMAX * QK_LD];
  __shared__ __align__(16) int8_t svt[HDMAX * VT_LD];        // v transposed: [d][token]
  __shared__ float sx[WMAX][WMAX + 1];
  __shared__ __align__(16) uint8_t sp[WMAX][VT_LD];
  __shared__ int tok[WMAX];
  __shared__ int lab[WMAX];
  __shared__ float srow[WMAX];                               // row max, then the row's factor
  const int ws = a.ws, N = ws * ws, hd = a.hd, res = a.res, nWx = res / ws, nW = nWx * nWx;
  const uint32_t magicN = 0xFFFFFFFFu / static_cast<uint32_t>(N) + 1u;   // t / N == umulhi(t, magicN) for t < 2^16
  const int h = blockIdx.x % a.nH;
  const int w = (blockIdx.x / a.nH) % nW;
  const int b = blockIdx.x / (a.nH * nW);
  const int C = a.nH * hd, L = res * res;
  const int wy = w / nWx, wx = w % nWx;
  const int N4 = (N + 3) & ~3;
  if (threadIdx.x < N) {
    const int iy = threadIdx.x / ws, ix = threadIdx.x % ws;
    const int ry = wy * ws + iy, rx = wx * ws + ix;                        // rolled coordinates
    // window order (module path: roll + window_partition already applied by the caller): row w * N + i of the image
    tok[threadIdx.x] = a.p.windowed ? w * N + static_cast<int>(threadIdx.x)
                                    : ((ry + a.shift) % res) * res + (rx + a.shift) % res;
    lab[threadIdx.x] = 3 * mask_region(ry, res, ws, a.shift) + mask_region(rx, res, ws, a.shift);
  }
  __syncthreads();
  // q, k rows and v^T of this (window, head): hd bytes per token, 4-byte words
  const int wpr = hd >> 2;
  for (int t = threadIdx.x; t < 3 * N * wpr; t += blockDim.x) {
    const int which = t / (N * wpr), r = (t / wpr) % N, c4 = t % wpr;
    const int8_t* src = a.qkv + (static_cast<long long>(b) * L + tok[r]) * 3 * C + which * C + h * hd + 4 * c4;
    const int v = __ldg(reinterpret_cast<const int*>(src));
    if (which < 2) {
      *reinterpret_cast<int*>((which == 0 ? sq : sk) + r * QK_LD + 4 * c4) = v;
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) svt[(4 * c4 + u) * VT_LD + r] = static_cast<int8_t>(v >> (8 * u));
    }
  }
  __syncthreads();
  // S = q k^T -> qact_attn1 -> qact2 (+ relative-position bias identity) -> (+ mask) -> x / s
  const int8_t* bias = a.p.bias_codes + static_cast<long long>(h) * N * N;
  const int32_t* bterms = a.p.bias_terms != nullptr ? a.p.bias_terms + static_cast<long long>(h) * N * N : nullptr;
  for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
    const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
    int acc = 0;
    const int* qi = reinterpret_cast<const int*>(sq + i * QK_LD);
    const int* kj = reinterpret_cast<const int*>(sk + j * QK_LD);
    for (int c = 0; c < wpr; ++c) acc = __dp4a(qi[c], kj[c], acc);
    const int c1 = clampi(requant_i32_fast(acc, a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn);        // qact_attn1
    const int bterm = bterms != nullptr ? __ldg(bterms + t) : requant_i32_fast(__ldg(bias + t), a.p.mult_bias);
    const int c2 = clampi(requant_i32_fast(c1, a.p.mult2) + bterm, a.p.qmin2, a.p.qmax2);            // qact2 (:147)
    if (TABLE) {
      sx[i][j] = __int_as_float((a.shift > 0 && lab[i] != lab[j]) ? kMasked : c2);
    } else {
      float x = __fmul_rn(static_cast<float>(c2), a.p.s2);                 // the float carrier
      if (a.shift > 0) x = __fadd_rn(x, lab[i] != lab[j] ? -100.f : 0.f);  // attn + mask (:151)
      sx[i][j] = __fdiv_rn(x, a.p.s2);                                     // ivit_modules.py:165
    }
  }
  __syncthreads();
  // Shiftmax (ivit_modules.py:164-179): row max | exp terms (all threads) | torch-order row sum | codes
  if (threadIdx.x < N) {
    const float* row = sx[threadIdx.x];
    if (TABLE) {
      int mx = __float_as_int(row[0]);
      for (int j = 1; j < N; ++j) mx = max(mx, __float_as_int(row[j]));
      srow[threadIdx.x] = __int_as_float(mx);
    } else {
      float mx = row[0];
      for (int j = 1; j < N; ++j) mx = fmaxf(mx, row[j]);
      srow[threadIdx.x] = mx;
    }
  }
  __syncthreads();
  if (TABLE) {
    const float e_mask = -a.x0;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
      const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
      const int c = __float_as_int(sx[i][j]);
      const float* trow = a.p.exp_table + (__float_as_int(srow[i]) + 128) * 256 + 128;
      sx[i][j] = c == kMasked ? e_mask : __ldg(trow + c);
    }
  } else {
    ExpShiftParams ep;
    ep.x0 = a.x0; ep.nx0 = a.nx0; ep.n = 15;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
      const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
      sx[i][j] = exp_shift(__fsub_rn(sx[i][j], srow[i]), ep);
    }
  }
  __syncthreads();
  if (threadIdx.x < N) {
    float sum = torch_row_sum_small(sx[threadIdx.x], N);
    sum = fminf(sum, 2147483648.f);                                        // :171-173
    srow[threadIdx.x] = __fmul_rn(floorf(__fdiv_rn(2147483648.f, sum)), a.out_shift);   // :174-175
  }
  __syncthreads();
  for (int t = threadIdx.x; t < N * N4; t += blockDim.x) {
    const int i = t / N4, j = t - i * N4;
    // floor of a value in [0, 2^22) through the 2^23 magic constant (round toward zero), not F2I
    sp[i][j] = j < N ? static_cast<uint8_t>(__float_as_uint(__fadd_rz(__fmul_rn(sx[i][j], srow[i]), 8388608.f)) & 0xffu)
                     : uint8_t(0);
  }
  __syncthreads();
  if (a.probs != nullptr) {
    uint8_t* pr = a.probs + (static_cast<long long>(b * nW + w) * a.nH + h) * N * N;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) pr[t] = sp[t / N][t % N];
  }
  // O = P v -> qact3, stored at the window token's image position (window_reverse + roll back)
  if (a.out != nullptr) {
    for (int t = threadIdx.x; t < N * hd; t += blockDim.x) {
      const int i = t / hd, d = t - i * hd;
      const uint32_t* pi = reinterpret_cast<const uint32_t*>(sp[i]);
      const int* vd = reinterpret_cast<const int*>(svt + d * VT_LD);
      int acc = 0;
      for (int j4 = 0; j4 < N4 / 4; ++j4) acc = dp4a_us(pi[j4], vd[j4], acc);
      const int c = clampi(requant_i32_fast(acc, a.p.mult_out), a.p.qmin_out, a.p.qmax_out);
      a.out[(static_cast<long long>(b) * L + tok[i]) * C + h * hd + d] = static_cast<int8_t>(c);
    }
  }
}

// ------------------------------------------------------------------ tensor-core variant (hd = 32)
// Same kernel with the two QuantMatMul products (quant_modules.py:404-409) on tcgen05 kind::i8:
//   S = Q K^T   one 128 x 64 x 64 tile: the window's 49 query rows in TMEM lanes 0..48, 49 key columns;
//               hd = 32 is zero-padded to the 64-byte rows of the SWIZZLE_64B K-major layout the ViT
//               attention kernel uses (two K = 32 MMAs, the second multiplies zeros)
//   O = P V     128 x 32 x 64: P (u8) rows of 64 bytes (columns past the window are zero), V^T [d][token]
// Operand tiles are written by the threads straight in the swizzled layout (the window rows are a
// gather, not a TMA box).  The score tile is copied TMEM -> shared memory by the two warps that can
// reach lanes 0..63 and the per-score epilogue then runs on all 256 threads exactly as above; the
// output rows are requantised by those two warps straight from TMEM.  64 TMEM columns per CTA.
__device__ __forceinline__ int sw64_off(int r, int byte) {       // 64-byte rows, 8-row groups 512 B apart
  return (r >> 3) * 512 + (r & 7) * 64 + ((((byte >> 4) ^ (r >> 1)) & 3) << 4) + (byte & 15);
}

template <bool TABLE>
__global__ void __launch_bounds__(256, 5)
window_attention_tc_kernel(const WinAttnArgs a) {
  __shared__ __align__(1024) uint8_t tiles[8192 + 4096 + 8192 + 2048];
  __shared__ float sx[WMAX][WMAX + 1];
  __shared__ int tok[WMAX];
  __shared__ int lab[WMAX];
  __shared__ float srow[WMAX];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  uint8_t* const sQ = tiles;                 // [128][64 B]  rows >= N never read back
  uint8_t* const sK = tiles + 8192;          // [64][64 B]
  uint8_t* const sP = tiles + 12288;         // [128][64 B]
  uint8_t* const sVt = tiles + 20480;        // [32][64 B]   v transposed: [d][token]
  const int ws = a.ws, N = ws * ws, res = a.res, nWx = res / ws, nW = nWx * nWx;
  const uint32_t magicN = 0xFFFFFFFFu / static_cast<uint32_t>(N) + 1u;   // t / N == umulhi(t, magicN) for t < 2^16
  constexpr int hd = 32;
  const int h = blockIdx.x % a.nH;
  const int w = (blockIdx.x / a.nH) % nW;
  const int b = blockIdx.x / (a.nH * nW);
  const int C = a.nH * hd, L = res * res;
  const int wy = w / nWx, wx = w % nWx;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 64);
    tmem_relinquish();
  }
  if (threadIdx.x < N) {
    const int iy = threadIdx.x / ws, ix = threadIdx.x % ws;
    const int ry = wy * ws + iy, rx = wx * ws + ix;                        // rolled coordinates
    // window order (module path: roll + window_partition already applied by the caller): row w * N + i of the image
    tok[threadIdx.x] = a.p.windowed ? w * N + static_cast<int>(threadIdx.x)
                                    : ((ry + a.shift) % res) * res + (rx + a.shift) % res;
    lab[threadIdx.x] = 3 * mask_region(ry, res, ws, a.shift) + mask_region(rx, res, ws, a.shift);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // gather q, k (K-major rows, bytes 32..63 zero) and v^T into the swizzled operand tiles
  constexpr int wpr = hd / 4;
  for (int t = threadIdx.x; t < 3 * N * wpr; t += blockDim.x) {
    const int rr = t / wpr, c4 = t % wpr;                                  // wpr is a compile-time 8
    const int which = static_cast<int>(__umulhi(static_cast<uint32_t>(rr), magicN)), r = rr - which * N;
    const int8_t* src = a.qkv + (static_cast<long long>(b) * L + tok[r]) * 3 * C + which * C + h * hd + 4 * c4;
    const int v = __ldg(reinterpret_cast<const int*>(src));
    if (which < 2) {
      *reinterpret_cast<int*>((which == 0 ? sQ : sK) + sw64_off(r, 4 * c4)) = v;
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) sVt[sw64_off(4 * c4 + u, r)] = static_cast<uint8_t>(v >> (8 * u));
    }
  }
  for (int t = threadIdx.x; t < 2 * N * 2; t += blockDim.x) {              // zero the K padding of q and k
    const int which = t / (2 * N), r = (t >> 1) % N, c = 2 + (t & 1);
    *reinterpret_cast<int4*>((which == 0 ? sQ : sK) + sw64_off(r, 16 * c)) = make_int4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    tc_fence_after();
    constexpr uint32_t idesc1 = make_idesc_i8(128, 64, true, true, false, false);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const uint64_t da = make_smem_desc(smem_u32(sQ) + k * 32, 16, 512, kLayoutSW64);
      const uint64_t db = make_smem_desc(smem_u32(sK) + k * 32, 16, 512, kLayoutSW64);
      mma_i8_ss(tmem_base, da, db, idesc1, k != 0 ? 1u : 0u);
    }
    tc_commit(&bars[0]);
  }
  __syncwarp();
  int (*sacc)[WMAX + 1] = reinterpret_cast<int (*)[WMAX + 1]>(sx);
  if (warp < 2) {                              // TMEM lanes 0..63 -> the score tile in shared memory
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const int row = 32 * warp + lane;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(32 * warp) << 16);
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {           // 16-column chunks: the kernel has to live in 48 registers
      uint32_t v[16];
      tmem_ld16(t_row + 16 * q4, v);
      tmem_ld_wait();
      if (row < N) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (16 * q4 + j < N) sacc[row][16 * q4 + j] = static_cast<int>(v[j]);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  // qact_attn1 -> qact2 (+ relative-position bias identity) -> (+ mask) -> x / s   (as in the kernel above)
  const int8_t* bias = a.p.bias_codes + static_cast<long long>(h) * N * N;
  const int32_t* bterms = a.p.bias_terms != nullptr ? a.p.bias_terms + static_cast<long long>(h) * N * N : nullptr;
  for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
    const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
    const int acc = sacc[i][j];
    const int c1 = clampi(requant_i32_fast(acc, a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn);
    const int bterm = bterms != nullptr ? __ldg(bterms + t) : requant_i32_fast(__ldg(bias + t), a.p.mult_bias);
    const int c2 = clampi(requant_i32_fast(c1, a.p.mult2) + bterm, a.p.qmin2, a.p.qmax2);
    if (TABLE) {
      sx[i][j] = __int_as_float((a.shift > 0 && lab[i] != lab[j]) ? kMasked : c2);
    } else {
      float x = __fmul_rn(static_cast<float>(c2), a.p.s2);
      if (a.shift > 0) x = __fadd_rn(x, lab[i] != lab[j] ? -100.f : 0.f);
      sx[i][j] = __fdiv_rn(x, a.p.s2);
    }
  }
  __syncthreads();
  if (threadIdx.x < N) {
    const float* row = sx[threadIdx.x];
    if (TABLE) {
      int mx = __float_as_int(row[0]);
      for (int j = 1; j < N; ++j) mx = max(mx, __float_as_int(row[j]));
      srow[threadIdx.x] = __int_as_float(mx);
    } else {
      float mx = row[0];
      for (int j = 1; j < N; ++j) mx = fmaxf(mx, row[j]);
      srow[threadIdx.x] = mx;
    }
  }
  __syncthreads();
  if (TABLE) {
    const float e_mask = -a.x0;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
      const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
      const int c = __float_as_int(sx[i][j]);
      const float* trow = a.p.exp_table + (__float_as_int(srow[i]) + 128) * 256 + 128;
      sx[i][j] = c == kMasked ? e_mask : __ldg(trow + c);
    }
  } else {
    ExpShiftParams ep;
    ep.x0 = a.x0; ep.nx0 = a.nx0; ep.n = 15;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) {
      const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)), j = t - i * N;
      sx[i][j] = exp_shift(__fsub_rn(sx[i][j], srow[i]), ep);
    }
  }
  __syncthreads();
  if (threadIdx.x < N) {
    float sum = torch_row_sum_small(sx[threadIdx.x], N);
    sum = fminf(sum, 2147483648.f);
    srow[threadIdx.x] = __fmul_rn(floorf(__fdiv_rn(2147483648.f, sum)), a.out_shift);
  }
  __syncthreads();
  // probabilities -> the u8 A operand (rows of 64 bytes; columns past the window are zero)
  for (int t = threadIdx.x; t < N * 16; t += blockDim.x) {       // four codes = one 32-bit store per step
    const int i = t >> 4, j0 = (t & 15) << 2;
    const float fs = srow[i];
    uint32_t word = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (j0 + u < N)
        word |= (__float_as_uint(__fadd_rz(__fmul_rn(sx[i][j0 + u], fs), 8388608.f)) & 0xffu) << (8 * u);
    *reinterpret_cast<uint32_t*>(sP + sw64_off(i, j0)) = word;
  }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    tc_fence_after();
    constexpr uint32_t idesc2 = make_idesc_i8(128, 32, false, true, false, false);     // P u8, v^T s8, both K-major
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const uint64_t da = make_smem_desc(smem_u32(sP) + k * 32, 16, 512, kLayoutSW64);
      const uint64_t db = make_smem_desc(smem_u32(sVt) + k * 32, 16, 512, kLayoutSW64);
      mma_i8_ss(tmem_base, da, db, idesc2, k != 0 ? 1u : 0u);           // over the dead score columns
    }
    tc_commit(&bars[1]);
  }
  __syncwarp();
  if (a.probs != nullptr) {
    uint8_t* pr = a.probs + (static_cast<long long>(b * nW + w) * a.nH + h) * N * N;
    for (int t = threadIdx.x; t < N * N; t += blockDim.x) { const int i = static_cast<int>(__umulhi(static_cast<uint32_t>(t), magicN)); pr[t] = sP[sw64_off(i, t - i * N)]; }
  }
  if (warp < 2) {                              // O rows from TMEM -> qact3 -> the token's image position
    mbar_wait(&bars[1], 0);
    tc_fence_after();
    const int row = 32 * warp + lane;
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      uint32_t o[16];
      tmem_ld16(tmem_base + (static_cast<uint32_t>(32 * warp) << 16) + 16 * half, o);
      tmem_ld_wait();
      if (row < N && a.out != nullptr) {
        uint32_t wds[4] = {0, 0, 0, 0};
#pragma unroll
        for (int d = 0; d < 16; ++d) {
          const int c = clampi(requant_i32_fast(static_cast<int>(o[d]), a.p.mult_out), a.p.qmin_out, a.p.qmax_out);
          wds[d >> 2] |= (static_cast<uint32_t>(c) & 0xffu) << (8 * (d & 3));
        }
        int4* op = reinterpret_cast<int4*>(a.out + (static_cast<long long>(b) * L + tok[row]) * C + h * hd);
        op[half] = make_int4(wds[0], wds[1], wds[2], wds[3]);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 64);
}

// ------------------------------------------------------------------ thread-per-row variant (7x7 windows, hd = 32)
// Round 2.  Two (image, window, head) units per 128-thread CTA, persistent over the unit pairs:
//   S = Q K^T   the A tile holds unit 0's 49 query rows in rows 0..48 and unit 1's in rows 64..112; one 128 x 64 x 64
//               MMA per unit (its own K tile) writes TMEM columns 64 u .. 64 u + 63 -- each thread then owns ONE
//               score row in its own TMEM lane (lanes 0..63 read unit 0's columns, lanes 64..127 unit 1's),
//   the per-score chain runs out of registers: qact_attn1 (FP64 dyadic), qact2 as a 256-entry shared table of
//               rint(c1 m2) plus the layer's precomputed bias term, mask as a per-row 49-bit set, row maximum,
//               exp term gathered from the layer's (row max, code) table, row sum in torch's order, probabilities
//               written as one swizzled 64-byte row of the P tile -- no score tile in shared memory, no barrier
//               between the softmax phases;
//   O = P V     one 128 x 32 x 64 MMA per unit into columns 32 u .. 32 u + 31 (over the dead scores), read back by
//               the row's thread, requantised (qact3) and stored as the token's 32 output bytes.
// Table path only (exp_table present and, under a shift mask, mask_clamps), int8 code ranges.
constexpr int RN = 49, RWS = 7;

template <bool MASKED>
__device__ __forceinline__ void rows_softmax(const WinAttnArgs& a, uint32_t t_row, const int* bt_row, const int* rq2,
                                             unsigned long long mbits, float (&e)[RN]) {
  int c2[RN];
  int mx = -0x7fffffff;
#pragma unroll
  for (int q4 = 0; q4 < 4; ++q4) {
    uint32_t v[16];
    tmem_ld16(t_row + 16 * q4, v);
    tmem_ld_wait();
#pragma unroll
    for (int jj = 0; jj < 16; ++jj) {
      const int j = 16 * q4 + jj;
      if (j < RN) {
        // int32 -> fp64 on the conversion pipe (idle here): two issue slots per requantisation instead of four
        const int c1 = clampi(requant_i32_cvt(static_cast<int>(v[jj]), a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn);
        c2[j] = clampi(rq2[c1 + 128] + bt_row[j], a.p.qmin2, a.p.qmax2);
        if (MASKED) mx = ((mbits >> j) & 1ull) ? mx : max(mx, c2[j]);
        else mx = max(mx, c2[j]);
      }
    }
  }
  mx = max(mx, a.p.qmin2);                       // (a real row always keeps its diagonal: no effect there)
  const float* trow = a.p.exp_table + (mx + 128) * 256 + 128;
  const float e_mask = -a.x0;
  // codes at least sat_k below the row maximum all hold the table's saturated value: no gather for them
  const int csat = a.p.sat_k > 0 ? mx - a.p.sat_k : -1000;
#pragma unroll
  for (int j = 0; j < RN; ++j) {
    const float t = c2[j] > csat ? __ldg(trow + c2[j]) : a.p.sat_e;
    e[j] = (MASKED && ((mbits >> j) & 1ull)) ? e_mask : t;
  }
}

// grid: a multiple of nH CTAs, so that CTA c keeps head c % nH for all of its pairs: the head's bias terms
// (rint(bias code x m_bias), [49][49] int32) live in shared memory.  A pair = the same head of two consecutive
// (image, window) positions bw = 2 m, 2 m + 1 of the B * nW windows.
__global__ void __launch_bounds__(128, 4)
window_attention_rows_kernel(const WinAttnArgs a, int BW) {
  __shared__ __align__(1024) uint8_t tiles[8192 + 2 * 4096 + 8192 + 2 * 2048];
  __shared__ int sbt[RN * RN];
  __shared__ int rq2[256];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  uint8_t* const sQ = tiles;                 // [128][64 B]: unit 0 rows 0..48, unit 1 rows 64..112
  uint8_t* const sK = tiles + 8192;          // two [64][64 B]
  uint8_t* const sP = tiles + 16384;         // [128][64 B]
  uint8_t* const sVt = tiles + 24576;        // two [64 tokens][32 B]: v as it lies in qkv (MN-major B operand)
  constexpr int hd = 32;
  const int res = a.res, nWx = res / RWS, nW = nWx * nWx;
  const int C = a.nH * hd, L = res * res;
  const int warp = threadIdx.x >> 5;
  const int u = threadIdx.x >> 6, i = threadIdx.x & 63;
  const int h = blockIdx.x % a.nH;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 128);
    tmem_relinquish();
  }
  for (int t = threadIdx.x; t < static_cast<int>(sizeof(tiles) / 16); t += blockDim.x)
    reinterpret_cast<int4*>(tiles)[t] = make_int4(0, 0, 0, 0);       // the K padding (bytes 32..63) stays zero
  for (int t = threadIdx.x; t < 256; t += blockDim.x)
    rq2[t] = requant_i32_fast(t - 128, a.p.mult2);                   // qact2's own branch per qact_attn1 code
  for (int t = threadIdx.x; t < RN * RN; t += blockDim.x)            // its identity branch (quant_utils.py:232-245)
    sbt[t] = a.p.bias_terms != nullptr ? __ldg(a.p.bias_terms + h * RN * RN + t)
                                       : requant_i32_fast(__ldg(a.p.bias_codes + h * RN * RN + t), a.p.mult_bias);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(32 * warp) << 16);
  const int iy = i / RWS, ix = i - iy * RWS;
  const int* bt_row = sbt + (i < RN ? i : 0) * RN;
  uint32_t phase = 0;
  const int step = gridDim.x / a.nH;                                  // window pairs between two iterations
  const int npairs = (BW + 1) >> 1;
  for (int m = blockIdx.x / a.nH; m < npairs; m += step, phase ^= 1u) {
    const int bw = 2 * m + u;
    const bool active = bw < BW && i < RN;
    const int bwc = min(bw, BW - 1);
    const int b = bwc / nW, w = bwc - b * nW;
    const int wy = w / nWx, wx = w - wy * nWx;
    const int ry = wy * RWS + iy, rx = wx * RWS + ix;                 // rolled coordinates
    int ty = ry + a.shift, tx = rx + a.shift;                         // < 2 res: one conditional subtract each
    ty -= ty >= res ? res : 0;
    tx -= tx >= res ? res : 0;
    const int tok = a.p.windowed ? w * RN + i : ty * res + tx;
    // shifted-window mask of this row as a 49-bit set (swin_quant.py:226-247): only the last window row / column
    // straddles two regions
    unsigned long long mbits = 0;
    const bool masked = a.shift > 0 && (wy == nWx - 1 || wx == nWx - 1);
    if (masked) {
      const int by = mask_region(ry, res, RWS, a.shift), bx = mask_region(rx, res, RWS, a.shift);
      uint32_t my = 0, mxb = 0;
#pragma unroll
      for (int t = 0; t < RWS; ++t) {
        my |= (mask_region(wy * RWS + t, res, RWS, a.shift) != by ? 1u : 0u) << t;
        mxb |= (mask_region(wx * RWS + t, res, RWS, a.shift) != bx ? 1u : 0u) << t;
      }
#pragma unroll
      for (int t = 0; t < RWS; ++t)
        mbits |= static_cast<unsigned long long>(((my >> t) & 1u) ? 0x7fu : mxb) << (RWS * t);
      if (i >= RN) mbits = 0;                    // padding lanes: no row of their own
    }
    const long long row_off = static_cast<long long>(b) * L + tok;
    if (active) {
      const int8_t* src = a.qkv + row_off * 3 * C + h * hd;
      const int4 q0 = __ldg(reinterpret_cast<const int4*>(src)), q1 = __ldg(reinterpret_cast<const int4*>(src) + 1);
      const int4 k0 = __ldg(reinterpret_cast<const int4*>(src + C)), k1 = __ldg(reinterpret_cast<const int4*>(src + C) + 1);
      const int4 v0 = __ldg(reinterpret_cast<const int4*>(src + 2 * C));
      const int4 v1 = __ldg(reinterpret_cast<const int4*>(src + 2 * C) + 1);
      const int row = 64 * u + i;
      *reinterpret_cast<int4*>(sQ + sw64_off(row, 0)) = q0;
      *reinterpret_cast<int4*>(sQ + sw64_off(row, 16)) = q1;
      *reinterpret_cast<int4*>(sK + 4096 * u + sw64_off(i, 0)) = k0;
      *reinterpret_cast<int4*>(sK + 4096 * u + sw64_off(i, 16)) = k1;
      // v stays in its natural [token][32 B] order: the B operand of O = P V is read MN-major (head-dim contiguous,
      // SWIZZLE_32B: inside each 8-token group of 256 bytes, tokens 4..7 have their two 16-byte halves exchanged)
      uint8_t* vt = sVt + 2048 * u + 32 * i;
      const int sw = ((i >> 2) & 1) << 4;
      *reinterpret_cast<int4*>(vt + sw) = v0;
      *reinterpret_cast<int4*>(vt + (sw ^ 16)) = v1;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      constexpr uint32_t idesc1 = make_idesc_i8(128, 64, true, true, false, false);
#pragma unroll
      for (int uu = 0; uu < 2; ++uu)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const uint64_t da = make_smem_desc(smem_u32(sQ) + k * 32, 16, 512, kLayoutSW64);
          const uint64_t db = make_smem_desc(smem_u32(sK) + 4096 * uu + k * 32, 16, 512, kLayoutSW64);
          mma_i8_ss(tmem_base + 64 * uu, da, db, idesc1, k != 0 ? 1u : 0u);
        }
      tc_commit(&bars[0]);
    }
    __syncwarp();
    {
      // the next pair's q / k / v rows of this thread: pull them into L2 while this pair is worked on
      const int bwn = 2 * (m + step) + u;
      if (bwn < BW && i < RN) {
        const int bn = bwn / nW, wn = bwn - bn * nW;
        const int wyn = wn / nWx, wxn = wn - wyn * nWx;
        int tyn = wyn * RWS + iy + a.shift, txn = wxn * RWS + ix + a.shift;
        tyn -= tyn >= res ? res : 0;
        txn -= txn >= res ? res : 0;
        const int tokn = a.p.windowed ? wn * RN + i : tyn * res + txn;
        const int8_t* nsrc = a.qkv + (static_cast<long long>(bn) * L + tokn) * 3 * C + h * hd;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nsrc));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nsrc + C));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nsrc + 2 * C));
      }
    }
    mbar_wait(&bars[0], phase);
    tc_fence_after();
    {
      float e[RN];
      if (masked) rows_softmax<true>(a, t_lane + 64 * u, bt_row, rq2, mbits, e);
      else rows_softmax<false>(a, t_lane + 64 * u, bt_row, rq2, 0ull, e);
      // torch.sum over the 49 terms (ATen vectorized_inner_sum: 32 accumulators take e[0..31], the two left-over
      // vectors go to accumulator row 0, rows merge in order, lanes add in order after the scalar tail e[48])
      float acc[8];
#pragma unroll
      for (int l = 0; l < 8; ++l) {
        acc[l] = __fadd_rn(__fadd_rn(e[l], e[32 + l]), e[40 + l]);
        acc[l] = __fadd_rn(acc[l], e[8 + l]);
        acc[l] = __fadd_rn(acc[l], e[16 + l]);
        acc[l] = __fadd_rn(acc[l], e[24 + l]);
      }
      float sum = e[48];
#pragma unroll
      for (int l = 0; l < 8; ++l) sum = __fadd_rn(sum, acc[l]);
      sum = fminf(sum, 2147483648.f);                                             // ivit_modules.py:171-173
      const float fs = __fmul_rn(floorf(__fdiv_rn(2147483648.f, sum)), a.out_shift);   // :174-175
      uint32_t pw[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        uint32_t lob[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int j = 4 * k + t;
          // floor of a value in [0, 2^22) through the 2^23 magic constant (round toward zero): the code is the low byte
          lob[t] = j < RN ? __float_as_uint(__fadd_rz(__fmul_rn(e[j], fs), 8388608.f)) : 0u;
        }
        pw[k] = __byte_perm(__byte_perm(lob[0], lob[1], 0x0040), __byte_perm(lob[2], lob[3], 0x0040), 0x5410);
      }
      if (i < RN) {
        const int row = 64 * u + i;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          *reinterpret_cast<int4*>(sP + sw64_off(row, 16 * c)) = make_int4(pw[4 * c], pw[4 * c + 1], pw[4 * c + 2], pw[4 * c + 3]);
        if (a.probs != nullptr && active) {
          uint8_t* pr = a.probs + ((static_cast<long long>(b) * nW + w) * a.nH + h) * RN * RN + i * RN;
#pragma unroll 1
          for (int j = 0; j < RN; ++j) pr[j] = sP[sw64_off(row, j)];
        }
      }
    }
    tc_fence_before();
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      constexpr uint32_t idesc2 = make_idesc_i8(128, 32, false, true, false, true);    // P u8 K-major, v s8 MN-major
#pragma unroll
      for (int uu = 0; uu < 2; ++uu)
#pragma unroll
        for (int k = 0; k < 2; ++k) {              // 32 tokens per MMA: four 8-token groups of 256 bytes
          const uint64_t da = make_smem_desc(smem_u32(sP) + k * 32, 16, 512, kLayoutSW64);
          const uint64_t db = make_smem_desc(smem_u32(sVt) + 2048 * uu + k * 1024, 16, 256, kLayoutSW32);
          mma_i8_ss(tmem_base + 32 * uu, da, db, idesc2, k != 0 ? 1u : 0u);
        }
      tc_commit(&bars[1]);
    }
    __syncwarp();
    mbar_wait(&bars[1], phase);
    tc_fence_after();
    {
      uint32_t o[32];
      tmem_ld16(t_lane + 32 * u, *reinterpret_cast<uint32_t(*)[16]>(&o[0]));
      tmem_ld16(t_lane + 32 * u + 16, *reinterpret_cast<uint32_t(*)[16]>(&o[16]));
      tmem_ld_wait();
      if (active && a.out != nullptr) {
        uint32_t wds[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          int c[4];
#pragma unroll
          for (int t = 0; t < 4; ++t)
            c[t] = clampi(requant_i32_cvt(static_cast<int>(o[4 * k + t]), a.p.mult_out), a.p.qmin_out, a.p.qmax_out);
          wds[k] = __byte_perm(__byte_perm(c[0], c[1], 0x0040), __byte_perm(c[2], c[3], 0x0040), 0x5410);
        }
        int4* op = reinterpret_cast<int4*>(a.out + row_off * C + h * hd);
        op[0] = make_int4(wds[0], wds[1], wds[2], wds[3]);
        op[1] = make_int4(wds[4], wds[5], wds[6], wds[7]);
      }
    }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// PatchMerging gather (swin_quant.py:339-344): [B, res, res, C] -> [B, (res/2)^2, 4C], channel block k
// of output token (i, j) = input token (2 i + k % 2, 2 j + k / 2).
// One 16-byte vector per thread (VEC elements); C % VEC == 0.
template <typename T>
__global__ void patch_merge_gather_kernel(const T* __restrict__ x, T* __restrict__ out, int B, int res, int C) {
  constexpr int VEC = 16 / sizeof(T);
  const int r2 = res / 2, cv = C / VEC;
  const long long total = static_cast<long long>(B) * r2 * r2 * 4 * cv;
  const int4* xin = reinterpret_cast<const int4*>(x);
  int4* o = reinterpret_cast<int4*>(out);
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % cv);
    const int k = static_cast<int>((t / cv) % 4);
    const long long tokq = t / (4LL * cv);
    const int j = static_cast<int>(tokq % r2), i = static_cast<int>((tokq / r2) % r2);
    const int b = static_cast<int>(tokq / (r2 * r2));
    const long long src = ((static_cast<long long>(b) * res + 2 * i + (k & 1)) * res + 2 * j + (k >> 1)) * cv + c;
    o[t] = __ldg(xin + src);
  }
}

// nn.AdaptiveAvgPool1d(1) over the tokens of the float carrier (swin_quant.py:554) followed by qact3:
// torch computes sum(dim) / L in its outer-reduction order (each output sums its L tokens
// sequentially, cascading every 2^lp tokens; oracle/swin_oracle.py::avgpool_tokens).
__global__ void avgpool_requant_kernel(const int8_t* __restrict__ x, int8_t* __restrict__ out, int B, int L, int C,
                                       float s_in, double mult, int qmin, int qmax, int lp) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B * C) return;
  const int b = t / C, c = t % C;
  const int8_t* col = x + static_cast<long long>(b) * L * C + c;
  const int step = 1 << lp, mask = step - 1;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  int i = 0;
  while (i + step <= L) {
    for (int u = 0; u < step; ++u, ++i)
      acc[0] = __fadd_rn(acc[0], __fmul_rn(static_cast<float>(col[static_cast<long long>(i) * C]), s_in));
    for (int j = 1; j < 4; ++j) {
      acc[j] = __fadd_rn(acc[j], acc[j - 1]);
      acc[j - 1] = 0.f;
      if ((i & (mask << (j * lp))) != 0) break;
    }
  }
  for (; i < L; ++i) acc[0] = __fadd_rn(acc[0], __fmul_rn(static_cast<float>(col[static_cast<long long>(i) * C]), s_in));
  for (int j = 1; j < 4; ++j) acc[0] = __fadd_rn(acc[0], acc[j]);
  const float mean = __fdiv_rn(acc[0], static_cast<float>(L));
  const float z = rintf(__fdiv_rn(mean, s_in));                              // quant_utils.py:220
  const double r = requant_d(static_cast<double>(z), mult);
  out[t] = static_cast<int8_t>(static_cast<int>(fmin(fmax(r, static_cast<double>(qmin)), static_cast<double>(qmax))));
}

}  // namespace ivit

using namespace ivit;

// 0: tcgen05 products where hd = 32 (default); 1: the dp4a kernel always (A/B timing, tests)
static int g_win_variant = 0;
extern "C" int ivit_window_attention_set_variant(int variant) {
  const int prev = g_win_variant;
  g_win_variant = variant;
  return prev;
}

extern "C" int ivit_window_attention_i8(const int8_t* qkv, int8_t* out, uint8_t* probs, int B, int res, int ws, int shift,
                                        int nH, int hd, const ivit_window_attention_params_t* p, void* stream) {
  if (!p || !p->bias_codes) return fail(IVIT_ERR_BAD_ARG, "window attention: params / bias codes NULL");
  if (ws < 1 || ws * ws > WMAX || res % ws != 0) return fail(IVIT_ERR_UNSUPPORTED, "window attention: ws^2 <= 64, ws | res");
  if (hd % 4 != 0 || hd > HDMAX) return fail(IVIT_ERR_UNSUPPORTED, "window attention: head_dim % 4 == 0, <= 64");
  if (shift < 0 || shift >= ws) return fail(IVIT_ERR_BAD_ARG, "window attention: 0 <= shift < ws");
  if (p->output_bit < 2 || p->output_bit > 8) return fail(IVIT_ERR_UNSUPPORTED, "window attention: softmax output_bit <= 8");
  const long long ctas = static_cast<long long>(B) * (res / ws) * (res / ws) * nH;
  if (ctas <= 0) return 0;
  if (ctas > 0x7fffffffLL) return fail(IVIT_ERR_UNSUPPORTED, "window attention: too many windows");
  WinAttnArgs a;
  a.qkv = qkv; a.out = out; a.probs = probs;
  a.res = res; a.ws = ws; a.shift = shift; a.nH = nH; a.hd = hd;
  a.p = *p;
  a.x0 = floorf(-1.0f / p->s2);                                              // ivit_modules.py:154
  a.nx0 = 15.f * a.x0;                                                       // :155
  a.out_shift = 1.f / ldexpf(1.f, 31 - p->output_bit + 1);
  const bool table = p->exp_table != nullptr && (shift == 0 || p->mask_clamps != 0) && p->qmin2 >= -128 && p->qmax2 <= 127;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = static_cast<int>(ctas);
  // 7x7 windows, hd = 32, table path, int8 code ranges: the thread-per-row kernel (variant 0)
  if (g_win_variant == 0 && table && ws == RWS && hd == 32 && p->qmin_attn >= -128 && p->qmax_attn <= 127 &&
      (reinterpret_cast<uintptr_t>(out) & 15) == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0) {
    static int attr_set[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    dev = dev < 0 || dev >= 16 ? 0 : dev;
    if (attr_set[dev] == 0) {
      // four CTAs per SM: 128 of the 512 TMEM columns each (the occupancy API does not know about TMEM), 4 x 41 KB of
      // shared memory -- ask for the large carve-out so that they fit
      cudaFuncSetAttribute(window_attention_rows_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                           cudaSharedmemCarveoutMaxShared);
      attr_set[dev] = 1;
    }
    const long long BW = static_cast<long long>(B) * (res / ws) * (res / ws);
    const long long pairs = ((BW + 1) / 2) * nH;
    long long g = pairs < 148LL * 4 ? pairs : 148LL * 4;
    g -= g % nH;                                   // CTA c keeps head c % nH
    if (g < nH) g = nH;
    window_attention_rows_kernel<<<static_cast<int>(g), 128, 0, st>>>(a, static_cast<int>(BW));
    return check_launch("window_attention_rows_kernel");
  }
  // hd = 32 (every Swin size): the two products on tcgen05; otherwise (or variant 1) the dp4a kernel
  const bool tc = g_win_variant != 1 && hd == 32 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(qkv) & 3) == 0;
  if (tc) {
    if (table) window_attention_tc_kernel<true><<<grid, 256, 0, st>>>(a);
    else window_attention_tc_kernel<false><<<grid, 256, 0, st>>>(a);
    return check_launch("window_attention_tc_kernel");
  }
  if (table) window_attention_i8_kernel<true><<<grid, 256, 0, st>>>(a);
  else window_attention_i8_kernel<false><<<grid, 256, 0, st>>>(a);
  return check_launch("window_attention_i8_kernel");
}

extern "C" int ivit_patch_merge_gather(const void* x, void* out, int elem_bits, int B, int res, int C, void* stream) {
  if (res % 2 != 0) return fail(IVIT_ERR_BAD_ARG, "patch merge: resolution must be even");
  if (elem_bits != 8 && elem_bits != 16) return fail(IVIT_ERR_BAD_ARG, "patch merge: elem_bits must be 8 or 16");
  if ((static_cast<long long>(C) * elem_bits / 8) % 16 != 0 || (reinterpret_cast<uintptr_t>(x) & 15) ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return fail(IVIT_ERR_UNSUPPORTED, "patch merge: rows of a multiple of 16 bytes, 16-byte aligned buffers");
  const long long total = static_cast<long long>(B) * res * res * C * (elem_bits / 8) / 16;
  if (total <= 0) return 0;
  long long g = (total + 255) / 256;
  if (g > 148LL * 16) g = 148LL * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (elem_bits == 16)
    patch_merge_gather_kernel<int16_t><<<static_cast<int>(g), 256, 0, st>>>(static_cast<const int16_t*>(x),
                                                                            static_cast<int16_t*>(out), B, res, C);
  else if (elem_bits == 8)
    patch_merge_gather_kernel<int8_t><<<static_cast<int>(g), 256, 0, st>>>(static_cast<const int8_t*>(x),
                                                                           static_cast<int8_t*>(out), B, res, C);
  else
    return fail(IVIT_ERR_BAD_ARG, "patch merge: elem_bits must be 8 or 16");
  return check_launch("patch_merge_gather_kernel");
}

extern "C" int ivit_avgpool_requant_i8(const int8_t* x, int8_t* out, int B, int L, int C, float s_in, double mult,
                                       int qmin, int qmax, void* stream) {
  if (B * C <= 0 || L <= 0) return 0;
  int ceil_log2 = 0;
  while ((1LL << ceil_log2) < L) ++ceil_log2;
  const int lp = ceil_log2 / 4 > 4 ? ceil_log2 / 4 : 4;
  avgpool_requant_kernel<<<(B * C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(x, out, B, L, C, s_in, mult,
                                                                                          qmin, qmax, lp);
  return check_launch("avgpool_requant_kernel");
}
