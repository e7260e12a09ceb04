// Fused integer attention for one (image, head, 128-query tile):
//
//   S = Q K^T            tcgen05 kind::i8, INT32 in TMEM          (QuantMatMul, quant_modules.py:404-409)
//   c = qact_attn1(S)    dyadic requant, scalar ratio              (vit_quant.py:74-76)
//   P = IntSoftmax(c)    Shiftmax (ivit_modules.py:150-179) or I-BERT int-exp (ibert_modules.py:275-319)
//   O = P V              tcgen05 kind::i8 (u8 x s8), V consumed MN-major straight from its [token][hd] layout
//   out = qact2(O)       dyadic requant, stored head-interleaved [B,N,H*hd] (vit_quant.py:83-85)
//
// The N x N score tile never leaves the SM: it lives in TMEM (256 columns), each of the 128
// threads owns one query row (thread = TMEM lane), so row max / row sum need no shuffles, and
// the float32 row sum follows torch's CPU order exactly (intmath.cuh).  TMEM doubles as the
// scratch for the per-row intermediates between the three softmax passes.
#include <cuda.h>
#include <cstdio>
#include <type_traits>

#include "intmath.cuh"
#include "ivit_b200.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ivit {
int fail(int code, const char* msg);
int check_launch(const char* what);
int make_tmap_2d_u8(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_cols,
                    uint32_t box_rows, CUtensorMapSwizzle swz);

struct AttnArgs {
  int B, N, H;
  int NKP;        // N rounded up to 16: columns of S (MMA1 N)
  int KP2;        // N rounded up to 32: K extent of P.V
  int crow;       // rows kernel: byte stride of a code row in shared memory (odd multiple of 16)
  int8_t* out;
  uint8_t* probs;  // optional debug tap [B,H,N,N] (uint8, or uint16 for a 16-bit softmax)
  float s_pv;      // 16-bit softmax: float32 scale of the P.V accumulator (s_softmax * s_v)
  const float* exp_table;  // optional [256][256]: softmax_term(c, cmax) tabulated (bit-identical)
  const float* exp_table_1d;   // optional [256]: the same over cmax - c, when the 2-D table collapses
  ivit_attention_params_t p;
  // derived on the host with float32 arithmetic, mirroring the reference's scalar ops
  float x0, nx0;              // ivit: floor(-1/s), 15*x0 ; ibert: floor(-0.6931/s), 30*x0
  float b_int, c_int;         // ibert polynomial (ibert_modules.py:277-278)
  float exp_scale;            // ibert: 0.35815147*s^2/2^30 (:282,:294)
  float num;                  // 2^31 (ivit, float32(2^31-1)) or 2^32 (ibert)
  float out_shift;            // 2^-(31-ob+1) or 2^-(32-ob+1)
};

constexpr int HD = 64;

#ifdef IVIT_ATTN_TIMING
// debug probe: clock64 deltas between the phases of attention_i8_kernel, summed over CTAs by thread 0
__device__ unsigned long long g_attn_t[16];
#define ATTN_STAMP(i) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_attn_t[i], static_cast<unsigned long long>(t_ - t_prev)); t_prev = t_; } } while (0)
#else
#define ATTN_STAMP(i) do { } while (0)
#endif

// exp term of one score, to be summed over the row.  c, cmax: int8 codes of qact_attn1.
__device__ __forceinline__ float softmax_term(int c, float xmax, const AttnArgs& a) {
  const float xi = x_eff(static_cast<float>(c), a.p.s_attn);       // x / s, un-rounded
  const float d = __fsub_rn(xi, xmax);                             // x_int - max
  if (a.p.kind == 0) {
    ExpShiftParams ep;
    ep.x0 = a.x0; ep.nx0 = a.nx0; ep.n = 15;
    return exp_shift(d, ep);                                       // ivit_modules.py:150-162
  }
  // ibert_modules.py:285-295 int_exp, then :310-311 inner QuantAct(16) and its un-rounded recovery
  const float t = fmaxf(d, a.nx0);
  const float q = floorf(__fdiv_rn(t, a.x0));
  const float r = __fsub_rn(t, __fmul_rn(a.x0, q));
  float z = __fadd_rn(r, a.b_int);
  z = __fmul_rn(r, z);
  z = __fadd_rn(z, a.c_int);
  const float e = fmaxf(floorf(__fmul_rn(z, pow2i(30 - static_cast<int>(q)))), 0.f);
  // QuantAct(16)(exp_int, exp_scale): the reference passes the *integers* as the value
  const float zi = rintf(__fdiv_rn(e, a.exp_scale));                              // quant_utils.py:220
  double rr = requant_d(static_cast<double>(zi), a.p.mult_exp16);                 // :229-230
  rr = fmin(fmax(rr, -32768.0), 32767.0);                                         // :247-249
  return x_eff(static_cast<float>(rr), a.p.s_exp16);                              // ibert_modules.py:311
}

// The exp term of one score depends only on (its int8 code, the row-max int8 code): tabulated once
// per layer at freeze time by the very same device function, as table[cmax+128][c+128].
__global__ void softmax_table_kernel(float* __restrict__ table, const AttnArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 65536) return;
  const int c = (idx & 255) - 128, mx = (idx >> 8) - 128;
  table[idx] = (c <= mx) ? softmax_term(c, x_eff(static_cast<float>(mx), a.p.s_attn), a) : 0.f;
}

// Does table[cmax][c] depend on cmax - c only?  (True only when fl(fl(c*s)/s) == c for EVERY int8 code: measured
// 5.5 % of random scales in [0.01, 0.5]; 2 of the 12 DeiT-S layers, 1 of 12 for DeiT-T, 3 of 12 for ViT-B.)  t1[d] = table[127][127 - d]; *ok is cleared on the first mismatch.
__global__ void softmax_table_collapse_init(const float* __restrict__ table, float* __restrict__ t1, int* ok) {
  t1[threadIdx.x] = table[255 * 256 + 255 - threadIdx.x];
  if (threadIdx.x == 0) *ok = 1;
}
__global__ void softmax_table_collapse_check(const float* __restrict__ table, const float* __restrict__ t1, int* ok) {
  const int m = blockIdx.x, j = threadIdx.x;
  if (j <= m && __float_as_uint(table[m * 256 + j]) != __float_as_uint(t1[m - j])) *ok = 0;
}

__device__ __forceinline__ void tmem_ld_w(uint32_t t, uint32_t (&r)[8]) { tmem_ld8(t, r); }
__device__ __forceinline__ void tmem_ld_w(uint32_t t, uint32_t (&r)[16]) { tmem_ld16(t, r); }
__device__ __forceinline__ void tmem_ld_w(uint32_t t, uint32_t (&r)[32]) { tmem_ld32(t, r); }
__device__ __forceinline__ void tmem_st_w(uint32_t t, const uint32_t (&r)[8]) { tmem_st8(t, r); }
__device__ __forceinline__ void tmem_st_w(uint32_t t, const uint32_t (&r)[16]) { tmem_st16(t, r); }

// for ci in [0, n): body(ci, v) over this thread's W TMEM columns of 32-column chunk ci, with the
// tcgen05.ld of chunk ci + 1 in flight while chunk ci is processed (two register buffers).
template <int W, class F>
__device__ __forceinline__ void chunk_pipeline(uint32_t t_row, int n, F&& body) {
  uint32_t v0[W], v1[W];
  if (n > 0) tmem_ld_w(t_row, v0);
  for (int ci = 0; ci < n; ci += 2) {
    tmem_ld_wait();
    if (ci + 1 < n) tmem_ld_w(t_row + (ci + 1) * 32, v1);
    body(ci, v0);
    if (ci + 1 < n) {
      tmem_ld_wait();
      if (ci + 2 < n) tmem_ld_w(t_row + (ci + 2) * 32, v0);
      body(ci + 1, v1);
    }
  }
}

// for ci in [0, n): body(ci, v), D chunks' tcgen05.ld in flight at once (one wait per D chunks): for the passes whose
// per-chunk work is far shorter than a TMEM round trip (raw maximum, probability scaling), where the two-deep
// pipeline above pays one load latency per chunk (the clock probe: 1760 cycles for pass 1 of a 7-chunk row).
template <int W, int D, class F>
__device__ __forceinline__ void chunk_batches(uint32_t t_row, int n, F&& body) {
  uint32_t v[D][W];
  for (int c0 = 0; c0 < n; c0 += D) {
#pragma unroll
    for (int d = 0; d < D; ++d)
      if (c0 + d < n) tmem_ld_w(t_row + (c0 + d) * 32, v[d]);
    tmem_ld_wait();
#pragma unroll
    for (int d = 0; d < D; ++d)
      if (c0 + d < n) body(c0 + d, v[d]);
  }
}

// CTA = one (image, head, 128-query tile); TPR threads per query row (2 or 4): thread q of a row
// (warps 4q .. 4q+3) owns columns [32i + q W, 32i + (q+1) W), W = 32 / TPR, of every 32-column
// chunk.  That split keeps torch's float32 summation order (accumulator e % 32 lives in exactly one
// thread) while multiplying the warps that hide the TMEM / table-gather latency (the kernel is
// occupancy-bound: TMEM allows two CTAs per SM).  TABLE 1: exp term gathered from the global
// (row max, code) table (32 sectors per warp load); TABLE 2: from the 256-entry (row max - code)
// table replicated 16x in shared memory (copy lane % 16: at most 2 lanes per bank); TABLE 0:
// computed.  TAP: also dump the softmax codes (tests).
// P16: 16-bit softmax output (softmax_bw = 16; codes up to 32767, I-BERT 32768): the u16 operand is
// split into two u8 planes P = 256 Phi + Plo, each its own SWIZZLE_128B tile and its own MMA into
// its own TMEM accumulator (O_lo columns [0,64), O_hi [64,128), both over the dead score tile);
// O = 256 O_hi + O_lo reaches 8e8, so attn.qact2 takes the general requant path that replays the
// float32 carrier of the reference (requant_any).  The two planes double sP; the row-sum exchange buffer
// is aliased onto the dead Q / K tiles so that two CTAs per SM still fit (110 KB each).
template <int TPR, int TABLE, bool TAP, bool P16 = false>
__global__ void __launch_bounds__(128 * TPR, 2)
attention_i8_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  constexpr int W = 32 / TPR;               // columns per thread per 32-column chunk
  constexpr int OW = HD / TPR;              // output head-dims per thread
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sP = smem;                       // 2 x [128 rows x 128 B], SWIZZLE_128B (written by threads)
  constexpr int P_BYTES = P16 ? 65536 : 32768;   // P16: low-byte plane, then high-byte plane
  uint8_t* sQ = smem + P_BYTES;             // [128 x 64 B]  SWIZZLE_64B (TMA)
  uint8_t* sK = sQ + 8192;                  // [<=256 x 64 B]
  uint8_t* sV = sK + 16384;                 // [<=256 x 64 B]
  // [32][128] accumulators of threads q > 0 -> thread 0.  P16: aliased onto sQ + sK (24 KB, dead once the
  // score MMA has completed, which every thread waits for before pass 1), so that two CTAs still fit an SM
  float* sAcc = P16 ? reinterpret_cast<float*>(sQ) : reinterpret_cast<float*>(sV + 16384);
  int* sMax = P16 ? reinterpret_cast<int*>(sV + 16384) : reinterpret_cast<int*>(sAcc + 32 * 128);   // [4][128]
  float* sFac = reinterpret_cast<float*>(sMax + 512);          // [128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sFac + 128);    // qk, v, s, o
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  float* sT = reinterpret_cast<float*>(bars + 8);              // TABLE 2: [256][16]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp >> 2;                  // which of the row's TPR threads
  const int quarter = warp & 3;             // TMEM lane quarter
  // the TPR threads of a query row sit in warps quarter, quarter + 4, ...: the row-max / row-sum exchanges only
  // need THOSE warps in step (named barrier 1 + quarter), not the whole CTA
  auto row_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(1 + quarter), "n"(32 * TPR) : "memory"); };
  // grid (image*head, query tile): every full 128-query tile is dispatched before the ragged last
  // tiles (69 of 197 rows), so the short CTAs fill the tail of the launch
  const int mt = blockIdx.y;
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int C = a.H * HD;
  const int N = a.N;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
#ifdef IVIT_ATTN_TIMING
  long long t_prev = clock64();
#endif
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                   // programmatic dependent launch: qkv is the predecessor's output
  ATTN_STAMP(0);                 // barrier init, TMEM allocation, first CTA barrier

  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], 128 * HD + a.NKP * HD);
    tma_load_2d(sQ, &tmQ, &bars[0], h * HD, b * N + mt * 128);
    tma_load_2d(sK, &tmK, &bars[0], C + h * HD, b * N);
    mbar_arrive_expect_tx(&bars[1], a.KP2 * HD);
    tma_load_2d(sV, &tmV, &bars[1], 2 * C + h * HD, b * N);
    // S = Q K^T : A, B K-major, 64-byte rows (SWIZZLE_64B: 8-row groups 512 B apart)
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t idesc1 = make_idesc_i8(128, a.NKP, true, true, false, false);
#pragma unroll
    for (int k = 0; k < HD / 32; ++k) {
      const uint64_t da = make_smem_desc(smem_u32(sQ) + k * 32, 16, 512, kLayoutSW64);
      const uint64_t db = make_smem_desc(smem_u32(sK) + k * 32, 16, 512, kLayoutSW64);
      mma_i8_ss(tmem_base, da, db, idesc1, k != 0 ? 1u : 0u);
    }
    tc_commit(&bars[2]);
  }
  __syncwarp();
  if (TABLE == 2 && threadIdx.x < 256) {
    // replicate while the TMA loads and the S MMA run; visible after the barrier that ends pass 1
#pragma unroll
    for (int i = 0; i < 16; ++i) sT[i * 256 + threadIdx.x] = __ldg(a.exp_table_1d + i * 16 + (threadIdx.x >> 4));
  }

  // ---------------------------------------------------------------- softmax
  mbar_wait(&bars[2], 0);
  ATTN_STAMP(1);                 // TMA of Q / K + S MMA latency
  tc_fence_after();
  const int r_loc = quarter * 32 + lane;
  const int row = mt * 128 + r_loc;                     // token index within the image
  const int cbase = q * W;                              // this thread's first column inside a chunk
  const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + cbase;
  const int nfull = N / 32;                             // chunks with all 32 columns valid
  const int nchunks = (N + 31) / 32;
  const int pchunks = a.KP2 / 32;
  // chunks in which THIS thread's W-column segment holds a valid column (warp-uniform: all threads
  // of a warp have the same q): for N = 197 the last segments skip the 7th chunk entirely
  const int my_chunks = (N - cbase + 31) / 32;
  const bool warp_active = mt * 128 + quarter * 32 < N; // warp-uniform: rows past the last token do no work

  // pass 1: row maximum of the RAW int32 scores (qact_attn1 is monotone: the row-max code is the requantised
  // raw maximum), nothing is written back
  int smax = -(1 << 30);
  if (warp_active) {
    chunk_batches<W, 64 / W>(t_row, my_chunks, [&](int ci, uint32_t (&v)[W]) {
      if (ci < nfull) {
#pragma unroll
        for (int j = 0; j < W; ++j) smax = max(smax, static_cast<int>(v[j]));
      } else {
#pragma unroll
        for (int j = 0; j < W; ++j)
          if (ci * 32 + cbase + j < N) smax = max(smax, static_cast<int>(v[j]));
      }
    });
    sMax[q * 128 + r_loc] = smax;
  }
  if (TABLE == 2) __syncthreads();      // also publishes the replicated 1-D table written by all warps above
  else row_sync();
  int cmax = 0;
  float xmax = 0.f;
  const float* trow = nullptr;
  if (warp_active) {
#pragma unroll
    for (int t = 0; t < TPR; ++t) smax = max(smax, sMax[t * 128 + r_loc]);
    cmax = clampi(requant_i32_fast(smax, a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn);
    xmax = x_eff(static_cast<float>(cmax), a.p.s_attn);   // max of un-rounded x/s (monotone in code)
    if (TABLE == 1) trow = a.exp_table + (cmax + 128) * 256 + 128;
    if (TABLE == 2) trow = sT + cmax * 16 + (lane & 15);           // entry (cmax - c) of copy lane % 16
  }
  // scores at least sat_k codes below the row maximum all hold the table's saturated value: no lookup (fewer
  // distinct lines per gather: the L1 tag stage is what bounds this pass)
  const int csat = a.p.sat_k > 0 ? cmax - a.p.sat_k : -1000;
  auto term = [&](int c) -> float {
    if (TABLE == 1) return c > csat ? __ldg(trow + c) : a.p.sat_e;
    if (TABLE == 2) return trow[-16 * c];
    return softmax_term(c, xmax, a);
  };
  // qact_attn1 of one raw score (vit_quant.py:74-76)
  auto code = [&](uint32_t sraw) -> int {
    return clampi(requant_i32_cvt(static_cast<int>(sraw), a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn);
  };

  ATTN_STAMP(2);                 // pass 1 (raw maximum) + exchange
  // pass 2: requant -> code -> exp term -> TMEM; float32 row sum in torch's CPU order: accumulator
  // (col % 32) is acc[col % W] of thread (col % 32) / W, fed in ascending column order.
  const int nvec = N / 8, size_ilp = nvec / 4;           // size_ilp == nfull
  float2 acc2[W / 2];
#pragma unroll
  for (int j = 0; j < W / 2; ++j) acc2[j] = make_float2(0.f, 0.f);
  if (warp_active) {
    // software pipeline: the table lookups of chunk ci + 1 are issued before the terms of chunk ci are summed
    // and stored, so their latency hides behind the next chunk's requantisation (two register sets)
    {
      uint32_t s0[W], s1[W];
      float e0[W], e1[W];
      auto lookups = [&](const uint32_t (&sv)[W], float (&ev)[W]) {
#pragma unroll
        for (int j = 0; j < W; ++j) ev[j] = term(code(sv[j]));
      };
      auto consume = [&](int ci, const float (&ev)[W]) {
        uint32_t v[W];
#pragma unroll
        for (int j = 0; j < W; j += 2) {
          acc2[j >> 1] = __fadd2_rn(acc2[j >> 1], make_float2(ev[j], ev[j + 1]));
          v[j] = __float_as_uint(ev[j]);
          v[j + 1] = __float_as_uint(ev[j + 1]);
        }
        tmem_st_w(t_row + ci * 32, v);
      };
      const int n = size_ilp;
      if (n > 0) {
        tmem_ld_w(t_row, s0);
        tmem_ld_wait();
        if (n > 1) tmem_ld_w(t_row + 32, s1);
        lookups(s0, e0);
      }
      for (int ci = 0; ci < n; ci += 2) {
        if (ci + 1 < n) {
          tmem_ld_wait();
          if (ci + 2 < n) tmem_ld_w(t_row + (ci + 2) * 32, s0);
          lookups(s1, e1);
        }
        consume(ci, e0);
        if (ci + 1 < n) {
          if (ci + 2 < n) {
            tmem_ld_wait();
            if (ci + 3 < n) tmem_ld_w(t_row + (ci + 3) * 32, s1);
            lookups(s0, e0);
          }
          consume(ci + 1, e1);
        }
      }
    }
    if (my_chunks > size_ilp) {        // partial chunk: exp terms only; its sum is thread 0's job below
      uint32_t v[W];
      tmem_ld_w(t_row + size_ilp * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < W; ++j)
        v[j] = (size_ilp * 32 + cbase + j < N) ? __float_as_uint(term(code(v[j]))) : 0u;
      tmem_st_w(t_row + size_ilp * 32, v);
    }
    tmem_st_wait();
  }
  float acc[W];
#pragma unroll
  for (int j = 0; j < W / 2; ++j) {
    acc[2 * j] = acc2[j].x;
    acc[2 * j + 1] = acc2[j].y;
  }
  if (warp_active && q > 0) {
#pragma unroll
    for (int j = 0; j < W; ++j) sAcc[(cbase + j) * 128 + r_loc] = acc[j];
  }
  ATTN_STAMP(3);                 // pass 2 (requant, lookups, sum, TMEM store)
  tc_fence_before();
  row_sync();
  tc_fence_after();
  if (warp_active && q == 0) {
    // torch keeps 4 rows of 8 accumulators (e % 32 = 8 k + l): ps[0] (row 0) += left-over whole
    // vectors, THEN rows 1..3 are added, then the tail scalars.  Accumulator i lives in this thread
    // for i < W, else in shared memory.
    auto acc32 = [&](int i) -> float { return i < W ? acc[i < W ? i : 0] : sAcc[i * 128 + r_loc]; };
    float p8[8];
#pragma unroll
    for (int l = 0; l < 8; ++l) p8[l] = acc[l];
    float tail = 0.f;
    if (nchunks > size_ilp) {
      uint32_t v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + size_ilp * 32, v);   // every thread's terms
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = size_ilp * 32 + j;
        const float e = __uint_as_float(v[j]);
        if (col < nvec * 8) p8[j & 7] = __fadd_rn(p8[j & 7], e);
        else if (col < N) tail = __fadd_rn(tail, e);
      }
    }
#pragma unroll
    for (int k = 1; k < 4; ++k)
#pragma unroll
      for (int l = 0; l < 8; ++l) p8[l] = __fadd_rn(p8[l], acc32(8 * k + l));
    float sum = tail;
#pragma unroll
    for (int l = 0; l < 8; ++l) sum = __fadd_rn(sum, p8[l]);
    if (a.p.kind == 0) sum = fminf(sum, 2147483648.f);               // ivit_modules.py:173
    if (a.p.kind == 2) sum = fmaxf(sum, 1.f);                        // ppoly_modules.py:318
    const float factor = floorf(__fdiv_rn(a.num, sum));              // :174 / ibert :314
    // floor(e*factor / 2^k) == floor(e * (factor / 2^k)): scaling by a power of two is exact
    sFac[r_loc] = __fmul_rn(factor, a.out_shift);
  }
  row_sync();

  ATTN_STAMP(4);                 // row sums, factor
  // pass 3: probabilities -> u8 operand tile in shared memory (K-major, SWIZZLE_128B)
  if (warp_active) {
    const float fs = sFac[r_loc];
    auto store_p = [&](uint8_t* plane, int ci, const uint32_t (&w)[W / 4]) {
      // W bytes inside one 16-byte unit of this row's 128-byte swizzle line
      const int col = ci * 32 + cbase;
      uint8_t* base = plane + (col >> 7) * 16384 + (r_loc >> 3) * 1024 + (r_loc & 7) * 128;
      const int u0 = (col & 127) >> 4;
      uint8_t* dst = base + ((u0 ^ (r_loc & 7)) << 4) + (col & 15);
      if (W == 16) *reinterpret_cast<int4*>(dst) = make_int4(w[0], w[1], w[W / 4 - 2], w[W / 4 - 1]);
      else *reinterpret_cast<uint2*>(dst) = make_uint2(w[0], w[1]);
    };
    chunk_batches<W, 48 / W>(t_row, my_chunks, [&](int ci, uint32_t (&v)[W]) {
      uint32_t w[W / 4], wh[W / 4];
      // :175 / :315 (terms past column N were stored as 0).  floor of a value in [0, 2^22) through the 2^23
      // magic constant (round toward zero) instead of F2I (quarter-rate XU pipe); two scores per instruction
      // (f32x2), the code bytes gathered by PRMT
      const float2 fs2 = make_float2(fs, fs), magic2 = make_float2(8388608.f, 8388608.f);
#pragma unroll
      for (int j = 0; j < W / 4; ++j) {
        const float2 p01 = __fadd2_rz(__fmul2_rn(make_float2(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1])), fs2), magic2);
        const float2 p23 = __fadd2_rz(__fmul2_rn(make_float2(__uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])), fs2), magic2);
        const uint32_t b0 = __float_as_uint(p01.x), b1 = __float_as_uint(p01.y), b2 = __float_as_uint(p23.x),
                       b3 = __float_as_uint(p23.y);
        w[j] = __byte_perm(__byte_perm(b0, b1, 0x0040), __byte_perm(b2, b3, 0x0040), 0x5410);
        if (P16) wh[j] = __byte_perm(__byte_perm(b0, b1, 0x0051), __byte_perm(b2, b3, 0x0051), 0x5410);   // codes <= 32768
        else wh[j] = 0;
      }
      if (TAP && a.probs != nullptr && row < N) {
        const long long po = ((static_cast<long long>(blockIdx.x) * N + row) * N) + ci * 32 + cbase;
        for (int j = 0; j < W && ci * 32 + cbase + j < N; ++j) {
          const uint32_t lo = (w[j >> 2] >> (8 * (j & 3))) & 0xffu, hi = (wh[j >> 2] >> (8 * (j & 3))) & 0xffu;
          if (P16) reinterpret_cast<uint16_t*>(a.probs)[po + j] = static_cast<uint16_t>(lo | (hi << 8));
          else a.probs[po + j] = static_cast<uint8_t>(lo);
        }
      }
      store_p(sP, ci, w);
      if (P16) store_p(sP + 32768, ci, wh);
    });
    uint32_t zero[W / 4];
#pragma unroll
    for (int j = 0; j < W / 4; ++j) zero[j] = 0;
    for (int ci = my_chunks; ci < pchunks; ++ci) {   // K padding of the P V product
      store_p(sP, ci, zero);
      if (P16) store_p(sP + 32768, ci, zero);
    }
  }
  fence_proxy_async_smem();     // generic-proxy smem writes -> visible to tcgen05.mma
  tc_fence_before();
  __syncthreads();

  ATTN_STAMP(5);                 // pass 3 + CTA barrier
  // ---------------------------------------------------------------- O = P V
  if (threadIdx.x == 0) {
    mbar_wait(&bars[1], 0);
    tc_fence_after();
    const uint32_t idesc2 = make_idesc_i8(128, HD, false, true, false, true);   // A u8 K-major, B s8 MN-major
    const int ksteps = a.KP2 / 32;
    for (int pl = 0; pl < (P16 ? 2 : 1); ++pl) {
      for (int k = 0; k < ksteps; ++k) {
        const uint32_t pa = smem_u32(sP) + pl * 32768 + (k >> 2) * 16384 + (k & 3) * 32;
        const uint64_t da = make_smem_desc(pa, 16, 1024, kLayoutSW128);
        // V tile [token][64 B]: MN-major, one 64-byte swizzle atom wide, 8-token groups 512 B apart
        const uint64_t db = make_smem_desc(smem_u32(sV) + k * 32 * HD, 512, 512, kLayoutSW64);
        mma_i8_ss(tmem_base + pl * HD, da, db, idesc2, k != 0 ? 1u : 0u);
      }
    }
    tc_commit(&bars[3]);
  }
  __syncwarp();
  mbar_wait(&bars[3], 0);
  ATTN_STAMP(6);                 // P.V MMA + commit latency
  tc_fence_after();
  if (warp_active) {
    // O tile: columns [0,64); thread q stores head-dims [q OW, (q+1) OW).  The TMEM load is
    // warp-collective (.sync.aligned): it must stay outside the per-lane row guard.
    uint32_t o[OW];
    tmem_ld_w(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + q * OW, o);
    if (P16) {
      uint32_t oh[OW];
      tmem_ld_w(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + HD + q * OW, oh);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < OW; ++j) o[j] += oh[j] << 8;          // O = 256 O_hi + O_lo  (< 2^31)
    }
    tmem_ld_wait();
    if (row < N && a.out != nullptr) {
      uint32_t w[OW / 4];
      auto oc = [&](uint32_t acc) -> uint32_t {
        if (P16) {
          const double r = requant_any(static_cast<int>(acc), a.s_pv, a.p.mult_out);
          return static_cast<uint32_t>(static_cast<int>(fmin(fmax(r, static_cast<double>(a.p.qmin_out)), static_cast<double>(a.p.qmax_out))));
        }
        return static_cast<uint32_t>(clampi(requant_i32_cvt(static_cast<int>(acc), a.p.mult_out), a.p.qmin_out, a.p.qmax_out));
      };
#pragma unroll
      for (int j = 0; j < OW / 4; ++j)
        w[j] = __byte_perm(__byte_perm(oc(o[4 * j]), oc(o[4 * j + 1]), 0x0040),
                           __byte_perm(oc(o[4 * j + 2]), oc(o[4 * j + 3]), 0x0040), 0x5410);
      int8_t* op = a.out + (static_cast<long long>(b) * N + row) * C + h * HD + q * OW;
#pragma unroll
      for (int j = 0; j < OW / 16; ++j)
        *reinterpret_cast<int4*>(op + 16 * j) = make_int4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
    }
  }
  ATTN_STAMP(7);                 // epilogue (TMEM load, requant, store)
  pdl_launch_dependents();      // late: the grid has more CTAs than fit, the successor must not crowd them out
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ---------------------------------------------------------------------------------------------------
// 8-bit softmax output, second kernel: the softmax runs ONE QUERY ROW PER WARP-STEP.
//
// The kernel above keeps thread = query row through the whole softmax, so the 32 lanes of a table lookup
// carry 32 different row maxima and hit 32 different 1 KB table rows: 32 tag look-ups per request, and ncu
// shows the L1 tag stage saturated during that pass (l1tex__t_output_wavefronts 52 % of peak over the whole
// kernel).  Here the scores are requantised once with thread = row (phase A: TMEM -> int8 codes -> shared
// memory, row maxima on packed bytes), and then each warp walks whole rows with lane = column % 32 (phase B):
// every lane of a lookup reads the SAME table row (<= 8 lines), the 32 lanes ARE torch's 32 float32
// accumulators (intmath.cuh / oracle torch_sum_lastdim), the exp terms stay in registers (no TMEM round
// trips), and a row's probability bytes leave as 32 contiguous bytes of the swizzled P tile.
template <int TABLE, bool TAP, int NCH, int NFIX>
__global__ void __launch_bounds__(256, 2)
attention_rows_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const AttnArgs a) {
  constexpr int TPR = 2, W = 16, OW = HD / TPR;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sP = smem;                       // 2 x [128 rows x 128 B], SWIZZLE_128B (written by the lanes)
  uint8_t* sQ = smem + 32768;               // [128 x 64 B]  SWIZZLE_64B (TMA)
  uint8_t* sK = sQ + 8192;                  // [<=256 x 64 B]
  uint8_t* sV = sK + 16384;                 // [<=256 x 64 B]
  int8_t* sC = reinterpret_cast<int8_t*>(sV + 16384);            // [128][crow] qact_attn1 codes; crow = an odd number of
  int* sMax = reinterpret_cast<int*>(sC + 128 * a.crow);         // 16-byte units: conflict-free 16-byte stores.  [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sMax + 256);      // qk, v, s, o
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp >> 2, quarter = warp & 3;
  const int mt = blockIdx.y;
  const int b = blockIdx.x / a.H, h = blockIdx.x % a.H;
  const int C = a.H * HD;
  // NFIX: the token count as a compile-time constant (197 = 224 x 224 / 16 x 16 + class token: every DeiT / ViT
  // of the reference's factories), so that the chunk / tail predicates of phase B fold; 0 = run-time N
  const int N = NFIX ? NFIX : a.N;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], 128 * HD + a.NKP * HD);
    tma_load_2d(sQ, &tmQ, &bars[0], h * HD, b * N + mt * 128);
    tma_load_2d(sK, &tmK, &bars[0], C + h * HD, b * N);
    mbar_arrive_expect_tx(&bars[1], a.KP2 * HD);
    tma_load_2d(sV, &tmV, &bars[1], 2 * C + h * HD, b * N);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t idesc1 = make_idesc_i8(128, a.NKP, true, true, false, false);
#pragma unroll
    for (int k = 0; k < HD / 32; ++k) {
      const uint64_t da = make_smem_desc(smem_u32(sQ) + k * 32, 16, 512, kLayoutSW64);
      const uint64_t db = make_smem_desc(smem_u32(sK) + k * 32, 16, 512, kLayoutSW64);
      mma_i8_ss(tmem_base, da, db, idesc1, k != 0 ? 1u : 0u);
    }
    tc_commit(&bars[2]);
  }
  __syncwarp();

  // ---------------------------------------------------------------- phase A: thread = query row
  mbar_wait(&bars[2], 0);
  tc_fence_after();
  const int CROW = a.crow;
  const int r_loc = quarter * 32 + lane;
  const int row = mt * 128 + r_loc;                     // token index within the image
  const int cbase = q * W;
  const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + cbase;
  const int nfull = N / 32;                             // chunks with all 32 columns valid
  const int nchunks = (N + 31) / 32;
  const int my_chunks = (N - cbase + 31) / 32;          // chunks in which this thread's segment has a valid column
  const int rows_here = min(128, N - mt * 128);
  const bool warp_active = mt * 128 + quarter * 32 < N; // warp-uniform
  auto code = [&](uint32_t sraw) -> uint32_t {          // qact_attn1 of one raw score (vit_quant.py:74-76)
    return static_cast<uint32_t>(
        clampi(requant_i32_cvt(static_cast<int>(sraw), a.p.mult_attn), a.p.qmin_attn, a.p.qmax_attn));
  };
  if (warp_active) {
    uint32_t m4 = 0x80808080u;                          // four int8 lanes of -128
    const uint32_t crow = smem_u32(sC) + r_loc * CROW + cbase;
    chunk_pipeline<W>(t_row, my_chunks, [&](int ci, uint32_t (&v)[W]) {
      uint32_t w[4];
      if (ci < nfull) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          w[j] = __byte_perm(__byte_perm(code(v[4 * j]), code(v[4 * j + 1]), 0x0040),
                             __byte_perm(code(v[4 * j + 2]), code(v[4 * j + 3]), 0x0040), 0x5410);
      } else {                                          // columns past the last key never win the maximum
        uint32_t c[W];
#pragma unroll
        for (int j = 0; j < W; ++j) c[j] = (ci * 32 + cbase + j < N) ? code(v[j]) : 0x80u;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          w[j] = __byte_perm(__byte_perm(c[4 * j], c[4 * j + 1], 0x0040), __byte_perm(c[4 * j + 2], c[4 * j + 3], 0x0040),
                             0x5410);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) m4 = __vmaxs4(m4, w[j]);
      sts_v4(crow + ci * 32, w[0], w[1], w[2], w[3]);
    });
    sts_s32(smem_u32(sMax) + 4 * (q * 128 + r_loc), max(max(static_cast<int>(static_cast<int8_t>(m4)), static_cast<int>(static_cast<int8_t>(m4 >> 8))),
                                max(static_cast<int>(static_cast<int8_t>(m4 >> 16)), static_cast<int>(static_cast<int8_t>(m4 >> 24)))));
  }
  tc_fence_before();
  __syncthreads();

  // ---------------------------------------------------------------- phase B: warp = one query row per step
  // rows warp, warp + 8, ...: (row & 7) is a constant of the warp, so is the XOR term of the SWIZZLE_128B
  // address.  Two rows per iteration, branch-free (NCH = compile-time bound of the 32-column chunks; predicates
  // are warp-uniform except the column guard), so the two rows' dependent chains interleave.
  {
    const int left = (N / 8) & 3;                       // whole 8-vectors of the partial chunk (torch's size_ilp rest)
    const int ntail0 = left * 8, ntail1 = N - nfull * 32;
    const uint32_t sw = static_cast<uint32_t>(warp & 7);
    uint32_t poff[NCH];                                 // byte offset of column 32 i + lane in a P row (no row term)
    const bool lane_tail = lane < ntail1;               // this lane's column of the partial chunk is a key
    auto cvalid = [&](int i) -> bool { return i < nfull || (i == nfull && lane_tail); };
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const uint32_t col = 32u * i + lane;
      poff[i] = smem_u32(sP) + (col >> 7) * 16384u + ((((col & 127u) >> 4) ^ sw) << 4) + (col & 15u);
    }
    const uint32_t sC32 = smem_u32(sC) + lane, sMax32 = smem_u32(sMax);
    constexpr int U = 2;
    for (int r0 = warp; r0 < rows_here; r0 += 8 * U) {
      int rr[U];
      bool rv[U];
      float e[U][NCH], acc[U], elast[U], pl[U], sum[U], fs[U];
      uint32_t tidx[U];                                 // element index of table entry (row max, code 0)
      int csat[U];                                      // codes <= csat hold the table's saturated value
      float xmax[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        rv[u] = r0 + 8 * u < rows_here;
        rr[u] = rv[u] ? r0 + 8 * u : r0;
        const int cmax = max(lds_s32(sMax32 + 4 * rr[u]), lds_s32(sMax32 + 4 * (128 + rr[u])));
        xmax[u] = TABLE == 0 ? x_eff(static_cast<float>(cmax), a.p.s_attn) : 0.f;
        tidx[u] = TABLE == 1 ? static_cast<uint32_t>((cmax + 128) * 256 + 128) : static_cast<uint32_t>(cmax);
        csat[u] = a.p.sat_k > 0 ? cmax - a.p.sat_k : -1000;
      }
      // every lane loads unconditionally: a column past the last key reads a stale code byte (any int8 is a
      // valid table index) and its term is zeroed afterwards -- no predicated address arithmetic
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t crow = sC32 + rr[u] * CROW;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          const int c = lds_s8(crow + 32 * i);
          if (TABLE == 1) e[u][i] = c > csat[u] ? __ldg(a.exp_table + (tidx[u] + static_cast<uint32_t>(c))) : a.p.sat_e;
          else if (TABLE == 2) e[u][i] = __ldg(a.exp_table_1d + ((tidx[u] - static_cast<uint32_t>(c)) & 255u));   // stale codes wrap
          else e[u][i] = softmax_term(c, xmax[u], a);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int i = 0; i < NCH; ++i)
          if (i >= NCH - 3) e[u][i] = cvalid(i) ? e[u][i] : 0.f;      // NCH - 3 <= nfull by the dispatch below
      // float32 row sum in torch's CPU order: lane = accumulator (col % 32), ascending columns; then the
      // cascade p8[l] = acc[l] (+ the partial chunk's whole vectors) + acc[l + 8] + acc[l + 16] + acc[l + 24],
      // sum = tail scalars + p8[0] + ... + p8[7]
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u] = 0.f;
        elast[u] = 0.f;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          acc[u] = i < nfull ? __fadd_rn(acc[u], e[u][i]) : acc[u];
          elast[u] = i == nfull ? e[u][i] : elast[u];
        }
        pl[u] = acc[u];
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          const float t = __shfl_sync(0xffffffffu, elast[u], (lane & 7) + 8 * v);
          pl[u] = v < left ? __fadd_rn(pl[u], t) : pl[u];
        }
#pragma unroll
        for (int k = 1; k < 4; ++k) pl[u] = __fadd_rn(pl[u], __shfl_sync(0xffffffffu, acc[u], (lane & 7) + 8 * k));
        sum[u] = 0.f;
#pragma unroll
        for (int j = 0; j < 7; ++j) {
          const float t = __shfl_sync(0xffffffffu, elast[u], (ntail0 + j) & 31);
          sum[u] = ntail0 + j < ntail1 ? __fadd_rn(sum[u], t) : sum[u];
        }
#pragma unroll
        for (int l = 0; l < 8; ++l) sum[u] = __fadd_rn(sum[u], __shfl_sync(0xffffffffu, pl[u], l));
        if (a.p.kind == 0) sum[u] = fminf(sum[u], 2147483648.f);         // ivit_modules.py:173
        if (a.p.kind == 2) sum[u] = fmaxf(sum[u], 1.f);                  // ppoly_modules.py:318
        const float factor = floorf(__fdiv_rn(a.num, sum[u]));            // :174 / ibert :314
        fs[u] = __fmul_rn(factor, a.out_shift);                           // power-of-two scaling is exact
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t prow = (rr[u] >> 3) * 1024 + (rr[u] & 7) * 128;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          // :175 / :315: floor of a value in [0, 2^22) through the 2^23 magic constant (round toward zero)
          const uint32_t pc = __float_as_uint(__fadd_rz(__fmul_rn(e[u][i], fs[u]), 8388608.f));
          if (rv[u] && i < nchunks) sts_u8(prow + poff[i], pc);
          if (TAP && a.probs != nullptr && rv[u] && cvalid(i))
            a.probs[(static_cast<long long>(blockIdx.x) * N + mt * 128 + rr[u]) * N + 32 * i + lane] = static_cast<uint8_t>(pc);
        }
      }
    }
  }
  fence_proxy_async_smem();     // generic-proxy smem writes -> visible to tcgen05.mma
  __syncthreads();

  // ---------------------------------------------------------------- O = P V
  if (threadIdx.x == 0) {
    mbar_wait(&bars[1], 0);
    tc_fence_after();
    const uint32_t idesc2 = make_idesc_i8(128, HD, false, true, false, true);   // A u8 K-major, B s8 MN-major
    const int ksteps = a.KP2 / 32;
    for (int k = 0; k < ksteps; ++k) {
      const uint32_t pa = smem_u32(sP) + (k >> 2) * 16384 + (k & 3) * 32;
      const uint64_t da = make_smem_desc(pa, 16, 1024, kLayoutSW128);
      const uint64_t db = make_smem_desc(smem_u32(sV) + k * 32 * HD, 512, 512, kLayoutSW64);
      mma_i8_ss(tmem_base, da, db, idesc2, k != 0 ? 1u : 0u);
    }
    tc_commit(&bars[3]);
  }
  __syncwarp();
  mbar_wait(&bars[3], 0);
  tc_fence_after();
  if (warp_active) {
    uint32_t o[OW];
    tmem_ld_w(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + q * OW, o);
    tmem_ld_wait();
    if (row < N && a.out != nullptr) {
      auto oc = [&](uint32_t acc) -> uint32_t {
        return static_cast<uint32_t>(clampi(requant_i32_cvt(static_cast<int>(acc), a.p.mult_out), a.p.qmin_out, a.p.qmax_out));
      };
      uint32_t w[OW / 4];
#pragma unroll
      for (int j = 0; j < OW / 4; ++j)
        w[j] = __byte_perm(__byte_perm(oc(o[4 * j]), oc(o[4 * j + 1]), 0x0040),
                           __byte_perm(oc(o[4 * j + 2]), oc(o[4 * j + 3]), 0x0040), 0x5410);
      int8_t* op = a.out + (static_cast<long long>(b) * N + row) * C + h * HD + q * OW;
#pragma unroll
      for (int j = 0; j < OW / 16; ++j)
        *reinterpret_cast<int4*>(op + 16 * j) = make_int4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
    }
  }
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

static void derive_softmax_consts(AttnArgs& a, const ivit_attention_params_t* p) {
  a.p = *p;
  const float s = p->s_attn;
  if (p->kind == 0) {
    a.x0 = floorf(-1.0f / s);                          // ivit_modules.py:154
    a.nx0 = 15.f * a.x0;                               // :155
    a.b_int = a.c_int = a.exp_scale = 0.f;
    a.num = 2147483648.f;                              // float32(2^31 - 1)
    a.out_shift = 1.f / ldexpf(1.f, 31 - p->output_bit + 1);
  } else {
    const double c0 = 0.35815147, c1 = 0.96963238 / c0, c2 = 1.0 / c0;   // ibert_modules.py:265-267
    a.x0 = floorf(static_cast<float>(-0.6931) / s);    // :287
    a.nx0 = 30.f * a.x0;                               // :288
    a.b_int = floorf(static_cast<float>(c1) / s);      // :277
    a.c_int = floorf(static_cast<float>(c2) / (s * s));   // :278
    a.exp_scale = (static_cast<float>(c0) * (s * s)) / 1073741824.f;     // :282, :294
    a.num = 4294967296.f;                              // 2^32 (:314)
    a.out_shift = 1.f / ldexpf(1.f, 32 - p->output_bit + 1);
  }
}

// threads per query row: 2 (8 warps per CTA; default: measured 1.60 ms/step on DeiT-S) or 4 (16 warps
// per CTA, 64 registers, no spills: 1.67 ms/step -- more warps do not help, the passes are throughput-bound)
static int g_attn_tpr = 2;
// 0 / 1: attention_i8_kernel (thread = query row; measured faster: 1.30 vs 1.58 ms per DeiT-S step);
// 2: an 8-bit softmax output runs attention_rows_kernel
static int g_attn_variant = 0;

static int attention_launch(const int8_t* qkv, int8_t* out, uint8_t* probs, int B, int N, int H, int hd,
                            const ivit_attention_params_t* p, cudaStream_t st) {
  if (!p) return fail(IVIT_ERR_BAD_ARG, "attention: params NULL");
  if (p->kind < 0 || p->kind > 2) return fail(IVIT_ERR_BAD_ARG, "attention: kind 0 (ivit), 1 (ibert) or 2 (ppoly)");
  if (p->kind == 2 && p->exp_table == nullptr && p->exp_table_1d == nullptr)
    return fail(IVIT_ERR_BAD_ARG, "attention: the ppoly softmax (kind 2) takes its exp term from a caller-built table");
  if (hd != HD) return fail(IVIT_ERR_UNSUPPORTED, "attention: head_dim must be 64");
  if (N < 1 || N > 256) return fail(IVIT_ERR_UNSUPPORTED, "attention: 1 <= N <= 256 tokens");
  if (p->output_bit < 2 || p->output_bit > 16)
    return fail(IVIT_ERR_UNSUPPORTED, "attention: softmax output_bit must be in [2, 16]");
  const bool p16 = p->output_bit > 8;
  if (p16 && !(p->s_pv > 0.f)) return fail(IVIT_ERR_BAD_ARG, "attention: s_pv (scale of the P.V accumulator) required at 16 bit");
  if (B * H <= 0) return 0;
  AttnArgs a;
  a.B = B; a.N = N; a.H = H;
  a.NKP = (N + 15) / 16 * 16;
  a.KP2 = (N + 31) / 32 * 32;
  a.out = out; a.probs = probs; a.exp_table = p->exp_table; a.exp_table_1d = p->exp_table_1d;
  derive_softmax_consts(a, p);
  a.s_pv = p->s_pv;
  const int C = H * hd;
  CUtensorMap tq, tk, tv;
  int rc;
  const uint64_t rows = static_cast<uint64_t>(B) * N, cols = 3ull * C;
  if ((rc = make_tmap_2d_u8(&tq, qkv, cols, rows, cols, HD, 128, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  if ((rc = make_tmap_2d_u8(&tk, qkv, cols, rows, cols, HD, a.NKP, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  if ((rc = make_tmap_2d_u8(&tv, qkv, cols, rows, cols, HD, a.KP2, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  const int table = a.exp_table_1d != nullptr ? 2 : (a.exp_table != nullptr ? 1 : 0);
  const bool tap = probs != nullptr;
  const size_t smem = 1024 + (p16 ? 65536 : 32768) + 8192 + 16384 + 16384 + (p16 ? 0 : 32 * 128 * 4) + 512 * 4 + 128 * 4 + 64 +
                      (table == 2 ? 256 * 16 * 4 : 0);
  const int tpr = p16 ? 2 : g_attn_tpr;
  a.crow = a.NKP + (((a.NKP >> 4) & 1) ? 0 : 16);
  if (!p16 && g_attn_variant == 2) {
    // 8-bit softmax output: the row-per-warp-step kernel
    const size_t smem2 = 1024 + 32768 + 8192 + 16384 + 16384 + 128 * static_cast<size_t>(a.crow) + 256 * 4 + 64;
    auto launch2 = [&](auto kernel) -> int {
      cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem2));
      if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
      e = launch_kernel(kernel, dim3(B * H, (N + 127) / 128), dim3(256), smem2, st, tq, tk, tv, a);
      if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
      return 0;
    };
    const int nch = (N + 31) / 32;
    auto pick = [&](auto nchc, auto nfixc) -> int {
      constexpr int NC = decltype(nchc)::value, NF = decltype(nfixc)::value;
      if (table == 2) return tap ? launch2(attention_rows_kernel<2, true, NC, NF>) : launch2(attention_rows_kernel<2, false, NC, NF>);
      if (table == 1) return tap ? launch2(attention_rows_kernel<1, true, NC, NF>) : launch2(attention_rows_kernel<1, false, NC, NF>);
      return tap ? launch2(attention_rows_kernel<0, true, NC, NF>) : launch2(attention_rows_kernel<0, false, NC, NF>);
    };
    using I0 = std::integral_constant<int, 0>;
    if (N == 197) rc = pick(std::integral_constant<int, 7>(), std::integral_constant<int, 197>());
    else
    if (nch <= 2) rc = pick(std::integral_constant<int, 2>(), I0());
    else if (nch <= 4) rc = pick(std::integral_constant<int, 4>(), I0());
    else if (nch <= 7) rc = pick(std::integral_constant<int, 7>(), I0());
    else rc = pick(std::integral_constant<int, 8>(), I0());
    if (rc) return rc;
    return check_launch("attention_rows_kernel");
  }
  auto launch = [&](auto kernel) -> int {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    dim3 grid(B * H, (N + 127) / 128);
    e = launch_kernel(kernel, grid, dim3(128 * tpr), smem, st, tq, tk, tv, a);
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    return 0;
  };
  if (p16) {
    if (table == 2) rc = tap ? launch(attention_i8_kernel<2, 2, true, true>) : launch(attention_i8_kernel<2, 2, false, true>);
    else if (table == 1) rc = tap ? launch(attention_i8_kernel<2, 1, true, true>) : launch(attention_i8_kernel<2, 1, false, true>);
    else rc = tap ? launch(attention_i8_kernel<2, 0, true, true>) : launch(attention_i8_kernel<2, 0, false, true>);
  } else if (tpr == 4) {
    if (table == 2) rc = tap ? launch(attention_i8_kernel<4, 2, true>) : launch(attention_i8_kernel<4, 2, false>);
    else if (table == 1) rc = tap ? launch(attention_i8_kernel<4, 1, true>) : launch(attention_i8_kernel<4, 1, false>);
    else rc = tap ? launch(attention_i8_kernel<4, 0, true>) : launch(attention_i8_kernel<4, 0, false>);
  } else {
    if (table == 2) rc = tap ? launch(attention_i8_kernel<2, 2, true>) : launch(attention_i8_kernel<2, 2, false>);
    else if (table == 1) rc = tap ? launch(attention_i8_kernel<2, 1, true>) : launch(attention_i8_kernel<2, 1, false>);
    else rc = tap ? launch(attention_i8_kernel<2, 0, true>) : launch(attention_i8_kernel<2, 0, false>);
  }
  if (rc) return rc;
  return check_launch("attention_i8_kernel");
}

}  // namespace ivit

using namespace ivit;

extern "C" int ivit_attention_set_threads_per_row(int tpr) {
  const int prev = g_attn_tpr;
  if (tpr == 2 || tpr == 4) g_attn_tpr = tpr;
  return prev;
}

#ifdef IVIT_ATTN_TIMING
extern "C" int ivit_debug_attn_timing(unsigned long long* out16, int reset) {
  cudaError_t e = cudaMemcpyFromSymbol(out16, g_attn_t, sizeof(unsigned long long) * 16);
  if (e == cudaSuccess && reset) {
    unsigned long long z[16] = {};
    e = cudaMemcpyToSymbol(g_attn_t, z, sizeof(z));
  }
  return static_cast<int>(e);
}
#endif

extern "C" int ivit_attention_set_variant(int variant) {
  const int prev = g_attn_variant;
  if (variant >= 0 && variant <= 2) g_attn_variant = variant;
  return prev;
}

extern "C" int ivit_attention_i8(const int8_t* qkv, int8_t* out, int B, int N, int H, int hd,
                                 const ivit_attention_params_t* p, void* stream) {
  return attention_launch(qkv, out, nullptr, B, N, H, hd, p, static_cast<cudaStream_t>(stream));
}

extern "C" int ivit_softmax_table_build(const ivit_attention_params_t* p, float* table65536, void* stream) {
  if (!p || !table65536) return fail(IVIT_ERR_BAD_ARG, "softmax table: NULL argument");
  AttnArgs a{};
  derive_softmax_consts(a, p);
  softmax_table_kernel<<<256, 256, 0, static_cast<cudaStream_t>(stream)>>>(table65536, a);
  return check_launch("softmax_table_kernel");
}

extern "C" int ivit_softmax_table_collapse(const float* table65536, float* table256, int* collapsible, void* stream) {
  if (!table65536 || !table256 || !collapsible) return fail(IVIT_ERR_BAD_ARG, "softmax table collapse: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  softmax_table_collapse_init<<<1, 256, 0, st>>>(table65536, table256, collapsible);
  softmax_table_collapse_check<<<256, 256, 0, st>>>(table65536, table256, collapsible);
  return check_launch("softmax_table_collapse");
}

extern "C" int ivit_attention_probs_i8(const int8_t* qkv, uint8_t* probs, int B, int N, int H, int hd,
                                       const ivit_attention_params_t* p, void* stream) {
  // the regular output is discarded into a scratch row range of `probs`' owner: callers pass a
  // valid `out` through ivit_attention_i8 when they need both; here only the tap is written.
  return attention_launch(qkv, nullptr, probs, B, N, H, hd, p, static_cast<cudaStream_t>(stream));
}
