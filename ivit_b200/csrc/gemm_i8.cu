// INT8 x INT8 -> INT32 GEMM on the 5th-gen tensor cores (tcgen05.mma kind::i8, accumulators in
// TMEM, operands staged by TMA into 128B-swizzled shared memory) with the QuantAct
// requantisation(s) that follow QuantLinear fused into the epilogue.
//
//   C[M,N] = A[M,K] * W[N,K]^T      A, W int8, K contiguous (both operands K-major)
//
// CTA = one 128 x BN output tile, 4 warps: warp 0 lane* = TMA producer, warp 1 lane* = MMA
// issuer, then all four warps drain their 32 TMEM lanes (thread = output row) through the
// epilogue.  Several CTAs are resident per SM (TMEM: BN columns each), so one CTA's
// epilogue overlaps another's main loop.
#include <cuda.h>
#include <cmath>
#include <cstdio>
#include <type_traits>

#include "intmath.cuh"
#include "ivit_b200.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace ivit {
int fail(int code, const char* msg);
int check_launch(const char* what);

constexpr int BM = 128;
constexpr int BK = 128;          // bytes of K per stage = one 128B swizzle row
constexpr int UMMA_K = 32;       // K per tcgen05.mma for 8-bit operands

struct GemmArgs {
  int M, N, K;
  int stages;
  int allow_fast;
  void* out;
  long long ldo;                 // in elements of the output type
  ivit_gemm_epilogue_t e;
};

// ------------------------------------------------------------------ shared epilogue math
__device__ __forceinline__ long long out_row_of(int r, const ivit_gemm_epilogue_t& e) {
  if (e.group <= 0) return r;
  return static_cast<long long>(r / e.group) * e.out_group_stride + (r % e.group) + e.out_row_off;
}
__device__ __forceinline__ long long id_row_of(int r, const ivit_gemm_epilogue_t& e) {
  if (e.group <= 0) return r;
  if (e.id_broadcast) return (r % e.group) + e.out_row_off;
  return out_row_of(r, e);
}

// One element, general (slow-path capable) form.  Used by the simple kernel and by the
// tcgen05 epilogue when a chunk leaves the fast range.
__device__ __forceinline__ int epi_requant_elem(int z, int n, const ivit_gemm_epilogue_t& e) {
  const double r = requant_any(z, __ldg(e.s_in + n), __ldg(e.mult + n));
  return static_cast<int>(fmin(fmax(r, static_cast<double>(e.qmin)), static_cast<double>(e.qmax)));
}
__device__ __forceinline__ int epi_residual_elem(int c1, int id, const ivit_gemm_epilogue_t& e) {
  const double r = requant_d(static_cast<double>(c1), e.mult2) + requant_d(static_cast<double>(id), e.mult_id);
  return static_cast<int>(fmin(fmax(r, static_cast<double>(e.qmin2)), static_cast<double>(e.qmax2)));
}

// ------------------------------------------------------------------ tcgen05 kernel
// Persistent, warp-specialised: one CTA per SM walks output tiles t = blockIdx.x + i*gridDim.x
// (n fastest, so the CTAs running together share A row-blocks in L2).
//   warp 0      TMA producer: A [128 x 128 B] + W [BN x 128 B] per stage, SWIZZLE_128B, STAGES deep
//   warp 1      MMA issuer (one elected lane): 4 x tcgen05.mma kind::i8 (K = 32) per stage into one
//               of TWO TMEM accumulators, so tile i+1 accumulates while tile i drains
//   warps 2..   epilogue: warp w reads TMEM lanes 32*(w%4).. (thread = output row); the BN/64
//               warps of a lane quarter take 64 columns each, in 16-column chunks with the next
//               chunk's tcgen05.ld in flight while this one is requantised; 16-byte stores.
//               (BN = 192: 12 epilogue warps, 14 in all -- registers are granted per 4 warps, so
//               16 + 2 warps would cap the epilogue at 96 registers and spill.)

template <int BN>
struct WsCfg {
  static constexpr int A_BYTES = BM * BK;
  static constexpr int B_BYTES = BN * BK;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = BN <= 128 ? 6 : (BN <= 192 ? 5 : 4);
  static constexpr int ACC_STRIDE = BN <= 128 ? 128 : 256;       // TMEM columns between the two accumulators
  static constexpr uint32_t TMEM_COLS = 2 * ACC_STRIDE;
  static constexpr int PART = 64;                                // columns per epilogue warp
  static constexpr int NPARTS = BN / PART;
  static constexpr int EPI_WARPS = 4 * NPARTS;
  static constexpr int THREADS = 64 + 32 * EPI_WARPS;
  static constexpr int NCH = PART / 16;                          // 16-column chunks per warp
  static constexpr int PAR_STRIDE = PART * 28;                   // per warp: {mult, kb} f64x2, mult f64, bias i32
  static constexpr size_t SMEM = 1024 + STAGES * STAGE_BYTES + 256 + EPI_WARPS * PAR_STRIDE;
  static_assert(BN % PART == 0, "BN must be a multiple of 64");
};

// Two int32 -> two saturated int8, packed below `hi16` (PTX cvt.pack.sat: d = hi16 << 16 | sat(a) << 8 | sat(b)).
__device__ __forceinline__ uint32_t pack_sat_s8(int a, int b, uint32_t hi16) {
  uint32_t d;
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(hi16));
  return d;
}
__device__ __forceinline__ uint32_t pack4_sat_s8(int v0, int v1, int v2, int v3) {
  return pack_sat_s8(v1, v0, pack_sat_s8(v3, v2, 0u));
}
// first requant of the channels-on-lanes epilogue: 0 = XOR + DADD + DFMA (default); 1 = I2F.F64 + DFMA -- one FP64 slot
// less per output, but the conversion pipe (16 lanes / clk / SM) then bounds the epilogue: measured slower
// (qkv 0.436 -> 0.459 ms, fc1 0.546 -> 0.584 ms per DeiT-S step)
#ifndef IVIT_CL_I2F
#define IVIT_CL_I2F 0
#endif
constexpr double kMagicInt = 4503601774854144.0;      // 2^52 + 2^31
constexpr double kMagicRnd = 6755399441055744.0;      // 1.5 * 2^52

// General 16-column chunk of one output row (thread = row): every epilogue, range-checked fast
// requant with the per-element replay path behind it.
__device__ __noinline__ void epilogue_chunk_generic(const uint32_t (&acc)[16], const int32_t* __restrict__ s_bias,
                                                    const double* __restrict__ s_mult, int nb, bool row_ok,
                                                    long long orow, long long irow, const GemmArgs& args) {
  const ivit_gemm_epilogue_t& e = args.e;
  const bool full_chunk = nb + 16 <= args.N;
  int z[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) z[j] = static_cast<int>(acc[j]) + s_bias[j];

  if (e.epilogue == IVIT_EPI_ACC32) {
    if (row_ok) {
      int32_t* o = static_cast<int32_t*>(args.out) + orow * args.ldo + nb;
      if (full_chunk) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) *reinterpret_cast<int4*>(o + j) = make_int4(z[j], z[j + 1], z[j + 2], z[j + 3]);
      } else {
        for (int j = 0; j < 16 && nb + j < args.N; ++j) o[j] = z[j];
      }
    }
    return;
  }
  if (e.epilogue == IVIT_EPI_LOGITS) {
    if (row_ok) {
      float* o = static_cast<float*>(args.out) + orow * args.ldo + nb;
      for (int j = 0; j < 16 && nb + j < args.N; ++j)
        o[j] = __fmul_rn(static_cast<float>(z[j]), __ldg(e.scale_out + nb + j));
    }
    return;
  }
  // REQUANT / RESIDUAL: first fixedpoint_mul (per-channel)
  uint32_t bits = 0;                 // OR of |z| (one's complement for negatives): < 2^22 <=> all in range
#pragma unroll
  for (int j = 0; j < 16; ++j) bits |= static_cast<uint32_t>(z[j] ^ (z[j] >> 31));
  const bool ok = args.allow_fast != 0 && bits < (1u << 22);
  int c[16];
  if (__all_sync(0xffffffffu, ok)) {
#pragma unroll
    for (int j = 0; j < 16; ++j) c[j] = clampi(requant_i32_fast(z[j], s_mult[j]), e.qmin, e.qmax);
  } else {
    for (int j = 0; j < 16; ++j) c[j] = (nb + j < args.N) ? epi_requant_elem(z[j], nb + j, e) : 0;
  }
  if (e.epilogue == IVIT_EPI_RESIDUAL) {
    int idv[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) idv[j] = 0;
    if (row_ok) {
      if (e.id_bits == 16) {               // 16-bit residual stream: identity int16, ld_id in elements
        const int16_t* ip = reinterpret_cast<const int16_t*>(e.identity) + irow * e.ld_id + nb;
        if (full_chunk && (reinterpret_cast<uintptr_t>(ip) & 15) == 0) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int4 a = __ldg(reinterpret_cast<const int4*>(ip) + h);
            const uint32_t w4[4] = {static_cast<uint32_t>(a.x), static_cast<uint32_t>(a.y), static_cast<uint32_t>(a.z),
                                    static_cast<uint32_t>(a.w)};
#pragma unroll
            for (int j = 0; j < 8; ++j) idv[8 * h + j] = static_cast<int16_t>(w4[j >> 1] >> (16 * (j & 1)));
          }
        } else {
          for (int j = 0; j < 16 && nb + j < args.N; ++j) idv[j] = ip[j];
        }
      } else {
        const int8_t* ip = e.identity + irow * e.ld_id + nb;
        if (full_chunk && (reinterpret_cast<uintptr_t>(ip) & 15) == 0) {
          const int4 a = __ldg(reinterpret_cast<const int4*>(ip));
          const uint32_t w4[4] = {static_cast<uint32_t>(a.x), static_cast<uint32_t>(a.y), static_cast<uint32_t>(a.z),
                                  static_cast<uint32_t>(a.w)};
#pragma unroll
          for (int j = 0; j < 16; ++j) idv[j] = static_cast<int8_t>(w4[j >> 2] >> (8 * (j & 3)));
        } else {
          for (int j = 0; j < 16 && nb + j < args.N; ++j) idv[j] = ip[j];
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int id = idv[j];
      // second fixedpoint_mul (scalar ratio) + identity branch, rounded separately then added
      if (args.allow_fast) {
        c[j] = clampi(requant_i32_fast(c[j], e.mult2) + requant_i32_fast(id, e.mult_id), e.qmin2, e.qmax2);
      } else {
        c[j] = epi_residual_elem(c[j], id, e);
      }
    }
  }
  if (row_ok && e.out_bits == 16) {        // 16-bit activation location: int16 codes, ldo in elements
    int16_t* o = static_cast<int16_t*>(args.out) + orow * args.ldo + nb;
    if (full_chunk && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          w[j] = (static_cast<uint32_t>(c[8 * h + 2 * j]) & 0xffffu) | (static_cast<uint32_t>(c[8 * h + 2 * j + 1]) << 16);
        reinterpret_cast<int4*>(o)[h] = make_int4(w[0], w[1], w[2], w[3]);
      }
    } else {
      for (int j = 0; j < 16 && nb + j < args.N; ++j) o[j] = static_cast<int16_t>(c[j]);
    }
    return;
  }
  if (row_ok) {
    int8_t* o = static_cast<int8_t*>(args.out) + orow * args.ldo + nb;
    if (full_chunk && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        w[j] = (c[4 * j] & 0xff) | ((c[4 * j + 1] & 0xff) << 8) | ((c[4 * j + 2] & 0xff) << 16) |
               (static_cast<uint32_t>(c[4 * j + 3]) << 24);
      *reinterpret_cast<int4*>(o) = make_int4(w[0], w[1], w[2], w[3]);
    } else {
      for (int j = 0; j < 16 && nb + j < args.N; ++j) o[j] = static_cast<int8_t>(c[j]);
    }
  }
}

// Streamlined 16-column chunk for the hot cases: int8 saturating outputs, full chunk, |acc + bias|
// < 2^22 (proven by the caller, or CHECKED here per chunk: returns false -> the caller re-runs the
// chunk through the general path).  MODE 1 = REQUANT, MODE 2 = RESIDUAL.
//   acc -> fp64 through the 2^52 exponent trick with the bias folded into the subtrahend
//   (kb = 2^52 + 2^31 - bias), one DFMA against 1.5*2^52 leaves rint(z*mult) in the low word,
//   cvt.pack.sat clamps to [-128,127] and packs two codes per instruction.
//   s_par[j] = {mult, kb} of column j.
template <int MODE, bool CHECKED>
__device__ __forceinline__ bool epilogue_chunk_fast(const uint32_t (&acc)[16], const double2* __restrict__ s_par,
                                                    const int32_t* __restrict__ s_bias, int nb, bool row_ok,
                                                    long long orow, long long irow, const GemmArgs& args) {
  const ivit_gemm_epilogue_t& e = args.e;
  if (CHECKED) {
    uint32_t bits = 0;                 // OR of |z| (one's complement for negatives): < 2^22 <=> all in range
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int z = static_cast<int>(acc[j]) + s_bias[j];
      bits |= static_cast<uint32_t>(z ^ (z >> 31));
    }
    if (!__all_sync(0xffffffffu, bits < (1u << 22))) return false;
  }
  int4 idv = make_int4(0, 0, 0, 0);
  if (MODE == 2 && row_ok) idv = __ldg(reinterpret_cast<const int4*>(e.identity + irow * e.ld_id + nb));
  int c[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const double2 pm = s_par[j];
    const double zd = __hiloint2double(0x43300000, static_cast<int>(acc[j] ^ 0x80000000u)) - pm.y;
    c[j] = __double2loint(__fma_rn(zd, pm.x, kMagicRnd));
  }
  if (MODE == 2) {
    const uint32_t idw[4] = {static_cast<uint32_t>(idv.x), static_cast<uint32_t>(idv.y), static_cast<uint32_t>(idv.z),
                             static_cast<uint32_t>(idv.w)};
    const double m2 = e.mult2, mi = e.mult_id;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int c1 = clampi(c[j], -128, 127);
      const int id = static_cast<int8_t>(idw[j >> 2] >> (8 * (j & 3)));
      const double cd = __hiloint2double(0x43300000, c1 ^ 0x80000000) - kMagicInt;
      const double dd = __hiloint2double(0x43300000, id ^ 0x80000000) - kMagicInt;
      c[j] = __double2loint(__fma_rn(cd, m2, kMagicRnd)) + __double2loint(__fma_rn(dd, mi, kMagicRnd));
    }
  }
  if (row_ok) {
    int8_t* o = static_cast<int8_t*>(args.out) + orow * args.ldo + nb;
    *reinterpret_cast<int4*>(o) =
        make_int4(pack4_sat_s8(c[0], c[1], c[2], c[3]), pack4_sat_s8(c[4], c[5], c[6], c[7]),
                  pack4_sat_s8(c[8], c[9], c[10], c[11]), pack4_sat_s8(c[12], c[13], c[14], c[15]));
  }
  return true;
}

// MODE 0: generic epilogues; 1 / 2: streamlined REQUANT / RESIDUAL (host checks the preconditions);
// CHECKED: the streamlined chunk verifies |acc + bias| < 2^22 itself.
template <int BN, int MODE, bool CHECKED>
__global__ void __launch_bounds__(WsCfg<BN>::THREADS, 1)
gemm_i8_ws_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const GemmArgs args) {
  using Cfg = WsCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  // SWIZZLE_128B tiles need 1024-byte alignment; indexing the __shared__ array directly (no
  // integer round-trip) keeps the address space, so parameter reads compile to LDS, not generic LD
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* tail = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);            // [STAGES] TMA -> MMA
  uint64_t* empty_bar = full_bar + STAGES;                           // [STAGES] MMA -> TMA
  uint64_t* tfull_bar = empty_bar + STAGES;                          // [2] MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + 2;                              // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint8_t* par = tail + 256;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntn = (args.N + BN - 1) / BN;
  const int ntiles = ((args.M + BM - 1) / BM) * ntn;
  const int kblocks = (args.K + BK - 1) / BK;
  const ivit_gemm_epilogue_t& e = args.e;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], Cfg::EPI_WARPS);
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (elect_one()) {
      int it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int m0 = (t / ntn) * BM, n0 = (t % ntn) * BN;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
          uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
          tma_load_2d(sa, &tmA, &full_bar[s], kb * BK, m0);
          tma_load_2d(sa + Cfg::A_BYTES, &tmW, &full_bar[s], kb * BK, n0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_i8(BM, BN, true, true, false, false);
      int it = 0, i = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++i) {
        const int a = i & 1;
        mbar_wait(&tempty_bar[a], ((i >> 1) & 1) ^ 1);        // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full_bar[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + Cfg::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = make_smem_desc(sa + k * UMMA_K, 16, 1024, kLayoutSW128);
            const uint64_t db = make_smem_desc(sb + k * UMMA_K, 16, 1024, kLayoutSW128);
            mma_i8_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit(&empty_bar[s]);          // smem slot reusable once these MMAs retire
        }
        tc_commit(&tfull_bar[a]);            // accumulator complete
      }
    }
    __syncwarp();
  } else {
    // ===================================================== epilogue
    const int ew = warp - 2;
    const int quarter = warp & 3;            // TMEM lanes this warp may read
    const int part = ew >> 2;                // which 64 columns
    double2* s_par = reinterpret_cast<double2*>(par + ew * Cfg::PAR_STRIDE);      // [PART] {mult, kb}
    double* s_mult = reinterpret_cast<double*>(s_par + Cfg::PART);                // [PART] (general path)
    int32_t* s_bias = reinterpret_cast<int32_t*>(s_mult + Cfg::PART);             // [PART]
    // this lane's two columns of the tile's parameters, fetched one tile ahead
    int pb[2];
    double pm[2];
    auto fetch_params = [&](int t) {
      const int n0 = (t % ntn) * BN + part * Cfg::PART;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int n = n0 + lane + 32 * u;
        pb[u] = (e.bias != nullptr && n < args.N) ? __ldg(e.bias + n) : 0;
        pm[u] = (e.mult != nullptr && n < args.N) ? __ldg(e.mult + n) : 0.0;
      }
    };
    if (blockIdx.x < ntiles) fetch_params(blockIdx.x);
    int i = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++i) {
      const int a = i & 1;
      const int m0 = (t / ntn) * BM, n0 = (t % ntn) * BN + part * Cfg::PART;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int j = lane + 32 * u;
        s_bias[j] = pb[u];
        s_mult[j] = pm[u];
        s_par[j] = make_double2(pm[u], kMagicInt - static_cast<double>(pb[u]));
      }
      __syncwarp();
      if (t + gridDim.x < ntiles) fetch_params(t + gridDim.x);     // lands while this tile is processed
      const int r = m0 + quarter * 32 + lane;
      const bool row_ok = r < args.M;
      const long long orow = row_ok ? out_row_of(r, e) : 0;
      const long long irow = row_ok ? id_row_of(r, e) : 0;
      mbar_wait(&tfull_bar[a], (i >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + a * Cfg::ACC_STRIDE + part * Cfg::PART +
                             (static_cast<uint32_t>(quarter * 32) << 16);
      uint32_t acc[2][16];
      tmem_ld16(t_row, acc[0]);
#pragma unroll
      for (int ch = 0; ch < Cfg::NCH; ++ch) {
        tmem_ld_wait();
        if (ch + 1 < Cfg::NCH) {
          tmem_ld16(t_row + (ch + 1) * 16, acc[(ch + 1) & 1]);
        } else {
          // every TMEM read of this warp has landed: hand the accumulator back before the math
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[a]);
        }
        const int nb = n0 + ch * 16;
        if (nb < args.N) {                   // warp-uniform
          bool done = false;
          if (MODE != 0 && nb + 16 <= args.N)
            done = epilogue_chunk_fast<MODE == 0 ? 1 : MODE, CHECKED>(acc[ch & 1], s_par + ch * 16, s_bias + ch * 16, nb,
                                                                     row_ok, orow, irow, args);
          if (!done) {
            uint32_t tmp[16];                // the out-of-line call takes its operand through memory:
#pragma unroll                               // keep that away from the register-resident hot arrays
            for (int j = 0; j < 16; ++j) tmp[j] = acc[ch & 1][j];
            epilogue_chunk_generic(tmp, s_bias + ch * 16, s_mult + ch * 16, nb, row_ok, orow, irow, args);
          }
        }
      }
      __syncwarp();                          // s_* are rewritten at the top of the next tile
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ channels-on-lanes kernel
// Same pipeline with the operand roles swapped: D^T[n, token] = W[n,:] . A[token,:], i.e. the
// WEIGHT tile (128 output channels) is the M operand and a 192-token tile the N operand.  In TMEM
// a lane is now an output CHANNEL and a column a token, so every per-channel epilogue constant
// (mult, bias) is a per-thread register loaded once per tile -- the row-major kernel above pays a
// 16-byte shared-memory broadcast per column per warp for them, which is what bounds it (l1tex).
// Codes of 4 tokens are packed per register and 4x4-byte-transposed across lane quads (two
// SHFL.BFLY + two PRMT), after which lane 4k+i holds channels 4k..4k+3 of token 4g+i: 4-byte
// stores, 32 contiguous bytes per token per warp.  The residual identity comes in the same way.
// Streamlined REQUANT / RESIDUAL only (no row remap); everything else stays on the kernel above.
template <int TB_>
struct ClCfg {
  static constexpr int CB = 128;                                 // channels per tile (UMMA M)
  static constexpr int TB = TB_;                                 // tokens per tile (UMMA N): 192 or 256
  static constexpr int W_BYTES = CB * BK;
  static constexpr int A_BYTES = TB * BK;
  static constexpr int STAGE_BYTES = W_BYTES + A_BYTES;
  static constexpr int STAGES = TB <= 192 ? 5 : 4;
  static constexpr int ACC_STRIDE = 256;
  static constexpr uint32_t TMEM_COLS = 512;
  static constexpr int PART = 64;                                // tokens per epilogue warp
  static constexpr int NPARTS = TB / PART;
  static constexpr int EPI_WARPS = 4 * NPARTS;                   // 12 (<= 128 registers) or 16 (<= 96)
  static constexpr int THREADS = 64 + 32 * EPI_WARPS;
  static constexpr int NCH = PART / 16;
  static constexpr size_t SMEM = STAGES * STAGE_BYTES + 256;
};

// Walks t = blockIdx.x + i * gridDim.x as (channel block, token block) without dividing per tile.
struct TileWalk {
  int cb, tb, dq, dr, ncb;
  __device__ __forceinline__ TileWalk(int t0, int step, int ncb_) : ncb(ncb_) {
    cb = t0 % ncb_; tb = t0 / ncb_; dq = step / ncb_; dr = step % ncb_;
  }
  __device__ __forceinline__ void next() {
    cb += dr; tb += dq;
    if (cb >= ncb) { cb -= ncb; ++tb; }
  }
};

// 4x4 byte transpose across the 4 lanes of a quad: in: lane i holds bytes T[i][0..3]; out: T[0..3][i]
__device__ __forceinline__ uint32_t quad_transpose(uint32_t x, uint32_t sel1, uint32_t sel2) {
  uint32_t y = __shfl_xor_sync(0xffffffffu, x, 1);
  x = __byte_perm(x, y, sel1);
  y = __shfl_xor_sync(0xffffffffu, x, 2);
  return __byte_perm(x, y, sel2);
}

__device__ __noinline__ int cl_requant_exact(int z, float s_in, double mult, int qmin, int qmax) {
  const double r = requant_any(z, s_in, mult);
  return static_cast<int>(fmin(fmax(r, static_cast<double>(qmin)), static_cast<double>(qmax)));
}

template <int TB, int MODE, bool CHECKED>
__global__ void __launch_bounds__(ClCfg<TB>::THREADS, 1)
gemm_i8_cl_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const GemmArgs args) {
  using Cfg = ClCfg<TB>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* tail = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  // warp index through a shuffle: the compiler then knows the role branches are warp-uniform
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int ncb = (args.N + Cfg::CB - 1) / Cfg::CB;
  const int ntiles = ((args.M + Cfg::TB - 1) / Cfg::TB) * ncb;
  const int kblocks = (args.K + BK - 1) / BK;
  const ivit_gemm_epilogue_t& e = args.e;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], Cfg::EPI_WARPS);
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // programmatic dependent launch: everything above overlapped the predecessor's tail; the grid is one
  // CTA per SM and fully resident, so the successor may be scheduled as soon as CTAs here retire
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================================================== TMA producer (channel block fastest:
    // the CTAs running together read the same token tile; the weights stay in L2 throughout)
    if (elect_one()) {
      int it = 0;
      TileWalk tw(blockIdx.x, gridDim.x, ncb);
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, tw.next()) {
        const int c0 = tw.cb * Cfg::CB, t0 = tw.tb * Cfg::TB;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
          uint8_t* sw = smem + s * Cfg::STAGE_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
          tma_load_2d(sw, &tmW, &full_bar[s], kb * BK, c0);
          tma_load_2d(sw + Cfg::W_BYTES, &tmA, &full_bar[s], kb * BK, t0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_i8(Cfg::CB, Cfg::TB, true, true, false, false);
      int it = 0, i = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++i) {
        const int a = i & 1;
        mbar_wait(&tempty_bar[a], ((i >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
          const int s = it % STAGES;
          mbar_wait(&full_bar[s], (it / STAGES) & 1);
          tc_fence_after();
          const uint32_t sw = smem_u32(smem + s * Cfg::STAGE_BYTES);
          const uint32_t sa = sw + Cfg::W_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t dw = make_smem_desc(sw + k * UMMA_K, 16, 1024, kLayoutSW128);
            const uint64_t da = make_smem_desc(sa + k * UMMA_K, 16, 1024, kLayoutSW128);
            mma_i8_ss(d_tmem, dw, da, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit(&empty_bar[s]);
        }
        tc_commit(&tfull_bar[a]);
      }
    }
    __syncwarp();
  } else {
    // ===================================================== epilogue (thread = output channel)
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int part = ew >> 2;
    const uint32_t sel1 = (lane & 1) ? 0x3715u : 0x6240u;
    const uint32_t sel2 = (lane & 2) ? 0x3276u : 0x5410u;
    const int qi = lane & 3;                 // after the transpose: this lane's token within a group of 4
    const int qc = (lane >> 2) * 4;          //                      and its first channel within the warp's 32
    int pb = 0;
    double pm = 0.0;
    float ps = 0.f;                          // MODE 3: dequantisation scale of the channel
    auto fetch_params = [&](int cb) {
      const int n = cb * Cfg::CB + quarter * 32 + lane;
      pb = (e.bias != nullptr && n < args.N) ? __ldg(e.bias + n) : 0;
      if (MODE == 3) ps = n < args.N ? __ldg(e.scale_out + n) : 0.f;
      else pm = n < args.N ? __ldg(e.mult + n) : 0.0;
    };
    TileWalk tw(blockIdx.x, gridDim.x, ncb), tw_next = tw;
    if (blockIdx.x < ntiles) fetch_params(tw.cb);
    tw_next.next();
    int8_t* const outp = static_cast<int8_t*>(args.out);
    const double m2 = e.mult2, mi = e.mult_id;
    int i = 0;
    const uint32_t ldo = static_cast<uint32_t>(args.ldo), ldi = static_cast<uint32_t>(e.ld_id);   // offsets < 2^31 (host)
    // MODE 4 (RESIDUAL with an int16 stream: int16 identity in, int16 codes out): 2 x 2 transposes of
    // 16-bit halves across lane pairs instead of the 4 x 4 byte transposes; after it the even lane
    // of a pair holds channels (n, n+1) of the even token, the odd lane the same channels of the odd one
    const uint32_t sel16 = (lane & 1) ? 0x3276u : 0x5410u;
    const int16_t* const id16 = reinterpret_cast<const int16_t*>(e.identity);
    int16_t* const out16 = static_cast<int16_t*>(args.out);
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++i, tw.next(), tw_next.next()) {
      const int a = i & 1;
      const int c0 = tw.cb * Cfg::CB + quarter * 32;           // the warp's first channel
      const int tok_w = tw.tb * Cfg::TB + part * Cfg::PART;    // the warp's first token
      const int n = c0 + lane;
      const double mult = pm;
      const float scale = ps;
      const int bias_now = pb;
      const double kb = kMagicInt - static_cast<double>(pb);
      const int brange = pb + (1 << 22);                       // CHECKED: 0 <= acc + brange < 2^23
      if (t + gridDim.x < ntiles) fetch_params(tw_next.cb);
      const int nq = c0 + qc;                                  // channels nq..nq+3 after the transpose
      const bool nq_ok = nq < args.N;                          // N % 4 == 0 (host-checked)

      // lane-private byte offsets of (token tok_w + qi, channel nq) in the output / identity
      const bool all_ok = tok_w + Cfg::PART <= args.M && c0 + 32 <= args.N;   // warp-uniform: no guards
      const uint32_t off_o = MODE == 4 ? static_cast<uint32_t>(tok_w + (lane & 1)) * ldo + (c0 + (lane & ~1))
                                       : static_cast<uint32_t>(tok_w + qi) * ldo + nq;
      const uint32_t off_i = MODE == 4 ? static_cast<uint32_t>(tok_w + (lane & 1)) * ldi + (c0 + (lane & ~1))
                                       : static_cast<uint32_t>(tok_w + qi) * ldi + nq;
      const bool np_ok = c0 + (lane & ~1) < args.N;            // MODE 4: this lane pair's channels exist (N % 4 == 0)
      const uint32_t t_row = tmem_base + a * Cfg::ACC_STRIDE + part * Cfg::PART +
                             (static_cast<uint32_t>(quarter * 32) << 16);
      // One tile.  FULL: every token / channel of the warp's 64 x 32 block exists, no guards.
      // The transposes + stores of chunk k-1 sit in the same basic block as the requantisation
      // of chunk k, so the scheduler can overlap the FP64 pipe with the ALU / LSU work.
      auto process = [&](auto full_tag) {
        constexpr bool FULL = decltype(full_tag)::value;
        auto load_identity = [&](int ch, uint32_t (&w)[4]) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const bool ok = FULL || (nq_ok && tok_w + ch * 16 + 4 * g + qi < args.M);
            w[g] = ok ? __ldg(reinterpret_cast<const uint32_t*>(e.identity + (off_i + (ch * 16 + 4 * g) * ldi))) : 0u;
          }
        };
        auto store_chunk = [&](int ch, const uint32_t (&packed)[4]) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t w = quad_transpose(packed[g], sel1, sel2);
            if (FULL || (nq_ok && tok_w + ch * 16 + 4 * g + qi < args.M))
              *reinterpret_cast<uint32_t*>(outp + (off_o + (ch * 16 + 4 * g) * ldo)) = w;
          }
        };
        // MODE 4: register g of a chunk = tokens (2g, 2g+1); elements (int16) instead of bytes
        auto load_identity16 = [&](int ch, uint32_t (&w)[8]) {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const bool ok = FULL || (np_ok && tok_w + ch * 16 + 2 * g + (lane & 1) < args.M);
            w[g] = ok ? __ldg(reinterpret_cast<const uint32_t*>(id16 + (off_i + (ch * 16 + 2 * g) * ldi))) : 0u;
          }
        };
        auto store_chunk16 = [&](int ch, const uint32_t (&pk)[8]) {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const uint32_t y = __shfl_xor_sync(0xffffffffu, pk[g], 1);
            const uint32_t w = __byte_perm(pk[g], y, sel16);
            if (FULL || (np_ok && tok_w + ch * 16 + 2 * g + (lane & 1) < args.M))
              *reinterpret_cast<uint32_t*>(out16 + (off_o + (ch * 16 + 2 * g) * ldo)) = w;
          }
        };
        uint32_t idw[2][MODE == 4 ? 8 : 4];
        uint32_t packed16[MODE == 4 ? 8 : 1];
        if (MODE == 2) load_identity(0, reinterpret_cast<uint32_t(&)[4]>(idw[0]));
        if (MODE == 4) load_identity16(0, reinterpret_cast<uint32_t(&)[8]>(idw[0]));
        mbar_wait(&tfull_bar[a], (i >> 1) & 1);
        tc_fence_after();
        uint32_t acc[2][16];
        uint32_t packed[4];
        tmem_ld16(t_row, acc[0]);
#pragma unroll
        for (int ch = 0; ch < Cfg::NCH; ++ch) {
          tmem_ld_wait();
          if (ch + 1 < Cfg::NCH) {
            tmem_ld16(t_row + (ch + 1) * 16, acc[(ch + 1) & 1]);
            if (MODE == 2) load_identity(ch + 1, reinterpret_cast<uint32_t(&)[4]>(idw[(ch + 1) & 1]));
            if (MODE == 4) load_identity16(ch + 1, reinterpret_cast<uint32_t(&)[8]>(idw[(ch + 1) & 1]));
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[a]);
          }
          const uint32_t(&z)[16] = acc[ch & 1];        // rows past M are TMA zero fill: harmless, stores are guarded
          if (MODE == 3) {
            // LOGITS (head, vit_quant.py:311): float32(acc + bias) * scale[n]; lane = channel, so a
            // token's 32 values leave as one 128-byte row segment
            float* const outf = static_cast<float*>(args.out);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int tok = tok_w + ch * 16 + j;
              if (n < args.N && tok < args.M)
                outf[static_cast<uint32_t>(tok) * ldo + n] =
                    __fmul_rn(static_cast<float>(static_cast<int>(z[j]) + bias_now), scale);
            }
            continue;
          }
          int c[16];
          bool fast = true;
          if (CHECKED) {
            uint32_t bits = 0;
#pragma unroll
            for (int j = 0; j < 16; ++j) bits |= static_cast<uint32_t>(static_cast<int>(z[j]) + brange);
            fast = __all_sync(0xffffffffu, (bits & 0xff800000u) == 0u);
          }
          if (fast) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              // int32 -> fp64 by I2F.F64 (conversion pipe) + one DFMA, or XOR + DADD + DFMA (two FP64 slots)
              const double zd = IVIT_CL_I2F ? __int2double_rn(static_cast<int>(z[j]) + bias_now)
                                            : __hiloint2double(0x43300000, static_cast<int>(z[j] ^ 0x80000000u)) - kb;
              c[j] = __double2loint(__fma_rn(zd, mult, kMagicRnd));
            }
          } else {
            const float s_in = n < args.N ? __ldg(e.s_in + n) : 1.0f;      // rare: out-of-range chunk
            const int bias = brange - (1 << 22);
#pragma unroll
            for (int j = 0; j < 16; ++j)
              c[j] = cl_requant_exact(static_cast<int>(z[j]) + bias, s_in, mult, e.qmin, e.qmax);
          }
          if (MODE == 4) {
            if (ch > 0) store_chunk16(ch - 1, reinterpret_cast<const uint32_t(&)[8]>(packed16));
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const uint32_t y = __shfl_xor_sync(0xffffffffu, idw[ch & 1][g], 1);
              const uint32_t iw = __byte_perm(idw[ch & 1][g], y, sel16);     // this channel's identity, tokens 2g, 2g+1
              int r2[2];
#pragma unroll
              for (int b = 0; b < 2; ++b) {
                const int c1 = clampi(c[2 * g + b], e.qmin, e.qmax);
                const int id = static_cast<int16_t>(iw >> (16 * b));
                // codes -> fp64 through I2F.F64 (conversion pipe) instead of XOR + MOV + DADD each, as in MODE 2
                const double cd = __int2double_rn(c1), dd = __int2double_rn(id);
                r2[b] = clampi(__double2loint(__fma_rn(cd, m2, kMagicRnd)) + __double2loint(__fma_rn(dd, mi, kMagicRnd)),
                               e.qmin2, e.qmax2);
              }
              packed16[g] = (static_cast<uint32_t>(r2[0]) & 0xffffu) | (static_cast<uint32_t>(r2[1]) << 16);
            }
            continue;
          }
          if (ch > 0) store_chunk(ch - 1, packed);
          if (MODE == 2) {
            // identity words arrive token-major (lane = token, bytes = channels): transpose to
            // channel-major (lane = channel, bytes = tokens) to meet the accumulators
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint32_t iw = quad_transpose(idw[ch & 1][g & 3], sel1, sel2);
#pragma unroll
              for (int b = 0; b < 4; ++b) {
                const int j = 4 * g + b;
                const int c1 = clampi(c[j], -128, 127);
                const int id = static_cast<int8_t>(iw >> (8 * b));
                // int8 codes -> fp64 through I2F.F64 (one slot on the idle conversion pipe each) instead of
                // XOR + MOV + DADD: the epilogue is issue-bound (23 instructions per output before)
                const double cd = __int2double_rn(c1), dd = __int2double_rn(id);
                c[j] = __double2loint(__fma_rn(cd, m2, kMagicRnd)) + __double2loint(__fma_rn(dd, mi, kMagicRnd));
              }
            }
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) packed[g] = pack4_sat_s8(c[4 * g], c[4 * g + 1], c[4 * g + 2], c[4 * g + 3]);
        }
        if (MODE == 4) store_chunk16(Cfg::NCH - 1, reinterpret_cast<const uint32_t(&)[8]>(packed16));
        else if (MODE != 3) store_chunk(Cfg::NCH - 1, packed);
      };
      if (all_ok) process(std::true_type{});
      else process(std::false_type{});
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ activation-stationary kernel
// QuantLinear -> QuantAct (REQUANT) for K <= 768, optionally with the GELU that follows it (mlp.fc1 ->
// qact_gelu -> ShiftGELU | I-BERT GELU -> mlp.qact1: layers_quant.py:145-154) fused into the same launch.
//
// The channels-on-lanes kernel above walks (channel block, token tile) pairs, so every 128 x 256 output tile
// pulls its own (128 + 256) x K operand bytes through L2 -> shared memory: at K = 384 that feed, not the tensor
// pipe, bounds it.  Here a CTA owns a TOKEN tile (TB tokens) for ALL channel blocks: the activation tile is
// loaded once and stays in shared memory (KB slabs of TB x 128 B), only the 128-channel weight slabs stream
// through the ring -- (TB + ncb * 128) x K bytes per token tile instead of ncb * (128 + TB) x K.  TB = 176
// gives 287 token tiles for 256 DeiT images = 1.94 rounds of 148 CTAs.
//
// Because all channels of a token pass through one CTA, the ShiftGELU row maximum (ivit_modules.py:110, over
// the whole hidden dimension) is available inside the launch: when the last channel block of a token tile has
// been stored, the epilogue warps re-read the tile's rows (L2 hits), take the row maximum, apply the layer's
// (row max, code) table -- GELU + mlp.qact1, built at freeze time -- and write the fc2 operand.  The separate
// GELU launch and its HBM round trip are gone; gelu_out may alias out (in place).
template <int TB_, int KB_>
struct AsCfg {
  static constexpr int CB = 128;
  static constexpr int TB = TB_;
  static constexpr int KB = KB_;                                 // K slabs of 128 bytes held for the token tile
  static constexpr int A_SLAB = TB * BK;
  static constexpr int W_BYTES = CB * BK;
  static constexpr int STAGES = KB <= 3 ? 8 : 4;
  static constexpr int ACC_STRIDE = 256;
  static constexpr uint32_t TMEM_COLS = 512;
  static constexpr int PART = 64;
  static constexpr int NPARTS = (TB + PART - 1) / PART;
  static constexpr int EPI_WARPS = 4 * NPARTS;
  static constexpr int THREADS = 64 + 32 * EPI_WARPS;
  static constexpr int NBAR = 2 * KB + 2 * STAGES + 4 + 1;
  static constexpr size_t SMEM = KB * A_SLAB + STAGES * W_BYTES + 8 * NBAR + 64 + EPI_WARPS * 512;
  static_assert(TB % 16 == 0 && TB <= 256 && A_SLAB % 1024 == 0, "token tile: UMMA N, whole swizzle atoms");
};

template <int TB, int KB, bool GELU>
__global__ void __launch_bounds__(AsCfg<TB, KB>::THREADS, 1)
gemm_i8_as_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const GemmArgs args) {
  using Cfg = AsCfg<TB, KB>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sA = smem;                                            // KB slabs [TB x 128 B], SWIZZLE_128B
  uint8_t* sW = smem + KB * Cfg::A_SLAB;                         // STAGES x [128 x 128 B]
  uint64_t* a_full = reinterpret_cast<uint64_t*>(sW + STAGES * Cfg::W_BYTES);
  uint64_t* a_empty = a_full + KB;
  uint64_t* w_full = a_empty + KB;
  uint64_t* w_empty = w_full + STAGES;
  uint64_t* tfull_bar = w_empty + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* tile_done = tempty_bar + 2;                          // GELU: every channel block of a token tile is stored
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tile_done + 1);
  uint8_t* srow = reinterpret_cast<uint8_t*>(tmem_slot) + 64;    // GELU: [EPI_WARPS][2][256] table lines

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int ncb = (args.N + Cfg::CB - 1) / Cfg::CB;
  const int ntt = (args.M + TB - 1) / TB;
  const int kblocks = (args.K + BK - 1) / BK;                    // <= KB (host-checked)
  const ivit_gemm_epilogue_t& e = args.e;

  if (threadIdx.x == 0) {
    for (int k = 0; k < KB; ++k) {
      mbar_init(&a_full[k], 1);
      mbar_init(&a_empty[k], 1);
    }
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], Cfg::EPI_WARPS);
    }
    mbar_init(tile_done, Cfg::EPI_WARPS);
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================================================== TMA producer
    if (elect_one()) {
      int it = 0, ti = 0;
      for (int tt = blockIdx.x; tt < ntt; tt += gridDim.x, ++ti) {
        // the token tile: slab kb is free once the LAST channel block of the previous tile has consumed it
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&a_empty[kb], (ti & 1) ^ 1);
          mbar_arrive_expect_tx(&a_full[kb], Cfg::A_SLAB);
          tma_load_2d(sA + kb * Cfg::A_SLAB, &tmA, &a_full[kb], kb * BK, tt * TB);
        }
        for (int cb = 0; cb < ncb; ++cb) {
          for (int kb = 0; kb < kblocks; ++kb, ++it) {
            const int s = it % STAGES;
            mbar_wait(&w_empty[s], ((it / STAGES) & 1) ^ 1);
            mbar_arrive_expect_tx(&w_full[s], Cfg::W_BYTES);
            tma_load_2d(sW + s * Cfg::W_BYTES, &tmW, &w_full[s], kb * BK, cb * Cfg::CB);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_i8(Cfg::CB, TB, true, true, false, false);
      int it = 0, ci = 0, ti = 0;
      for (int tt = blockIdx.x; tt < ntt; tt += gridDim.x, ++ti) {
        for (int cb = 0; cb < ncb; ++cb, ++ci) {
          const int a = ci & 1;
          mbar_wait(&tempty_bar[a], ((ci >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + a * Cfg::ACC_STRIDE;
          for (int kb = 0; kb < kblocks; ++kb, ++it) {
            const int s = it % STAGES;
            if (cb == 0) mbar_wait(&a_full[kb], ti & 1);
            mbar_wait(&w_full[s], (it / STAGES) & 1);
            tc_fence_after();
            const uint32_t sw = smem_u32(sW + s * Cfg::W_BYTES);
            const uint32_t sa = smem_u32(sA + kb * Cfg::A_SLAB);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t dw = make_smem_desc(sw + k * UMMA_K, 16, 1024, kLayoutSW128);
              const uint64_t da = make_smem_desc(sa + k * UMMA_K, 16, 1024, kLayoutSW128);
              mma_i8_ss(d_tmem, dw, da, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            tc_commit(&w_empty[s]);
            if (cb == ncb - 1) tc_commit(&a_empty[kb]);
          }
          tc_commit(&tfull_bar[a]);
        }
      }
    }
    __syncwarp();
  } else if (warp < 2 + Cfg::EPI_WARPS) {
    // ===================================================== epilogue (thread = output channel)
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int part = ew >> 2;
    const int nch = min(4, (TB - part * Cfg::PART) / 16);          // 16-token chunks of this warp's part
    const uint32_t sel1 = (lane & 1) ? 0x3715u : 0x6240u;
    const uint32_t sel2 = (lane & 2) ? 0x3276u : 0x5410u;
    const int qi = lane & 3;                 // after the transpose: this lane's token within a group of 4
    const int qc = (lane >> 2) * 4;          //                      and its first channel within the warp's 32
    int pb = 0;
    double pm = 0.0;
    auto fetch_params = [&](int cb) {
      const int n = cb * Cfg::CB + quarter * 32 + lane;
      pb = (e.bias != nullptr && n < args.N) ? __ldg(e.bias + n) : 0;
      pm = n < args.N ? __ldg(e.mult + n) : 0.0;
    };
    fetch_params(0);
    int8_t* const outp = static_cast<int8_t*>(args.out);
    const uint32_t ldo = static_cast<uint32_t>(args.ldo);
    int ci = 0;
    for (int tt = blockIdx.x; tt < ntt; tt += gridDim.x) {
      const int tok_w = tt * TB + part * Cfg::PART;              // the warp's first token
      for (int cb = 0; cb < ncb; ++cb, ++ci) {
        const int a = ci & 1;
        const int c0 = cb * Cfg::CB + quarter * 32;              // the warp's first channel
        const double mult = pm;
        const double kb = kMagicInt - static_cast<double>(pb);
        fetch_params(cb + 1 < ncb ? cb + 1 : 0);
        const int nq = c0 + qc;                                  // channels nq..nq+3 after the transpose
        const bool nq_ok = nq < args.N;                          // N % 4 == 0 (host-checked)
        const bool all_ok = tok_w + 16 * nch <= args.M && c0 + 32 <= args.N;   // warp-uniform: no guards
        const uint32_t off_o = static_cast<uint32_t>(tok_w + qi) * ldo + nq;
        const uint32_t t_row = tmem_base + a * Cfg::ACC_STRIDE + part * Cfg::PART +
                               (static_cast<uint32_t>(quarter * 32) << 16);
        auto process = [&](auto full_tag) {
          constexpr bool FULL = decltype(full_tag)::value;
          auto store_chunk = [&](int ch, const uint32_t (&packed)[4]) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint32_t w = quad_transpose(packed[g], sel1, sel2);
              if (FULL || (nq_ok && tok_w + ch * 16 + 4 * g + qi < args.M))
                *reinterpret_cast<uint32_t*>(outp + (off_o + (ch * 16 + 4 * g) * ldo)) = w;
            }
          };
          mbar_wait(&tfull_bar[a], (ci >> 1) & 1);
          tc_fence_after();
          uint32_t acc[2][16];
          uint32_t packed[4];
          tmem_ld16(t_row, acc[0]);
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            if (ch < nch) {
              tmem_ld_wait();
              if (ch + 1 < nch) {
                tmem_ld16(t_row + (ch + 1) * 16, acc[(ch + 1) & 1]);
              } else {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty_bar[a]);
              }
              const uint32_t(&z)[16] = acc[ch & 1];      // rows past M are TMA zero fill: harmless, stores are guarded
              int c[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const double zd = __hiloint2double(0x43300000, static_cast<int>(z[j] ^ 0x80000000u)) - kb;
                c[j] = __double2loint(__fma_rn(zd, mult, kMagicRnd));
              }
              if (ch > 0) store_chunk(ch - 1, packed);
#pragma unroll
              for (int g = 0; g < 4; ++g) packed[g] = pack4_sat_s8(c[4 * g], c[4 * g + 1], c[4 * g + 2], c[4 * g + 3]);
            }
          }
          store_chunk(nch - 1, packed);
        };
        if (all_ok) process(std::true_type{});
        else process(std::false_type{});
      }
      if (GELU) {
        // every channel block of this token tile is stored (by all epilogue warps): GELU + mlp.qact1 over the tile's
        // rows (L2 hits) -- row maximum (ivit_modules.py:110), then byte lookups in the row's 256-byte line of the
        // layer's (row max, code) table.  Two rows per warp-step, loads first.  (Measured: a dedicated group of four
        // tail warps is slower, 1.56 vs 1.13 ms per DeiT-S step: the pass needs many warps to hide its latencies.)
        asm volatile("bar.sync 1, %0;" ::"n"(32 * Cfg::EPI_WARPS) : "memory");
        const int nv = args.N / 16;
        uint8_t* trow0 = srow + ew * 512;
        const int rows_here = min(TB, args.M - tt * TB);
        for (int r0 = 2 * ew; r0 < rows_here; r0 += 2 * Cfg::EPI_WARPS) {
        const bool two = r0 + 1 < rows_here;
        int4 v[2][6];
        uint32_t m[2] = {0x80808080u, 0x80808080u};
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const long long ro = static_cast<long long>(tt * TB + r0 + (two ? u : 0)) * args.ldo;
          const int4* xr = reinterpret_cast<const int4*>(outp + ro);
#pragma unroll
          for (int i = 0; i < 6; ++i) {
            const int idx = i * 32 + lane;
            if (idx < nv) v[u][i] = __ldcg(xr + idx);      // written by the epilogue warps a moment ago: L2, not L1
          }
        }
        int mx[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
#pragma unroll
          for (int i = 0; i < 6; ++i) {
            if (i * 32 + lane < nv) {
              m[u] = __vmaxs4(m[u], v[u][i].x); m[u] = __vmaxs4(m[u], v[u][i].y);
              m[u] = __vmaxs4(m[u], v[u][i].z); m[u] = __vmaxs4(m[u], v[u][i].w);
            }
          }
          mx[u] = max(max(static_cast<int8_t>(m[u]), static_cast<int8_t>(m[u] >> 8)),
                      max(static_cast<int8_t>(m[u] >> 16), static_cast<int8_t>(m[u] >> 24)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          mx[0] = max(mx[0], __shfl_xor_sync(0xffffffffu, mx[0], o));
          mx[1] = max(mx[1], __shfl_xor_sync(0xffffffffu, mx[1], o));
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 2; ++u)
          *reinterpret_cast<uint2*>(trow0 + u * 256 + 8 * lane) =
              __ldg(reinterpret_cast<const uint2*>(e.gelu_lut + (mx[u] + 128) * 256) + lane);
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (u == 1 && !two) break;
          const uint8_t* trow = trow0 + u * 256;
          auto look4 = [&](uint32_t w) -> uint32_t {
            w ^= 0x80808080u;
            const uint32_t b0 = trow[w & 0xffu], b1 = trow[(w >> 8) & 0xffu], b2 = trow[(w >> 16) & 0xffu], b3 = trow[w >> 24];
            return __byte_perm(__byte_perm(b0, b1, 0x0040), __byte_perm(b2, b3, 0x0040), 0x5410);
          };
          int4* orow = reinterpret_cast<int4*>(e.gelu_out + static_cast<long long>(tt * TB + r0 + u) * args.ldo);
#pragma unroll
          for (int i = 0; i < 6; ++i) {
            const int idx = i * 32 + lane;
            if (idx < nv) {
              int4 o;
              o.x = look4(v[u][i].x); o.y = look4(v[u][i].y); o.z = look4(v[u][i].z); o.w = look4(v[u][i].w);
              orow[idx] = o;
            }
          }
        }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

// ------------------------------------------------------------------ simple cross-check kernel
__global__ void gemm_i8_simple_kernel(const int8_t* __restrict__ A, long long lda, const int8_t* __restrict__ W,
                                      const GemmArgs args) {
  const ivit_gemm_epilogue_t& e = args.e;
  const long long total = static_cast<long long>(args.M) * args.N;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int n = static_cast<int>(t % args.N), r = static_cast<int>(t / args.N);
    const int8_t* a = A + r * lda;
    const int8_t* w = W + static_cast<long long>(n) * args.K;
    int acc = 0;
    int k = 0;
    if (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(w)) & 3) == 0)
      for (; k + 4 <= args.K; k += 4)
        acc = __dp4a(*reinterpret_cast<const int*>(a + k), *reinterpret_cast<const int*>(w + k), acc);
    for (; k < args.K; ++k) acc += static_cast<int>(a[k]) * static_cast<int>(w[k]);
    const int z = acc + (e.bias ? e.bias[n] : 0);
    const long long orow = out_row_of(r, e);
    if (e.epilogue == IVIT_EPI_ACC32) {
      static_cast<int32_t*>(args.out)[orow * args.ldo + n] = z;
    } else if (e.epilogue == IVIT_EPI_LOGITS) {
      static_cast<float*>(args.out)[orow * args.ldo + n] = __fmul_rn(static_cast<float>(z), e.scale_out[n]);
    } else {
      int c = epi_requant_elem(z, n, e);
      if (e.epilogue == IVIT_EPI_RESIDUAL) {
        const long long io = id_row_of(r, e) * e.ld_id + n;
        const int id = e.id_bits == 16 ? static_cast<int>(reinterpret_cast<const int16_t*>(e.identity)[io])
                                       : static_cast<int>(e.identity[io]);
        c = epi_residual_elem(c, id, e);
      }
      if (e.out_bits == 16) static_cast<int16_t*>(args.out)[orow * args.ldo + n] = static_cast<int16_t>(c);
      else static_cast<int8_t*>(args.out)[orow * args.ldo + n] = static_cast<int8_t>(c);
    }
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The driver entry point is fetched at run time so that the library itself links only the
// CUDA runtime and loads (for symbol checks) on machines without a driver.
EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// 2-D uint8 tensor map: dim0 = contiguous bytes (cols), dim1 = rows with pitch `ld` bytes.
int make_tmap_2d_u8(CUtensorMap* m, const void* base, uint64_t cols, uint64_t rows, uint64_t ld, uint32_t box_cols,
                    uint32_t box_rows, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return fail(IVIT_ERR_DRIVER, "cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[96];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
    return fail(IVIT_ERR_DRIVER, buf);
  }
  return 0;
}

static int fill_args(GemmArgs& a, int M, int N, int K, void* out, int64_t ldo, const ivit_gemm_epilogue_t* epi) {
  if (!epi) return fail(IVIT_ERR_BAD_ARG, "gemm: epilogue is NULL");
  a.M = M; a.N = N; a.K = K; a.out = out; a.ldo = ldo; a.e = *epi;
  a.allow_fast = epi->force_slow ? 0 : 1;
  if ((epi->epilogue == IVIT_EPI_REQUANT || epi->epilogue == IVIT_EPI_RESIDUAL) && (!epi->mult || !epi->s_in))
    return fail(IVIT_ERR_BAD_ARG, "gemm: mult / s_in required for requant epilogues");
  if (epi->epilogue == IVIT_EPI_RESIDUAL && !epi->identity) return fail(IVIT_ERR_BAD_ARG, "gemm: identity is NULL");
  if (epi->epilogue == IVIT_EPI_LOGITS && !epi->scale_out) return fail(IVIT_ERR_BAD_ARG, "gemm: scale_out is NULL");
  if (epi->out_bits != 0 && epi->out_bits != 8 && epi->out_bits != 16)
    return fail(IVIT_ERR_BAD_ARG, "gemm: out_bits must be 0/8 or 16");
  if (epi->id_bits != 0 && epi->id_bits != 8 && epi->id_bits != 16)
    return fail(IVIT_ERR_BAD_ARG, "gemm: id_bits must be 0/8 or 16");
  if (epi->epilogue == IVIT_EPI_REQUANT || epi->epilogue == IVIT_EPI_RESIDUAL) {
    // the stored codes must fit the output type; RESIDUAL's first clamp (qmin, qmax) is an in-register
    // intermediate of up to 16 bits
    const int lim = epi->out_bits == 16 ? 32768 : 128;
    const bool res = epi->epilogue == IVIT_EPI_RESIDUAL;
    const int lo = res ? epi->qmin2 : epi->qmin, hi = res ? epi->qmax2 : epi->qmax;
    if (lo < -lim || hi > lim - 1) return fail(IVIT_ERR_UNSUPPORTED, "gemm: output clamp range exceeds the output type");
    if (res && (epi->qmin < -32768 || epi->qmax > 32767))
      return fail(IVIT_ERR_UNSUPPORTED, "gemm: intermediate clamp range exceeds 16 bits");
  }
  return 0;
}

}  // namespace ivit

using namespace ivit;

constexpr int kMaxDevices = 64;
static int current_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = 0;
  return dev;
}

static int sm_count() {
  static int n_dev[kMaxDevices] = {};
  int& n = n_dev[current_device_slot()];
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

template <int BN, int MODE, bool CHECKED>
static int launch_ws(const int8_t* A, int64_t lda, const int8_t* W, GemmArgs& args, cudaStream_t st) {
  using Cfg = WsCfg<BN>;
  CUtensorMap tmA, tmW;
  int rc = make_tmap_2d_u8(&tmA, A, args.K, args.M, lda, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d_u8(&tmW, W, args.K, args.N, args.K, BK, BN, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  args.stages = Cfg::STAGES;
  static bool attr_set_dev[kMaxDevices] = {};      // the attribute is per-device state
  bool& attr_set = attr_set_dev[current_device_slot()];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_i8_ws_kernel<BN, MODE, CHECKED>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(Cfg::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    attr_set = true;
  }
  const int ntiles = ((args.M + BM - 1) / BM) * ((args.N + BN - 1) / BN);
  const int grid = ntiles < sm_count() ? ntiles : sm_count();
  gemm_i8_ws_kernel<BN, MODE, CHECKED><<<grid, Cfg::THREADS, Cfg::SMEM, st>>>(tmA, tmW, args);
  return check_launch("gemm_i8_ws_kernel");
}

template <int TB, int MODE, bool CHECKED>
static int launch_cl(const int8_t* A, int64_t lda, const int8_t* W, GemmArgs& args, cudaStream_t st) {
  using Cfg = ClCfg<TB>;
  CUtensorMap tmA, tmW;
  int rc = make_tmap_2d_u8(&tmA, A, args.K, args.M, lda, BK, Cfg::TB, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d_u8(&tmW, W, args.K, args.N, args.K, BK, Cfg::CB, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  args.stages = Cfg::STAGES;
  static bool attr_set_dev[kMaxDevices] = {};      // the attribute is per-device state
  bool& attr_set = attr_set_dev[current_device_slot()];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_i8_cl_kernel<TB, MODE, CHECKED>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(Cfg::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    attr_set = true;
  }
  const int ntiles = ((args.M + Cfg::TB - 1) / Cfg::TB) * ((args.N + Cfg::CB - 1) / Cfg::CB);
  const int grid = ntiles < sm_count() ? ntiles : sm_count();
  cudaError_t le = launch_kernel(gemm_i8_cl_kernel<TB, MODE, CHECKED>, dim3(grid), dim3(Cfg::THREADS), Cfg::SMEM, st, tmA,
                                 tmW, args);
  if (le != cudaSuccess) return fail(static_cast<int>(le), cudaGetErrorString(le));
  return check_launch("gemm_i8_cl_kernel");
}

template <int TB, int KB, bool GELU>
static int launch_as(const int8_t* A, int64_t lda, const int8_t* W, GemmArgs& args, cudaStream_t st) {
  using Cfg = AsCfg<TB, KB>;
  CUtensorMap tmA, tmW;
  int rc = make_tmap_2d_u8(&tmA, A, args.K, args.M, lda, BK, TB, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d_u8(&tmW, W, args.K, args.N, args.K, BK, Cfg::CB, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  args.stages = Cfg::STAGES;
  static bool attr_set_dev[kMaxDevices] = {};
  bool& attr_set = attr_set_dev[current_device_slot()];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_i8_as_kernel<TB, KB, GELU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(Cfg::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    attr_set = true;
  }
  const int ntt = (args.M + TB - 1) / TB;
  const int grid = ntt < sm_count() ? ntt : sm_count();
  cudaError_t le = launch_kernel(gemm_i8_as_kernel<TB, KB, GELU>, dim3(grid),
                                 dim3(Cfg::THREADS), Cfg::SMEM, st, tmA, tmW, args);
  if (le != cudaSuccess) return fail(static_cast<int>(le), cudaGetErrorString(le));
  return check_launch("gemm_i8_as_kernel");
}

// 0: pick per call (default); 1: keep every GEMM on the row-major-epilogue kernel (A/B timing, tests)
static int g_gemm_layout = 0;
extern "C" int ivit_gemm_set_layout(int layout) {
  const int prev = g_gemm_layout;
  g_gemm_layout = layout;
  return prev;
}
static bool channels_on_lanes_enabled() { return g_gemm_layout != 1; }
static bool activation_stationary_enabled() { return g_gemm_layout == 0 || g_gemm_layout == 3; }   // 3: also without GELU

template <int BN>
static int launch_mode(int mode, bool checked, const int8_t* A, int64_t lda, const int8_t* W, GemmArgs& args,
                       cudaStream_t st) {
  if (mode == 1) return checked ? launch_ws<BN, 1, true>(A, lda, W, args, st) : launch_ws<BN, 1, false>(A, lda, W, args, st);
  if (mode == 2) return checked ? launch_ws<BN, 2, true>(A, lda, W, args, st) : launch_ws<BN, 2, false>(A, lda, W, args, st);
  return launch_ws<BN, 0, true>(A, lda, W, args, st);
}

extern "C" int ivit_gemm_i8(const int8_t* A, int64_t lda, const int8_t* W, int M, int N, int K, void* out,
                            int64_t ldo, const ivit_gemm_epilogue_t* epi, void* stream) {
  if (M <= 0 || N <= 0) return 0;
  GemmArgs args;
  int rc = fill_args(args, M, N, K, out, ldo, epi);
  if (rc) return rc;
  if (K % 16 != 0 || lda % 16 != 0 || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(W) & 15))
    return fail(IVIT_ERR_UNSUPPORTED, "gemm: K, lda multiples of 16 and 16-byte aligned operands required (TMA)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // Streamlined epilogue: int8 saturating clamps, 16-byte aligned output / identity rows, |mult| small
  // enough for the one-DFMA requant; |acc + bias| < 2^22 either proven by the caller (z_bound_ok:
  // sum_k |W[n,k]| * 128 + |bias| bound) or checked per chunk with the general path behind it.
  int mode = 0;
  const bool req = epi->epilogue == IVIT_EPI_REQUANT, res = epi->epilogue == IVIT_EPI_RESIDUAL;
  const bool wide = epi->out_bits == 16 || (res && epi->id_bits == 16);
  if (!wide && (req || res) && args.allow_fast && epi->qmin == -128 && epi->qmax == 127 && ldo % 16 == 0 &&
      (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    if (req) mode = 1;
    else if (epi->qmin2 == -128 && epi->qmax2 == 127 && epi->ld_id % 16 == 0 &&
             (reinterpret_cast<uintptr_t>(epi->identity) & 15) == 0 && fabs(epi->mult2) < 65536.0 &&
             fabs(epi->mult_id) < 65536.0)
      mode = 2;
  }
  // LOGITS without a row remap: channels-on-lanes too (coalesced float rows; the tokens-on-lanes
  // epilogue scatters one 4-byte store per lane per row)
  if (epi->epilogue == IVIT_EPI_LOGITS && epi->group <= 0 && channels_on_lanes_enabled() &&
      static_cast<int64_t>(M) * ldo + N < (1ll << 31))
    return launch_cl<192, 3, false>(A, lda, W, args, st);
  const bool checked = epi->z_bound_ok == 0;
  // RESIDUAL on the int16 residual stream (int16 identity in, int16 codes out; any clamp ranges): the
  // channels-on-lanes kernel with 16-bit lane-pair transposes
  if (res && wide && epi->out_bits == 16 && epi->id_bits == 16 && args.allow_fast && epi->group <= 0 && N % 4 == 0 &&
      ldo % 2 == 0 && epi->ld_id % 2 == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0 &&
      (reinterpret_cast<uintptr_t>(epi->identity) & 3) == 0 && fabs(epi->mult2) < 65536.0 &&
      fabs(epi->mult_id) < 65536.0 && static_cast<int64_t>(M) * ldo + N < (1ll << 31) &&
      static_cast<int64_t>(M) * epi->ld_id + N < (1ll << 31) && channels_on_lanes_enabled())
    return checked ? launch_cl<192, 4, true>(A, lda, W, args, st) : launch_cl<192, 4, false>(A, lda, W, args, st);
  // REQUANT with K <= 768 and a proven accumulator bound: the activation-stationary kernel (and the fused GELU)
  const bool gelu = req && epi->gelu_lut != nullptr;
  if (gelu && (epi->gelu_out == nullptr || N % 16 != 0))
    return fail(IVIT_ERR_BAD_ARG, "gemm: fused GELU needs gelu_out and N % 16 == 0");
  {
    const int64_t span = static_cast<int64_t>(M) * ldo + N;
    if (mode == 1 && !checked && epi->group <= 0 && N % 4 == 0 && ldo % 16 == 0 && span < (1ll << 31) && K <= 768 &&
        (gelu || g_gemm_layout == 3) && activation_stationary_enabled() && (!gelu || (N <= 3072 && (reinterpret_cast<uintptr_t>(epi->gelu_out) & 15) == 0))) {
      if (K <= 384) return gelu ? launch_as<176, 3, true>(A, lda, W, args, st) : launch_as<176, 3, false>(A, lda, W, args, st);
      return gelu ? launch_as<176, 6, true>(A, lda, W, args, st) : launch_as<176, 6, false>(A, lda, W, args, st);
    }
  }
  if (gelu) {
    // not eligible: the plain GEMM, then the table pass (same results, two launches)
    ivit_gemm_epilogue_t plain = *epi;
    plain.gelu_lut = nullptr;
    plain.gelu_out = nullptr;
    rc = ivit_gemm_i8(A, lda, W, M, N, K, out, ldo, &plain, stream);
    if (rc) return rc;
    if (ldo != N) return fail(IVIT_ERR_UNSUPPORTED, "gemm: the two-launch GELU fallback needs contiguous rows");
    return ivit_gelu_lut_apply_i8(static_cast<const int8_t*>(out), epi->gelu_out, epi->gelu_lut, M, N, stream);
  }
  // Hot shapes: channels-on-lanes kernel (per-channel constants in registers).  Needs 4-byte
  // aligned output / identity rows and no row remap.
  const int64_t span_o = static_cast<int64_t>(M) * ldo + N, span_i = static_cast<int64_t>(M) * epi->ld_id + N;
  if (mode != 0 && epi->group <= 0 && N % 4 == 0 && ldo % 4 == 0 && span_o < (1ll << 31) &&
      (mode == 1 || (epi->ld_id % 4 == 0 && span_i < (1ll << 31))) && channels_on_lanes_enabled()) {
    // REQUANT: 256-token tiles, 16 epilogue warps (fits 96 registers); RESIDUAL: 192 tokens, 12 warps
    if (mode == 1) return checked ? launch_cl<256, 1, true>(A, lda, W, args, st) : launch_cl<256, 1, false>(A, lda, W, args, st);
    return checked ? launch_cl<192, 2, true>(A, lda, W, args, st) : launch_cl<192, 2, false>(A, lda, W, args, st);
  }
  // BN = 192 divides every DeiT/ViT width (3C, C, 4C for C = 192/384/768); 128 otherwise
  if (N % 192 == 0 || N > 512) return launch_mode<192>(mode, checked, A, lda, W, args, st);
  return launch_mode<128>(mode, checked, A, lda, W, args, st);
}

extern "C" int ivit_gemm_i8_simple(const int8_t* A, int64_t lda, const int8_t* W, int M, int N, int K, void* out,
                                   int64_t ldo, const ivit_gemm_epilogue_t* epi, void* stream) {
  if (M <= 0 || N <= 0) return 0;
  GemmArgs args;
  int rc = fill_args(args, M, N, K, out, ldo, epi);
  if (rc) return rc;
  args.stages = 0;
  const long long total = static_cast<long long>(M) * N;
  long long g = (total + 255) / 256;
  if (g > 148 * 32) g = 148 * 32;
  gemm_i8_simple_kernel<<<static_cast<int>(g), 256, 0, static_cast<cudaStream_t>(stream)>>>(A, lda, W, args);
  return check_launch("gemm_i8_simple_kernel");
}
