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
#include <cstdio>

#include "intmath.cuh"
#include "ivit_b200.h"
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
//   warps 2..9  epilogue: warp w reads TMEM lanes 32*(w%4).. (thread = output row), the two warps
//               of a lane quarter split the BN columns; requant / residual / logits, 32-byte stores
constexpr int WS_THREADS = 320;
constexpr int EPI_WARPS = 8;

template <int BN>
struct WsCfg {
  static constexpr int A_BYTES = BM * BK;
  static constexpr int B_BYTES = BN * BK;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = BN <= 128 ? 6 : (BN <= 192 ? 5 : 4);
  static constexpr int ACC_STRIDE = BN <= 128 ? 128 : 256;       // TMEM columns between the two accumulators
  static constexpr uint32_t TMEM_COLS = 2 * ACC_STRIDE;
  static constexpr int HALF = BN / 2;                            // columns per epilogue warp
  static constexpr int PAR_BYTES = EPI_WARPS * HALF * 12;        // per-warp bias (int32) + mult (double)
  static constexpr size_t SMEM = 1024 + STAGES * STAGE_BYTES + 256 + PAR_BYTES;
};

// One 32-column chunk of one output row (thread = row): z = acc + bias, then the epilogue.
template <bool CHECKED>
__device__ __forceinline__ void epilogue_chunk(const uint32_t (&acc)[32], const int32_t* __restrict__ s_bias,
                                               const double* __restrict__ s_mult, int nb, bool row_ok,
                                               long long orow, long long irow, const GemmArgs& args) {
  const ivit_gemm_epilogue_t& e = args.e;
  const bool full_chunk = nb + 32 <= args.N;
  int z[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) z[j] = static_cast<int>(acc[j]) + s_bias[j];

  if (e.epilogue == IVIT_EPI_ACC32) {
    if (row_ok) {
      int32_t* o = static_cast<int32_t*>(args.out) + orow * args.ldo + nb;
      if (full_chunk) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<int4*>(o + j) = make_int4(z[j], z[j + 1], z[j + 2], z[j + 3]);
      } else {
        for (int j = 0; j < 32 && nb + j < args.N; ++j) o[j] = z[j];
      }
    }
    return;
  }
  if (e.epilogue == IVIT_EPI_LOGITS) {
    if (row_ok) {
      float* o = static_cast<float*>(args.out) + orow * args.ldo + nb;
      for (int j = 0; j < 32 && nb + j < args.N; ++j)
        o[j] = __fmul_rn(static_cast<float>(z[j]), __ldg(e.scale_out + nb + j));
    }
    return;
  }

  // REQUANT / RESIDUAL: first fixedpoint_mul (per-channel)
  bool ok = args.allow_fast != 0;
  if (CHECKED) {
    uint32_t bits = 0;                 // OR of |z| (one's complement for negatives): < 2^22 <=> all in range
#pragma unroll
    for (int j = 0; j < 32; ++j) bits |= static_cast<uint32_t>(z[j] ^ (z[j] >> 31));
    ok = ok && bits < (1u << 22);
  }
  int c[32];
  if (__all_sync(0xffffffffu, ok)) {
#pragma unroll
    for (int j = 0; j < 32; ++j) c[j] = clampi(requant_i32_fast(z[j], s_mult[j]), e.qmin, e.qmax);
  } else {
    for (int j = 0; j < 32; ++j) c[j] = (nb + j < args.N) ? epi_requant_elem(z[j], nb + j, e) : 0;
  }
  if (e.epilogue == IVIT_EPI_RESIDUAL) {
    uint32_t idw[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (row_ok) {
      const int8_t* ip = e.identity + irow * e.ld_id + nb;
      if (full_chunk) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(ip));
        const int4 b = __ldg(reinterpret_cast<const int4*>(ip + 16));
        idw[0] = a.x; idw[1] = a.y; idw[2] = a.z; idw[3] = a.w;
        idw[4] = b.x; idw[5] = b.y; idw[6] = b.z; idw[7] = b.w;
      } else {
        for (int j = 0; j < 32 && nb + j < args.N; ++j)
          idw[j >> 2] |= static_cast<uint32_t>(static_cast<uint8_t>(ip[j])) << (8 * (j & 3));
      }
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int id = static_cast<int8_t>(idw[j >> 2] >> (8 * (j & 3)));
      // second fixedpoint_mul (scalar ratio) + identity branch, rounded separately then added
      if (args.allow_fast) {
        const int a = requant_i32_fast(c[j], e.mult2);
        const int b = requant_i32_fast(id, e.mult_id);
        c[j] = clampi(a + b, e.qmin2, e.qmax2);
      } else {
        c[j] = epi_residual_elem(c[j], id, e);
      }
    }
  }
  if (row_ok) {
    int8_t* o = static_cast<int8_t*>(args.out) + orow * args.ldo + nb;
    if (full_chunk) {
      uint32_t w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        w[j] = (c[4 * j] & 0xff) | ((c[4 * j + 1] & 0xff) << 8) | ((c[4 * j + 2] & 0xff) << 16) |
               (static_cast<uint32_t>(c[4 * j + 3]) << 24);
      *reinterpret_cast<int4*>(o) = make_int4(w[0], w[1], w[2], w[3]);
      *reinterpret_cast<int4*>(o + 16) = make_int4(w[4], w[5], w[6], w[7]);
    } else {
      for (int j = 0; j < 32 && nb + j < args.N; ++j) o[j] = static_cast<int8_t>(c[j]);
    }
  }
}

template <int BN, bool CHECKED>
__global__ void __launch_bounds__(WS_THREADS, 1)
gemm_i8_ws_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const GemmArgs args) {
  using Cfg = WsCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
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
      mbar_init(&tempty_bar[a], EPI_WARPS);
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
    // ===================================================== epilogue (8 warps)
    const int ew = warp - 2;
    const int quarter = warp & 3;            // TMEM lanes this warp may read
    const int half = ew >> 2;                // which BN/2 columns
    int32_t* s_bias = reinterpret_cast<int32_t*>(par + ew * Cfg::HALF * 12);
    double* s_mult = reinterpret_cast<double*>(par + ew * Cfg::HALF * 12 + Cfg::HALF * 4);
    // keep the doubles 8-byte aligned: HALF*4 is a multiple of 8 for every BN used (64, 96, 128)
    int i = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++i) {
      const int a = i & 1;
      const int m0 = (t / ntn) * BM, n0 = (t % ntn) * BN + half * Cfg::HALF;
      for (int j = lane; j < Cfg::HALF; j += 32) {
        const int n = n0 + j;
        s_bias[j] = (e.bias != nullptr && n < args.N) ? __ldg(e.bias + n) : 0;
        s_mult[j] = (e.mult != nullptr && n < args.N) ? __ldg(e.mult + n) : 0.0;
      }
      __syncwarp();
      const int r = m0 + quarter * 32 + lane;
      const bool row_ok = r < args.M;
      const long long orow = row_ok ? out_row_of(r, e) : 0;
      const long long irow = row_ok ? id_row_of(r, e) : 0;
      mbar_wait(&tfull_bar[a], (i >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + a * Cfg::ACC_STRIDE + half * Cfg::HALF +
                             (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int ch = 0; ch < Cfg::HALF / 32; ++ch) {
        const int nb = n0 + ch * 32;
        if (nb >= args.N) break;             // warp-uniform
        uint32_t acc[32];
        tmem_ld32(t_row + ch * 32, acc);
        tmem_ld_wait();
        epilogue_chunk<CHECKED>(acc, s_bias + ch * 32, s_mult + ch * 32, nb, row_ok, orow, irow, args);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[a]);
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
      if (e.epilogue == IVIT_EPI_RESIDUAL) c = epi_residual_elem(c, e.identity[id_row_of(r, e) * e.ld_id + n], e);
      static_cast<int8_t*>(args.out)[orow * args.ldo + n] = static_cast<int8_t>(c);
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
  if (epi->epilogue == IVIT_EPI_REQUANT || epi->epilogue == IVIT_EPI_RESIDUAL) {
    if (epi->qmin < -128 || epi->qmax > 127) return fail(IVIT_ERR_UNSUPPORTED, "gemm: int8 outputs only");
    if (epi->epilogue == IVIT_EPI_RESIDUAL && (epi->qmin2 < -128 || epi->qmax2 > 127))
      return fail(IVIT_ERR_UNSUPPORTED, "gemm: int8 outputs only");
  }
  return 0;
}

}  // namespace ivit

using namespace ivit;

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

template <int BN, bool CHECKED>
static int launch_ws(const int8_t* A, int64_t lda, const int8_t* W, GemmArgs& args, cudaStream_t st) {
  using Cfg = WsCfg<BN>;
  CUtensorMap tmA, tmW;
  int rc = make_tmap_2d_u8(&tmA, A, args.K, args.M, lda, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_2d_u8(&tmW, W, args.K, args.N, args.K, BK, BN, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  args.stages = Cfg::STAGES;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_i8_ws_kernel<BN, CHECKED>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(Cfg::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), cudaGetErrorString(e));
    attr_set = true;
  }
  const int ntiles = ((args.M + BM - 1) / BM) * ((args.N + BN - 1) / BN);
  const int grid = ntiles < sm_count() ? ntiles : sm_count();
  gemm_i8_ws_kernel<BN, CHECKED><<<grid, WS_THREADS, Cfg::SMEM, st>>>(tmA, tmW, args);
  return check_launch("gemm_i8_ws_kernel");
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
  // |acc + bias| < 2^22 guaranteed by the caller (sum_k |W[n,k]| * 128 + |bias| bound): skip the range check
  const bool unchecked = epi->z_bound_ok != 0 && args.allow_fast;
  // BN = 192 divides every DeiT/ViT width (3C, C, 4C for C = 192/384/768); 128 otherwise
  if (N % 192 == 0 || N > 512)
    return unchecked ? launch_ws<192, false>(A, lda, W, args, st) : launch_ws<192, true>(A, lda, W, args, st);
  return unchecked ? launch_ws<128, false>(A, lda, W, args, st) : launch_ws<128, true>(A, lda, W, args, st);
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
