// INT8 tensor peak of this B200: back-to-back tcgen05.mma.kind::i8 from resident shared-memory operands.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ivit_b200/csrc -I include \
//        tools/micro/i8_mma_peak.cu -o tools/micro/i8_mma_peak && tools/micro/i8_mma_peak [ms_target]
//
// One CTA per SM (cta_group::1: 128x256x32 per instruction) or one CTA pair per TPC (cta_group::2:
// 256x256x32 per instruction, issued by the leader), operands in SWIZZLE_128B tiles written once (random
// bytes: data toggling costs power like real codes do), a single thread issuing K = 128 (4 instructions)
// into alternating TMEM accumulators.  Nothing is loaded or stored in the loop: this is the issue-rate
// ceiling of the tensor pipe, the denominator SURVEY section 6 asks for.  Prints one JSON line per variant with
// TOP/s, the SM clock measured inside the kernel (clock64 / globaltimer) and the run time.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#include "ptx.cuh"

using namespace ivit;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mma_i8_ss_2cta(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct Result {
  unsigned long long clk, ns;
};

template <int CTAS>   // 1: cta_group::1, 2: cta_group::2 (cluster of two)
__global__ void __launch_bounds__(128, 1) peak_kernel(int iters, Result* res) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;               // 128 rows x 128 B (K = 128), SW128: 8-row groups 1024 B apart
  uint8_t* sB = smem + 16384;       // cta_group::1: 256 rows x 128 B; cta_group::2: this CTA's 128 rows of N
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int nB = (CTAS == 1 ? 256 : 128) * 128;
  for (int i = threadIdx.x; i < (16384 + nB) / 4; i += blockDim.x) {
    uint32_t h = (i + 1) * 2654435761u + blockIdx.x * 40503u;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    reinterpret_cast<uint32_t*>(smem)[i] = h;
  }
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    if (CTAS == 1) {
      tmem_alloc(&slot, 512);
      tmem_relinquish();
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  if (CTAS == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  const bool leader = CTAS == 1 || cluster_ctarank() == 0;
  if (threadIdx.x == 0 && leader) {
    const uint32_t idesc = make_idesc_i8(CTAS == 1 ? 128 : 256, 256, true, true, false, false);
    const unsigned long long c0 = clock64(), t0 = global_ns();
    for (int it = 0; it < iters; ++it) {
      const uint32_t d = tmem + (it & 1) * 256;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = make_smem_desc(smem_u32(sA) + k * 32, 16, 1024, kLayoutSW128);
        const uint64_t db = make_smem_desc(smem_u32(sB) + k * 32, 16, 1024, kLayoutSW128);
        if (CTAS == 1) mma_i8_ss(d, da, db, idesc, k != 0 ? 1u : 0u);
        else mma_i8_ss_2cta(d, da, db, idesc, k != 0 ? 1u : 0u);
      }
    }
    if (CTAS == 1) tc_commit(&bar);
    else asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    mbar_wait(&bar, 0);
    const unsigned long long c1 = clock64(), t1 = global_ns();
    if (blockIdx.x == 0) {
      res->clk = c1 - c0;
      res->ns = t1 - t0;
    }
  }
  tc_fence_before();
  if (CTAS == 2) cluster_sync_all(); else __syncthreads();
  if (threadIdx.x < 32) {
    if (CTAS == 1) tmem_dealloc(tmem, 512);
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

template <int CTAS>
static double run(int sms, int iters, Result* dres, double* ms_out, double* mhz) {
  const size_t smem = 1024 + 16384 + 32768;
  cudaFuncSetAttribute(peak_kernel<CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sms);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CTAS;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, peak_kernel<CTAS>, iters / 8, dres);   // warm-up
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  cudaError_t le = cudaLaunchKernelEx(&cfg, peak_kernel<CTAS>, iters, dres);
  cudaEventRecord(e1);
  cudaError_t se = cudaDeviceSynchronize();
  if (le != cudaSuccess || se != cudaSuccess) {
    fprintf(stderr, "cta_group::%d failed: %s / %s\n", CTAS, cudaGetErrorString(le), cudaGetErrorString(se));
    return -1.0;
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  Result h;
  cudaMemcpy(&h, dres, sizeof(h), cudaMemcpyDeviceToHost);
  *ms_out = ms;
  *mhz = h.ns ? 1e3 * static_cast<double>(h.clk) / static_cast<double>(h.ns) : 0.0;
  // every instruction: 2 * M * 256 * 32 ops, M = 128 * CTAS; `sms / CTAS` issuers
  const double ops = static_cast<double>(sms / CTAS) * iters * 4.0 * 2.0 * (128.0 * CTAS) * 256.0 * 32.0;
  return ops / (ms * 1e-3) / 1e12;
}

int main(int argc, char** argv) {
  const double target_ms = argc > 1 ? atof(argv[1]) : 20.0;
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  Result* dres;
  cudaMalloc(&dres, sizeof(Result));
  // size the loop from a first guess of 4096 int8 MACs per clock per SM at 1.9 GHz
  int iters = static_cast<int>(target_ms * 1e-3 * 1.9e9 / 1024.0);
  for (int rep = 0; rep < 3; ++rep) {
    double ms, mhz;
    double t1 = run<1>(sms, iters, dres, &ms, &mhz);
    printf("{\"variant\": \"cta_group::1 128x256x32\", \"TOPs\": %.1f, \"ms\": %.3f, \"sm_mhz_in_kernel\": %.0f, \"sms\": %d, "
           "\"ops_per_clk_per_sm\": %.0f}\n", t1, ms, mhz, sms, t1 * 1e12 / (sms * mhz * 1e6));
    double t2 = run<2>(sms & ~1, iters, dres, &ms, &mhz);
    printf("{\"variant\": \"cta_group::2 256x256x32\", \"TOPs\": %.1f, \"ms\": %.3f, \"sm_mhz_in_kernel\": %.0f, \"sms\": %d, "
           "\"ops_per_clk_per_sm\": %.0f}\n", t2, ms, mhz, sms & ~1, t2 * 1e12 / ((sms & ~1) * mhz * 1e6));
  }
  return 0;
}
