// FP64 issue-rate probe for sm_100a: independent DFMA / DADD chains per thread, W warps per SM.
// Prints warp-instructions per clock per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dp_bench dp_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int KIND>
__global__ void probe(double* out, long long* clk, int iters, double m) {
  double a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
  int x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (KIND == 0) a[i] = __fma_rn(a[i], m, 1.5);
      if (KIND == 1) a[i] = __dadd_rn(a[i], m);
      if (KIND == 2) { a[i] = __dadd_rn(a[i], m); a[i] = __fma_rn(a[i], m, 1.5); }
      if (KIND == 3) { a[i] = __fma_rn(a[i], m, 1.5); x[i] = (x[i] ^ 0x5a5a) + it; x[i] = x[i] * 3 + i; x[i] ^= x[i] >> 3; }
      if (KIND == 4) { x[i] = (x[i] ^ 0x5a5a) + it; x[i] = x[i] * 3 + i; x[i] ^= x[i] >> 3; }
    }
  }
  long long t1 = clock64();
  double s = 0; int xs = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { s += a[i]; xs += x[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + xs;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
template <int KIND>
void run(const char* name, int warps, int dp_per_iter, int other_per_iter) {
  double* out; long long* clk;
  cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&clk, 148 * 8);
  const int iters = 4096;
  probe<KIND><<<148, warps * 32>>>(out, clk, iters, 1.0000001);
  probe<KIND><<<148, warps * 32>>>(out, clk, iters, 1.0000001);
  cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  double c = 0; for (int i = 0; i < 148; ++i) c += h[i]; c /= 148;
  printf("%-22s warps/SM=%2d  cycles=%9.0f  DP warp-inst/clk/SM=%.3f  other/clk/SM=%.3f\n", name, warps, c,
         (double)warps * iters * 8 * dp_per_iter / c, (double)warps * iters * 8 * other_per_iter / c);
  cudaFree(out); cudaFree(clk);
}
int main() {
  for (int w : {4, 8, 16, 32}) {
    run<0>("DFMA", w, 1, 0);
    run<1>("DADD", w, 1, 0);
    run<2>("DADD+DFMA dependent", w, 2, 0);
    run<3>("DFMA + ~6 int", w, 1, 6);
    run<4>("~6 int only", w, 0, 6);
  }
  return 0;
}
