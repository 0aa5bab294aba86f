// Micro-benchmark: throughput of ex2.approx.ftz.f32 vs ex2.approx.f16x2 vs ex2.approx.ftz.bf16x2 per SM sub-partition,
// and of an FMA-pipe degree-3 polynomial 2^x.  One block of 128 threads per SM (one warp per scheduler), 8 independent
// chains per thread.  Prints cycles per warp-instruction.   nvcc -arch=sm_100a -O3 -o mufu mufu.cu
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_bf16.h>

template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  float a[8];
  unsigned h[8];
  for (int i = 0; i < 8; ++i) { a[i] = -0.001f * (threadIdx.x + i); h[i] = 0xb800b400u + threadIdx.x + i; }
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
      if (MODE == 3) {   // Cody-Waite + degree-3 polynomial on the FMA pipe
        float x = fmaxf(a[i], -126.f);
        float y = x + 12582912.f;
        float f = x - (y - 12582912.f);
        float p = fmaf(f, 0.0555041f, 0.2402265f);
        p = fmaf(p, f, 0.6931472f);
        p = fmaf(p, f, 1.0f);
        a[i] = __int_as_float(__float_as_int(p) + (__float_as_int(y) << 23)) - 1.0f;
      }
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(h[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  const char* names[4] = {"ex2.f32", "ex2.f16x2", "ex2.bf16x2", "poly3 f32 (FMA pipe)"};
  for (int warps = 1; warps <= 4; warps *= 2) {
    for (int m = 0; m < 4; ++m) {
      for (int rep = 0; rep < 2; ++rep) {
        if (m == 0) k<0><<<148, 128 * warps>>>(out, iters, cyc);
        if (m == 1) k<1><<<148, 128 * warps>>>(out, iters, cyc);
        if (m == 2) k<2><<<148, 128 * warps>>>(out, iters, cyc);
        if (m == 3) k<3><<<148, 128 * warps>>>(out, iters, cyc);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      printf("%d warp(s)/scheduler  %-22s %.2f cycles per warp-instruction-group (%.2f per scheduler slot)\n", warps,
             names[m], (double)h / (iters * 8.0), (double)h / (iters * 8.0 * warps));
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
