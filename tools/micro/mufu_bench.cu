// Developer microbenchmark: how fast can ONE warp per SM sub-partition issue MUFU.EX2, alone and
// interleaved with FMA-pipe work, and what do two warps per sub-partition reach?  (Informs the
// softmax design of the dense attention kernel: profiles/r01_sm100_v3_experiments.md.)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mufu_bench tools/micro/mufu_bench.cu && ./mufu_bench
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int kFma>   // FFMAs issued between consecutive MUFUs
__global__ void k(float* out, long long* clk, int iters) {
  float v[16], f[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = -0.001f * (threadIdx.x + i);
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = 1.0f + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      v[i] = ex2(v[i]);                       // 16 independent chains
#pragma unroll
      for (int j = 0; j < kFma; ++j) f[j & 7] = fmaf(f[j & 7], 0.999f, 0.001f);
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int kFma>
void run(int threads, float* out, long long* clk) {
  const int iters = 2000;
  k<kFma><<<148, threads>>>(out, clk, iters);
  cudaDeviceSynchronize();
  k<kFma><<<148, threads>>>(out, clk, iters);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, clk, sizeof(c), cudaMemcpyDeviceToHost);
  printf("warps/SMSP %d, %d FFMA between MUFUs: %.2f clk per MUFU per warp (%.2f MUFU lanes/clk/SM)\n", threads / 128,
         kFma, (double)c / (iters * 16.0), 32.0 * (threads / 32) / ((double)c / (iters * 16.0)));
}

int main() {
  float* out;
  long long* clk;
  cudaMalloc(&out, 148 * 1024 * sizeof(float));
  cudaMalloc(&clk, sizeof(long long));
  for (int threads : {128, 256, 512}) {
    run<0>(threads, out, clk);
    run<2>(threads, out, clk);
    run<4>(threads, out, clk);
    run<6>(threads, out, clk);
    run<8>(threads, out, clk);
    run<12>(threads, out, clk);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
