// Micro-benchmark: legacy mma.sync m16n8k8 tf32 and packed fma.rn.f32x2 issue rates on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mma_probe tools/mma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_mma(float* out, int iters) {
  float c[6][4] = {};
  unsigned a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 7, a3 = 9, b0 = 5, b1 = blockIdx.x;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int t = 0; t < 6; ++t)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int t = 0; t < 6; ++t) for (int j = 0; j < 4; ++j) s += c[t][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma(float* out, int iters) {
  float c[24];
  for (int j = 0; j < 24; ++j) c[j] = j;
  float a = threadIdx.x * 1e-9f, b = 0.999f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 24; ++j) c[j] = fmaf(c[j], b, a);
  }
  float s = 0;
  for (int j = 0; j < 24; ++j) s += c[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float* out, int iters) {
  unsigned long long c[12];
  for (int j = 0; j < 12; ++j) c[j] = j;
  unsigned long long a = threadIdx.x, b = 0x3f7fbe773f7fbe77ull;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 12; ++j) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(c[j]) : "l"(b), "l"(a));
  }
  unsigned long long s = 0;
  for (int j = 0; j < 12; ++j) s += c[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = (float)s;
}

template <class F>
float time_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  float* out; cudaMalloc(&out, 148 * 1024 * 4 * 8);
  int dev_clock; cudaDeviceGetAttribute(&dev_clock, cudaDevAttrClockRate, 0);
  const int iters = 20000;
  for (int warps : {1, 2, 4, 8, 12}) {
    const int threads = 32 * warps;
    float ms = time_ms([&] { k_mma<<<148, threads>>>(out, iters); });
    double mmas = (double)iters * 6 * warps;  // per SM
    printf("mma.sync tf32 m16n8k8: %2d warps/SM  %.3f ms  -> %.2f MMA/us/SM, %.1f clk/MMA/SM @%.0f MHz (dense TF %.1f)\n", warps, ms,
           mmas / (ms * 1e3), ms * 1e-3 * dev_clock * 1e3 / mmas, dev_clock / 1e3, mmas * 148 * 2048 / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_ffma<<<148, threads>>>(out, iters); });
    printf("FFMA:   %2d warps/SM %.3f ms -> %.2f warp-FFMA/clk/SM\n", warps, ms, (double)iters * 24 * warps / (ms * 1e-3 * dev_clock * 1e3));
    ms = time_ms([&] { k_ffma2<<<148, threads>>>(out, iters); });
    printf("FFMA2:  %2d warps/SM %.3f ms -> %.2f warp-FFMA2/clk/SM\n", warps, ms, (double)iters * 12 * warps / (ms * 1e-3 * dev_clock * 1e3));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
