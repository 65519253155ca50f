// Cost of a warp-wide LDS.128 when lanes read duplicate 16-byte chunks (the dC pair sweep of
// sankoff_soft_bwd_uniform.cu reads 8 distinct chunks per instruction).  Prints cycles per LDS.128 per SM for
// 32 / 8 / 1 distinct chunks per instruction, and for LDS.64 / LDS.32 with 8 distinct words, at 1..16 warps per SM.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float sum(float v) { return v; }
__device__ __forceinline__ float sum(float2 v) { return v.x + v.y; }
__device__ __forceinline__ float sum(float4 v) { return v.x + v.y + v.z + v.w; }
template <int MODE, typename T>
__global__ void k(float* out, int iters) {
  extern __shared__ float4 sm4[];
  T* sm = reinterpret_cast<T*>(sm4);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int N = 2048 * 16 / sizeof(T);
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm4[i] = make_float4(i, 1, 2, 3);
  __syncthreads();
  int idx;
  if (MODE == 0) idx = lane;                                        // 32 distinct elements
  else if (MODE == 1) idx = ((lane % 16) / 4) * 17 + (lane / 16);   // 8 distinct (the sweep's (row block, site group))
  else idx = 0;                                                     // full broadcast
  idx += warp * 64;
  float acc = 0.f;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      T v = sm[(idx + u + (it & 7) * 16) & (N - 1)];
      acc += sum(v);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0) / (iters * 16);
  if (acc == 12345.f) out[1] = acc;
}
template <int MODE, typename T>
float run(float* d, int warps) {
  float r;
  for (int rep = 0; rep < 2; ++rep) {
    k<MODE, T><<<148, warps * 32, 2048 * 16>>>(d, 2000);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(&r, d, 4, cudaMemcpyDeviceToHost);
  return r / warps;  // SM cycles per warp-wide load instruction
}
int main() {
  float* d; cudaMalloc(&d, 64);
  for (int warps : {1, 4, 8, 16}) {
    printf("warps/SM %2d  SM-cycles per load:  LDS.128 distinct32 %.2f  sweep8 %.2f  bcast %.2f | LDS.64 distinct32 %.2f sweep8 %.2f | LDS.32 distinct32 %.2f sweep8 %.2f\n",
           warps, run<0, float4>(d, warps), run<1, float4>(d, warps), run<2, float4>(d, warps),
           run<0, float2>(d, warps), run<1, float2>(d, warps), run<0, float>(d, warps), run<1, float>(d, warps));
  }
  return 0;
}
