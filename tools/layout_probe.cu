// Micro-benchmark for the round-2 layout question: a warp walks the internal nodes of a
// 1024-leaf tree and, per node, reads the 20 state rows of one earlier node and writes the 20 rows of its own
// — the memory pattern of the uniform forward kernels — in two layouts:
//   strided  [node][state][site_stride]        row step = site_stride (run-time), rows of a node 400 KB apart
//   blocked  [node][site/64][state][64]        row step = 64 floats (compile-time), a warp's 20 rows contiguous
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/layout_probe.bin tools/layout_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int A = 20, V = 2;

template <bool BLOCKED>
__global__ void __launch_bounds__(128) walk(float* __restrict__ tab, int n_int, long long n_sites, long long stride) {
  const long long site0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (site0 >= n_sites) return;
  const size_t node_step = (size_t)A * stride;
  const size_t off = BLOCKED ? (size_t)(site0 / 64) * A * 64 + site0 % 64 : (size_t)site0;
  float* base = tab + off;
  float d[A][V];
#pragma unroll
  for (int s = 0; s < A; ++s) d[s][0] = d[s][1] = (float)s;
  for (int k = 0; k < n_int; ++k) {
    float* out = base + (size_t)k * node_step;
    const float* in = base + (size_t)(k > 8 ? k - 7 : 0) * node_step;  // an earlier node (L2 miss at this distance)
#pragma unroll
    for (int s = 0; s < A; ++s) {
      const size_t r = BLOCKED ? (size_t)s * 64 : (size_t)s * stride;
      const float2 x = __ldcs(reinterpret_cast<const float2*>(in + r));
      d[s][0] = fminf(d[s][0], x.x) + 1.f;
      d[s][1] = fminf(d[s][1], x.y) + 1.f;
      *reinterpret_cast<float2*>(out + r) = make_float2(d[s][0], d[s][1]);
    }
  }
}

int main() {
  const int n_int = 1023;
  const long long L = 100000, stride = (L + 63) / 64 * 64;
  float* tab;
  const size_t bytes = sizeof(float) * (size_t)n_int * A * stride;
  if (cudaMalloc(&tab, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tab, 0, bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int threads = 128, grid = (int)((L / V + threads - 1) / threads);
  for (int rep = 0; rep < 2; ++rep)
    for (int blocked = 0; blocked < 2; ++blocked) {
      auto run = [&] { if (blocked) walk<true><<<grid, threads>>>(tab, n_int, L, stride); else walk<false><<<grid, threads>>>(tab, n_int, L, stride); };
      run(); cudaDeviceSynchronize();
      cudaEventRecord(e0); run(); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      const double gb = 2.0 * n_int * A * L * 4 / 1e9;
      printf("%s: %.3f ms, %.0f GB/s (read + write %.2f GB)\n", blocked ? "blocked" : "strided", ms, gb / (ms * 1e-3), gb);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
