// Throughput probe for the min-plus inner loop of the general-cost hard Sankoff kernel on sm_100a:
//   0: FFMA + FMNMX3 on packed float keys (what hard_fwd_kernel does today: 1.5 instr per candidate)
//   1: __viaddmin_s32     (DPX: min(a + b, c), 1 instr per candidate)
//   2: __viaddmin_s16x2   (DPX, two 16-bit lanes per instr)
//   3: IADD3 + VIMNMX3-style (__vimin3_s32 on two sums)
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/_bin/dpx_probe tools/dpx_probe.cu ; run: tools/_bin/dpx_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

constexpr int kChains = 8, kT = 20, kIters = 20000;

template <int MODE>
__global__ void __launch_bounds__(128) probe(const int* __restrict__ in, int* __restrict__ out, long long* cyc) {
  const long long t0 = clock64();
  int c[kT];
  for (int t = 0; t < kT; ++t) c[t] = in[t];
  int d[kChains][kT];
  for (int k = 0; k < kChains; ++k)
    for (int t = 0; t < kT; ++t) d[k][t] = in[32 + k * kT + t] + threadIdx.x;
  int acc[kChains];
  for (int k = 0; k < kChains; ++k) acc[k] = 0;
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int k = 0; k < kChains; ++k) {
      if constexpr (MODE == 0) {
        float b = fmaf(__int_as_float(c[0]), 32.f, __int_as_float(d[k][0]));
#pragma unroll
        for (int t = 1; t < kT; ++t) b = fminf(b, fmaf(__int_as_float(c[t]), 32.f, __int_as_float(d[k][t])));
        acc[k] += __float_as_int(b);
      } else if constexpr (MODE == 1) {
        int b = c[0] + d[k][0];
#pragma unroll
        for (int t = 1; t < kT; ++t) b = __viaddmin_s32(c[t], d[k][t], b);
        acc[k] += b;
      } else if constexpr (MODE == 2) {
        unsigned b = __vadd2((unsigned)c[0], (unsigned)d[k][0]);
#pragma unroll
        for (int t = 1; t < kT; ++t) b = __viaddmin_s16x2((unsigned)c[t], (unsigned)d[k][t], b);
        acc[k] += (int)b;
      } else if constexpr (MODE == 3) {
        int b = c[0] + d[k][0];
#pragma unroll
        for (int t = 1; t + 1 < kT; t += 2) b = __vimin3_s32(c[t] + d[k][t], c[t + 1] + d[k][t + 1], b);
        b = min(b, c[kT - 1] + d[k][kT - 1]);
        acc[k] += b;
      } else if constexpr (MODE == 4) {  // FFMA + 2-input FMNMX
        float b = fmaf(__int_as_float(c[0]), 32.f, __int_as_float(d[k][0]));
#pragma unroll
        for (int t = 1; t < kT; ++t) {
          float x = fmaf(__int_as_float(c[t]), 32.f, __int_as_float(d[k][t]));
          asm("min.f32 %0, %0, %1;" : "+f"(b) : "f"(x));
        }
        acc[k] += __float_as_int(b);
      } else if constexpr (MODE == 5) {  // min only, 3-input float (rate of FMNMX3 alone)
        float b = __int_as_float(d[k][0]);
#pragma unroll
        for (int t = 1; t + 1 < kT; t += 2) b = fminf(fminf(b, __int_as_float(d[k][t] ^ c[t])), __int_as_float(d[k][t + 1]));
        acc[k] += __float_as_int(b);
      } else if constexpr (MODE == 6) {  // FFMA, then the min on the float bit patterns as integers (VIMNMX3)
        int b = __float_as_int(fmaf(__int_as_float(c[0]), 32.f, __int_as_float(d[k][0])));
#pragma unroll
        for (int t = 1; t + 1 < kT; t += 2)
          b = __vimin3_s32(__float_as_int(fmaf(__int_as_float(c[t]), 32.f, __int_as_float(d[k][t]))),
                           __float_as_int(fmaf(__int_as_float(c[t + 1]), 32.f, __int_as_float(d[k][t + 1]))), b);
        acc[k] += b;
      } else if constexpr (MODE == 7) {  // half the chains float (FFMA + FMNMX3), half DPX (VIADDMNMX)
        if (k & 1) {
          int b = c[0] + d[k][0];
#pragma unroll
          for (int t = 1; t < kT; ++t) b = __viaddmin_s32(c[t], d[k][t], b);
          acc[k] += b;
        } else {
          float b = fmaf(__int_as_float(c[0]), 32.f, __int_as_float(d[k][0]));
#pragma unroll
          for (int t = 1; t < kT; ++t) b = fminf(b, fmaf(__int_as_float(c[t]), 32.f, __int_as_float(d[k][t])));
          acc[k] += __float_as_int(b);
        }
      } else if constexpr (MODE == 8) {  // fp16x2: HADD2 + HMNMX2, two candidates per instruction pair
        __half2 b = __hadd2(*reinterpret_cast<const __half2*>(&c[0]), *reinterpret_cast<const __half2*>(&d[k][0]));
#pragma unroll
        for (int t = 1; t < kT; ++t)
          b = __hmin2(b, __hadd2(*reinterpret_cast<const __half2*>(&c[t]), *reinterpret_cast<const __half2*>(&d[k][t])));
        acc[k] += *reinterpret_cast<int*>(&b);
      } else {  // MODE 9: FFMA only (issue / FMA-pipe ceiling)
        float b = __int_as_float(d[k][0]);
#pragma unroll
        for (int t = 1; t < kT; ++t) b = fmaf(__int_as_float(c[t]), b, __int_as_float(d[k][t]));
        acc[k] += __float_as_int(b);
      }
    }
    // perturb so the compiler cannot hoist the chains
#pragma unroll
    for (int k = 0; k < kChains; ++k) d[k][0] ^= acc[k] & 1;
#pragma unroll
    for (int t = 0; t < kT; ++t) c[t] ^= acc[t % kChains] & 1;  // 20 LOP3 per 160 candidates, the same in every mode
  }
  int r = 0;
  for (int k = 0; k < kChains; ++k) r += acc[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = clock64() - t0;
}

template <int MODE>
static void run(const char* name, const int* in, int* out, int ctas) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  long long* cyc; cudaMalloc(&cyc, 8);
  probe<MODE><<<ctas, 128>>>(in, out, cyc);
  cudaEventRecord(e0);
  probe<MODE><<<ctas, 128>>>(in, out, cyc);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double cand = (double)ctas * 128 * kIters * kChains * kT;
  (void)h;  // CTA 0's cycles: not the kernel's (two CTAs per SM at ~200 registers, so two waves)
  printf("%-30s %8.3f ms  %8.1f G candidates/s  (%.1f per clk per SM at 1.9 GHz; ~30%% of the issue slots are the probe's own\n"
         "%-30s perturbation code, so read the modes against each other, not against a peak)\n", name, ms, cand / ms * 1e-6,
         cand / (ms * 1e-3) / 148 / 1.9e9, "");
}

int main() {
  int *in, *out;
  cudaMalloc(&in, 4096); cudaMemset(in, 0, 4096);
  const int ctas = 148 * 4;
  cudaMalloc(&out, ctas * 128 * 4);
  run<0>("ffma+fmnmx3 (float keys)", in, out, ctas);
  run<1>("viaddmin_s32", in, out, ctas);
  run<2>("viaddmin_s16x2 (per lane x2)", in, out, ctas);
  run<3>("iadd + vimin3_s32", in, out, ctas);
  run<4>("ffma + fmnmx (2-input)", in, out, ctas);
  run<5>("fmnmx3 alone (x2 per instr)", in, out, ctas);
  run<6>("ffma + vimnmx3 on float bits", in, out, ctas);
  run<7>("half float keys, half dpx", in, out, ctas);
  run<8>("hadd2 + hmnmx2 (per lane x2)", in, out, ctas);
  run<9>("ffma alone", in, out, ctas);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
