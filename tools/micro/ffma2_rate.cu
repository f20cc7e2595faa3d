// Issue rate of the packed fp32 FMA of sm_100 (fma.rn.f32x2 -> FFMA2) against the scalar FFMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_rate tools/micro/ffma2_rate.cu && /tmp/ffma2_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int PACKED>
__global__ void __launch_bounds__(256) k(float* out, int iters, float a, float b) {
  float x[8], y[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { x[j] = threadIdx.x + j; y[j] = threadIdx.x * 0.5f + j; }
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (PACKED) {
        unsigned long long v, s, t;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(x[j]), "f"(y[j]));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(s) : "f"(a));
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(t) : "f"(b));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(s), "l"(t));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(x[j]), "=f"(y[j]) : "l"(v));
      } else {
        x[j] = fmaf(x[j], a, b);
        y[j] = fmaf(y[j], a, b);
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += x[j] + y[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  float* out;
  cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  const int iters = 20000;
  for (int packed = 0; packed < 2; ++packed) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (packed) k<1><<<148 * 8, 256>>>(out, iters, 0.999f, 0.001f);
      else k<0><<<148 * 8, 256>>>(out, iters, 0.999f, 0.001f);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fma = 148.0 * 8 * 256 * (double)iters * 16;
    printf("%s: %.3f ms  %.1f TFMA/s  (%.1f FMA per clock per SM at 1.965 GHz)\n", packed ? "fma.rn.f32x2" : "fmaf scalar", ms,
           fma / ms / 1e9, fma / (ms * 1e-3) / 148 / 1.965e9);
  }
  return 0;
}
