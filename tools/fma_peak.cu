// fma_peak.cu -- measured non-tensor FMA peaks of the GPU: register-resident
// dependent-chain FFMA and DFMA loops, every SM filled, >= 1 s per precision, timed with CUDA events.
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_peak tools/fma_peak.cu
// Output: one JSON object on stdout (committed as profiles/r2/fma_peak.json; bench.py reads it).
#include <cstdio>
#include <cuda_runtime.h>

template <typename T, int ILP>
__global__ void __launch_bounds__(256) fma_loop(T* out, int iters, T a, T b) {
  T x[ILP];
#pragma unroll
  for (int k = 0; k < ILP; k++) x[k] = (T)(threadIdx.x + k) * (T)1e-3;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
      for (int k = 0; k < ILP; k++) x[k] = x[k] * a + b;  // contracted to one FMA
  }
  T s = 0;
#pragma unroll
  for (int k = 0; k < ILP; k++) s += x[k];
  if (s == (T)123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // never true: keeps the chain alive
}

template <typename T, int ILP>
static double measure(int sms, double* seconds) {
  T* out;
  cudaMalloc(&out, sizeof(T) * 1024);
  const int blocks = sms * 8, threads = 256;
  int iters = 1 << 12;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms = 0;
  for (;;) {  // grow until one launch takes >= 1 s
    fma_loop<T, ILP><<<blocks, threads>>>(out, iters, (T)0.999, (T)1e-4);  // warm-up at this size
    cudaEventRecord(e0);
    fma_loop<T, ILP><<<blocks, threads>>>(out, iters, (T)0.999, (T)1e-4);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms >= 1000.f || iters >= (1 << 26)) break;
    iters *= 2;
  }
  double flops = 2.0 * (double)blocks * threads * (double)iters * 8 * ILP;
  *seconds = ms * 1e-3;
  cudaFree(out);
  return flops / (ms * 1e-3);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  double t32, t64;
  double f32 = measure<float, 8>(p.multiProcessorCount, &t32);
  double f64 = measure<double, 8>(p.multiProcessorCount, &t64);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_max\": %.0f, \"fp32_fma_tflops\": %.2f, \"fp64_fma_tflops\": %.2f, "
         "\"fp32_seconds\": %.2f, \"fp64_seconds\": %.2f, "
         "\"how\": \"register-resident FMA chains, 8 independent accumulators per thread, 256 threads x 8 blocks per SM, one launch >= 1 s, CUDA events\"}\n",
         p.name, p.multiProcessorCount, clk_khz / 1000.0, f32 / 1e12, f64 / 1e12, t32, t64);
  return 0;
}
