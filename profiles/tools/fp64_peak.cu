// FP64 peak microbenchmark for B200 (sm_100a): register-resident DMMA.8x8x4 and DFMA loops.
// Measurement tool only ("FP64 tensor (DMMA) peak is NOT in MEASURED_PEAKS.json").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu
// Prints one JSON line with TFLOP/s for DMMA with NACC independent accumulators and for DFMA.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1);} } while (0)

template <int NACC>
__global__ void __launch_bounds__(256) dmma_loop(double* out, int iters, double a0, double b0) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x * 1e-9; c[i][1] = i * 1e-9; }
  double a = a0 + threadIdx.x * 1e-12, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) dfma_loop(double* out, int iters, double a0, double b0) {
  double c[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) c[i] = threadIdx.x * 1e-9 + i;
  double a = a0, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fma(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
static double time_ms(F launch) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  launch(); launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  int dev = 0; cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
  int sms = p.multiProcessorCount;
  double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 8 * 1024));
  const int iters = 20000;
  printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
  // DMMA: flops per instruction per warp = 2*8*8*4 = 512
  for (int wpb : {4, 8, 16}) {
    for (int bps : {1, 2}) {
      int grid = sms * bps, threads = wpb * 32;
      auto tf = [&](double ms, int nacc) {
        return (double)grid * wpb * iters * nacc * 512.0 / (ms * 1e-3) / 1e12; };
      double m2 = time_ms([&] { dmma_loop<2><<<grid, threads>>>(out, iters, 1.0, 1e-9); });
      double m4 = time_ms([&] { dmma_loop<4><<<grid, threads>>>(out, iters, 1.0, 1e-9); });
      double m8 = time_ms([&] { dmma_loop<8><<<grid, threads>>>(out, iters, 1.0, 1e-9); });
      double m16 = time_ms([&] { dmma_loop<16><<<grid, threads>>>(out, iters, 1.0, 1e-9); });
      printf(", \"dmma_w%d_b%d\": [%.2f, %.2f, %.2f, %.2f]", wpb, bps, tf(m2, 2), tf(m4, 4), tf(m8, 8),
             tf(m16, 16));
    }
  }
  for (int wpb : {8, 16}) {
    int grid = sms * 2, threads = wpb * 32;
    double m = time_ms([&] { dfma_loop<16><<<grid, threads>>>(out, iters, 1.0000001, 1e-9); });
    printf(", \"dfma_w%d\": %.2f", wpb, (double)grid * threads * iters * 16 * 2.0 / (m * 1e-3) / 1e12);
  }
  CK(cudaDeviceSynchronize());
  printf("}\n");
  return 0;
}
