// Dependent-chain latencies (cycles) of the operations on potf2's column-to-column critical path, one CTA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat_bench lat_bench.cu && ./lat_bench
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 256;

template <int OP>
__global__ void chain(double seed, double* out, long long* cyc) {
  __shared__ double sm[1024];
  double x = seed + threadIdx.x * 1e-9;
  double y = seed * 0.5;
  sm[threadIdx.x] = x;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < kIters; ++i) {
    if (OP == 0)
      x = fma(x, y, y);
    else if (OP == 1)
      x = x * y;
    else if (OP == 2)
      x = rsqrt(x) + 1.0;
    else if (OP == 3)
      x = 1.0 / x + 1.0;
    else if (OP == 4)
      x = sqrt(x) + 1.0;
    else if (OP == 5) {  // smem round trip through another thread's slot
      sm[threadIdx.x] = x;
      __syncthreads();
      x = sm[(threadIdx.x + 1) & (blockDim.x - 1)] + 1.0;
    }
    else if (OP == 6)
      x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31) + 1.0;
    else if (OP == 7) {  // warp-local smem round trip
      sm[threadIdx.x] = x;
      __syncwarp();
      x = sm[threadIdx.x ^ 1] + 1.0;
      __syncwarp();
    }
    else if (OP == 8) {
      float f = rsqrtf((float)x);
      double r = (double)f;
      double e = fma(-x * r, r, 1.0);
      r = fma(r * 0.5, e, r);
      e = fma(-x * r, r, 1.0);
      r = fma(r * 0.5, e, r);
      x = r + 1.0;
    }
    else if (OP == 9)
      x = x + y;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0)
    cyc[0] = t1 - t0;
  out[threadIdx.x] = x;
}

template <int OP>
void run(const char* name, int threads) {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1024 * sizeof(double));
  cudaMalloc(&cyc, sizeof(long long));
  for (int rep = 0; rep < 2; ++rep)
    chain<OP><<<1, threads>>>(1.37, out, cyc);
  long long h = 0;
  cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-28s threads=%4d  %7.1f cycles/iter\n", name, threads, (double)h / kIters);
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  for (int threads : {32, 512}) {
    run<0>("dfma", threads);
    run<1>("dmul", threads);
    run<9>("dadd", threads);
    run<2>("rsqrt(double)+dadd", threads);
    run<3>("1/x (double)+dadd", threads);
    run<4>("sqrt(double)+dadd", threads);
    run<8>("rsqrtf+2 newton+dadd", threads);
    run<5>("sts+bar+lds+dadd", threads);
    run<6>("shfl+dadd", threads);
    run<7>("sts+syncwarp+lds+dadd", threads);
  }
  return 0;
}
