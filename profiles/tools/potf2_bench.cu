// Times the critical-path tile operations alone: potf2 (modes factor / invert / both), potrf_tile(nb), trsm_panel.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I dla-future_b200/csrc -I include \
//        -o profiles/tools/potf2_bench profiles/tools/potf2_bench.cu -L dla-future_b200/lib -ldlaf_b200 \
//        -Xlinker -rpath -Xlinker $PWD/dla-future_b200/lib
#include <cstdio>
#include <vector>

#include "tile_ops.cuh"
#ifdef DLAFB_TIMING
#include "potf2.cu"  // instrumented private copy (phase time stamps)
#endif

using namespace dlafb;

template <class F>
static float time_us(F f, int reps = 20) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms * 1000.f / reps;
}

int main() {
  const int nb = 512, m = 16384;
  std::vector<double> h((size_t)nb * nb, 0.0);
  for (int j = 0; j < nb; ++j)
    for (int i = 0; i < nb; ++i)
      h[i + (size_t)j * nb] = (i == j) ? 2.0 * nb : 1.0 / (1 + ((i * 7 + j * 13) % 17));
  for (int j = 0; j < nb; ++j)
    for (int i = 0; i < j; ++i)
      h[i + (size_t)j * nb] = h[j + (size_t)i * nb];
  double *a, *a0, *w, *b;
  int* info;
  cudaMalloc(&a, sizeof(double) * nb * nb);
  cudaMalloc(&a0, sizeof(double) * nb * nb);
  cudaMalloc(&w, sizeof(double) * (nb * nb * 2 + 4 * 128 * 128));
  cudaMalloc(&b, sizeof(double) * (size_t)m * nb);
  cudaMalloc(&info, sizeof(int));
  cudaMemset(info, 0, sizeof(int));
  cudaMemcpy(a0, h.data(), sizeof(double) * nb * nb, cudaMemcpyHostToDevice);
  cudaMemset(b, 0, sizeof(double) * (size_t)m * nb);
  auto reset = [&] { cudaMemcpyAsync(a, a0, sizeof(double) * nb * nb, cudaMemcpyDeviceToDevice, 0); };
  const float t_reset = time_us(reset);
  printf("{\"reset_us\": %.1f", t_reset);
  for (int n : {32, 64, 128}) {
    for (int mode : {1, 2, 3}) {
      const float t = time_us([&] {
        reset();
        launch_potf2<double>(0, 'L', n, a, nb, w, 128, info, 0, mode);
      });
      printf(", \"potf2_n%d_mode%d_us\": %.1f", n, mode, t - t_reset);
    }
  }
  const float t_potrf = time_us([&] {
    reset();
    potrf_tile<double>(0, 'L', nb, a, nb, w, 128, info, 0);
  });
  printf(", \"potrf_tile_512_us\": %.1f", t_potrf - t_reset);
  for (int rows : {16384, 4096, 1024}) {
    const float t = time_us([&] { trsm_panel<double>(0, nb, a, nb, w, 128, b, rows, rows); });
    printf(", \"trsm_panel_m%d_us\": %.1f", rows, t);
  }
#ifdef DLAFB_TIMING
  reset();
  launch_potf2<double>(0, 'L', 128, a, nb, w, 128, info, 0, 3);
  cudaDeviceSynchronize();
  long long clk[64];
  cudaMemcpyFromSymbol(clk, g_potf2_clk, sizeof(clk));
  printf(", \"clk\": [");
  for (int i = 0; i < 63; ++i) printf("%lld%s", clk[i] ? clk[i] - clk[0] : -1, i < 62 ? ", " : "]");
#endif
  int hinfo = -1;
  cudaMemcpy(&hinfo, info, sizeof(int), cudaMemcpyDeviceToHost);
  printf(", \"info\": %d}\n", hinfo);
  return 0;
}
