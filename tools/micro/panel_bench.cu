// Micro-benchmark of chol_panel_kernel alone (few CTAs, like the envelope solve).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -DPANEL_TS \
//        -I droid-slam_b200/csrc -I include tools/micro/panel_bench.cu droid-slam_b200/build/runtime.o -o tools/micro/panel_bench
#include "../../droid-slam_b200/csrc/chol.cu"
#include <cstdio>
#include <vector>
int main() {
  const int n = 4096, j0 = 2048, jb = 64;
  std::vector<double> h((size_t)(n + 1) * n, 0.0);
  for (int r = 0; r <= n; r++)
    for (int c = 0; c < n && c <= r; c++) h[(size_t)r * n + c] = (r == c) ? 100.0 + r % 7 : 0.01 * ((r * 31 + c * 17) % 13 - 6);
  double* A; int* info; int* arrive; double* linv; cudaMalloc(&arrive, 4096); cudaMalloc(&linv, (size_t)n * 64 * 8);
  cudaMalloc(&A, h.size() * 8); cudaMalloc(&info, 4);
  cudaFuncSetAttribute(droid::chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, droid::kPanelSmem2);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int prev = 0; prev <= 64; prev += 64)
    for (int ncta = 2; ncta <= 16; ncta *= 2) {
      float best = 1e9;
      for (int it = 0; it < 6; it++) {
        cudaMemcpy(A, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
        cudaMemset(arrive, 0, 4096);
        cudaEventRecord(e0);
        droid::chol_panel_kernel<<<ncta, droid::kPanelThreads, prev ? droid::kPanelSmem2 : droid::kPanelSmem1>>>(
            A, n, j0, jb, prev, info, nullptr, arrive, 1, 0, 0, linv, nullptr);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      printf("prev %2d ctas %2d: panel kernel %.2f us (%s)", prev, ncta, best * 1e3, cudaGetErrorString(cudaGetLastError()));
#ifdef PANEL_TS
      long long ts[16]; cudaMemcpyFromSymbol(ts, droid::g_ts, sizeof(ts));
      printf("  stage+update %lld  elim %lld (%.0f clk/pivot)  rs %lld  wait+wb %lld clk", ts[1]-ts[0], ts[2]-ts[1], (ts[2]-ts[1]) / 64.0, ts[3]-ts[2], ts[4]-ts[3]);
#endif
      printf("\n");
    }
  return 0;
}
