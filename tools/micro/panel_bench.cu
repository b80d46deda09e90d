// Micro-benchmark of chol_panel2_kernel alone (few CTAs, like the envelope solve).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -DPANEL_TS \
//        -I droid-slam_b200/csrc -I include tools/micro/panel_bench.cu droid-slam_b200/build/geom.o -o tools/micro/panel_bench
#include "../../droid-slam_b200/csrc/chol.cu"
#include <cstdio>
#include <vector>
int main() {
  const int n = 4096, j0 = 2048, jb = 64, ncta = 3;
  std::vector<double> h((size_t)(n + 1) * n, 0.0);
  for (int r = 0; r <= n; r++)
    for (int c = 0; c < n && c <= r; c++) h[(size_t)r * n + c] = (r == c) ? 100.0 + r % 7 : 0.01 * ((r * 31 + c * 17) % 13 - 6);
  double* A; int* info; int* arrive; cudaMalloc(&arrive, 4096);
  cudaMalloc(&A, h.size() * 8); cudaMalloc(&info, 4);
  cudaFuncSetAttribute(droid::chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, droid::kPanelSmem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9;
  for (int it = 0; it < 6; it++) {
    cudaMemcpy(A, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    cudaMemset(arrive, 0, 4096);
    cudaEventRecord(e0);
    droid::chol_panel_kernel<<<ncta, 256, droid::kPanelSmem>>>(A, n, j0, jb, info, nullptr, arrive, 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  double s = 0; std::vector<double> o(64);
  cudaMemcpy(o.data(), A + (size_t)(j0 + 63) * n + j0, 64 * 8, cudaMemcpyDeviceToHost);
  for (double v : o) s += v;
  printf("panel kernel %.2f us  (%s) checksum %.6f\n", best * 1e3, cudaGetErrorString(cudaGetLastError()), s);
#ifdef PANEL_TS
  long long ts[16]; cudaMemcpyFromSymbol(ts, droid::g_ts, sizeof(ts));
  printf("  stage %lld  elim %lld  rs %lld  wait+wb %lld clk\n", ts[1]-ts[0], ts[2]-ts[1], ts[3]-ts[2], ts[4]-ts[3]);
#endif
  return 0;
}
