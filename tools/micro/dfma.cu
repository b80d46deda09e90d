// Microbenchmark: DFMA dependent latency and per-SM throughput on sm_100a (informs chol.cu design).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void lat(double* out, long long* cyc, double x, double y) {
  double a = x;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 64; i++) {
#pragma unroll
    for (int j = 0; j < 16; j++) a = fma(a, y, x);
  }
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int ILP>
__global__ void thr(double* out, long long* cyc, double x, double y) {
  double a[ILP];
#pragma unroll
  for (int k = 0; k < ILP; k++) a[k] = x + k;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) {
#pragma unroll
    for (int k = 0; k < ILP; k++) a[k] = fma(a[k], y, x);
  }
  __syncthreads();
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int k = 0; k < ILP; k++) s += a[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void rcp(double* out, long long* cyc, double x) {
  double a = x;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) a = 1.0 / a + 0.5;
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void bar(long long* cyc) {
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double* out; long long* cyc; long long h;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 64);
  lat<<<1, 32>>>(out, cyc, 1.0, 0.5); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("DFMA dependent latency: %.2f clk\n", h / 1024.0);
  for (int w = 1; w <= 16; w *= 2) {
    thr<16><<<1, 32 * w>>>(out, cyc, 1.0, 0.5); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("warps %2d ILP16: %.2f clk per warp-DFMA-round => %.1f DFMA lanes/clk/SM\n", w, h / 256.0 / 16, 256.0 * 16 * 32 * w / h);
  }
  rcp<<<1, 32>>>(out, cyc, 3.0); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("1.0/x + add dependent: %.1f clk\n", h / 256.0);
  for (int w = 2; w <= 8; w *= 2) { bar<<<1, 32 * w>>>(cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("syncthreads %d warps: %.1f clk\n", w, h / 256.0); }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
