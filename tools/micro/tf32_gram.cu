// Probe of the tf32 MN-major UMMA operand layout used by schur_tc.cu: G = F F^T for one 128 x 64 chunk.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I droid-slam_b200/csrc -I include \
//        tools/micro/tf32_gram.cu droid-slam_b200/build/geom.o -o tools/micro/tf32_gram
#include "../../droid-slam_b200/csrc/schur_tc.cu"
#include <cstdio>
#include <cmath>
#include <vector>
using namespace droid::stc;
static constexpr int kHalf = 32 * 1024, kStage = 2 * kHalf;

__device__ __host__ inline float fval(int r, int p) { return (float)(((r * 37 + p * 11) % 17) - 8) * 0.125f + (r == p ? 1.f : 0.f); }

__global__ void __launch_bounds__(128, 1) probe(float* out, int layout_type, int T, int split) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int lbo = T * 128;
  if (tid == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(128) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < kStage / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  for (int e = tid; e < 128 * T; e += 128) {
    const int r = e % 128, p = e / 128;
    put_split(smem, kHalf, lbo, p, r, fval(r, p));
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = slot;
  if (tid == 0) {
    const uint32_t hi = smem_u32(smem), lo = hi + kHalf;
    for (int kk = 0; kk < T / 8; kk++) {
      uint64_t a_hi = umma_desc(hi + kk * 1024, lbo), a_lo = umma_desc(lo + kk * 1024, lbo);
      if (layout_type != 1) {
        a_hi = (a_hi & ~((uint64_t)7 << 61)) | ((uint64_t)layout_type << 61);
        a_lo = (a_lo & ~((uint64_t)7 << 61)) | ((uint64_t)layout_type << 61);
      }
      umma_tf32(tmem, a_hi, a_hi, umma_idesc(), kk > 0 ? 1u : 0u);
      if (split) { umma_tf32(tmem, a_hi, a_lo, umma_idesc(), 1u); umma_tf32(tmem, a_lo, a_hi, umma_idesc(), 1u); }
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  float v[64];
  for (int h = 0; h < 2; h++) {
    const uint32_t taddr = tmem + 64 * h + ((uint32_t)(32 * warp) << 16);
    tmem_ld16(taddr, v); tmem_ld16(taddr + 16, v + 16); tmem_ld16(taddr + 32, v + 32); tmem_ld16(taddr + 48, v + 48);
    tmem_ld_wait();
    for (int j = 0; j < 64; j++) out[tid * 128 + 64 * h + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128) : "memory");
}

int main() {
  float* d; cudaMalloc(&d, 128 * 128 * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, kStage + 1024);
  std::vector<float> h(128 * 128);
  for (int T : {64, 32})
    for (int lt : {1})
      for (int split : {0, 1}) {
        cudaMemset(d, 0, 128 * 128 * 4);
        probe<<<1, 128, kStage + 1024>>>(d, lt, T, split);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(h.data(), d, 128 * 128 * 4, cudaMemcpyDeviceToHost);
        double worst = 0, scale = 0; int wi = 0, wj = 0;
        for (int i = 0; i < 128; i++)
          for (int j = 0; j < 128; j++) {
            double ref = 0;
            for (int p = 0; p < T; p++) ref += (double)fval(i, p) * fval(j, p);
            const double df = fabs(ref - h[i * 128 + j]);
            if (df > worst) { worst = df; wi = i; wj = j; }
            scale = fmax(scale, fabs(ref));
          }
        if (lt == 1 && split == 0) {
          // which output rows / columns are right, and do wrong entries match another (i', j')?
          printf("rows ok: ");
          for (int i = 0; i < 128; i++) {
            int ok = 1;
            for (int j = 0; j < 128; j++) {
              double ref = 0; for (int p = 0; p < T; p++) ref += (double)fval(i, p) * fval(j, p);
              if (fabs(ref - h[i * 128 + j]) > 1e-2) ok = 0;
            }
            printf("%d", ok);
          }
          printf("\nrow 0 cols ok: ");
          for (int j = 0; j < 128; j++) {
            double ref = 0; for (int p = 0; p < T; p++) ref += (double)fval(0, p) * fval(j, p);
            printf("%d", fabs(ref - h[j]) <= 1e-2);
          }
          printf("\nrow 0 first 12: ");
          for (int j = 0; j < 12; j++) printf("%.3f ", h[j]);
          printf("\nref        : ");
          for (int j = 0; j < 12; j++) { double ref = 0; for (int p = 0; p < T; p++) ref += (double)fval(0, p) * fval(j, p); printf("%.3f ", ref); }
          printf("\n");
        }
        printf("T=%d layout=%d split=%d : %s  max|err| %.3e (scale %.2f) at (%d,%d) got %.4f\n", T, lt, split,
               cudaGetErrorString(e), worst, scale, wi, wj, h[wi * 128 + wj]);
      }
  return 0;
}
