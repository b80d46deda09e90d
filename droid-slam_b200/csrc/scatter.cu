// Segmented mean over edges: the scatter_mean of GraphAgg.forward (droid_slam/droid_net.py:59-75,
// `net = scatter_mean(net, ix, dim=1)` with `_, ix = torch.unique(ii, return_inverse=True)`), i.e. the mean of the
// per-edge feature maps of all edges that leave the same source frame.  torch_scatter does it with a scatter-add of
// the values and one of ones (atomics, accumulation in the storage type, order undefined) followed by a division;
// here: a counting sort of the edges by group on the device (one CTA), then one pass in which a thread owns 8 or 4
// consecutive elements of a group's map, walks the group's edges in ascending edge order, accumulates in float32 and
// rounds once.  Bytes: N_e * D * s read, M * D * s written -- HBM-bound streaming.
#include <cuda_fp16.h>

#include "common.cuh"

namespace droid {
namespace {

// counting sort by group: start[m] .. start[m+1] index `order`, ascending edge id inside a group
__global__ void __launch_bounds__(1024) scatter_plan_kernel(const int64_t* __restrict__ ix, int n, int M,
                                                            int* __restrict__ start, int* __restrict__ order,
                                                            int* __restrict__ status) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  const int tid = threadIdx.x;
  for (int m = tid; m <= M; m += 1024) start[m] = 0;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int e = tid; e < n; e += 1024) {
    const long long g = ix[e];
    if (g < 0 || g >= M) { atomicOr(status, 1); continue; }
    atomicAdd(&start[g + 1], 1);
  }
  __syncthreads();
  // inclusive scan of the counts: start[m + 1] = number of edges in groups 0 .. m
  for (int m0 = 0; m0 < M; m0 += 1024) {
    const int m = m0 + tid;
    int x = m < M ? start[m + 1] : 0;
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
      s_warp[lane] = w;
    }
    __syncthreads();
    const int incl = x + (warp > 0 ? s_warp[warp - 1] : 0) + s_base;
    __syncthreads();
    if (m < M) start[m + 1] = incl;
    if (tid == 1023) s_base = incl;
    __syncthreads();
  }
  // ascending edge order inside a group: one thread per group walks the edge list (n is a few hundred)
  for (int m = tid; m < M; m += 1024) {
    int o = start[m];
    const int end = start[m + 1];
    for (int e = 0; e < n && o < end; e++)
      if (ix[e] == m) order[o++] = e;
  }
}

__device__ __forceinline__ void acc_add(float (&a)[4], const float4& v) { a[0] += v.x; a[1] += v.y; a[2] += v.z; a[3] += v.w; }
__device__ __forceinline__ void acc_add(float (&a)[8], const uint4& v) {
  const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int k = 0; k < 4; k++) { const float2 f = __half22float2(h[k]); a[2 * k] += f.x; a[2 * k + 1] += f.y; }
}
__device__ __forceinline__ float4 acc_out(const float (&a)[4], float inv, float4*) {
  return make_float4(a[0] * inv, a[1] * inv, a[2] * inv, a[3] * inv);
}
__device__ __forceinline__ uint4 acc_out(const float (&a)[8], float inv, uint4*) {
  uint4 r;
  __half2* h = reinterpret_cast<__half2*>(&r);
#pragma unroll
  for (int k = 0; k < 4; k++) h[k] = __floats2half2_rn(a[2 * k] * inv, a[2 * k + 1] * inv);
  return r;
}

// T = element type, VT = 16-byte vector of V elements; D = elements per map (multiple of V);
// thread = V consecutive elements of group blockIdx.y
template <typename T, typename VT, int V>
__global__ void __launch_bounds__(256) scatter_mean_kernel(const T* __restrict__ src, const int* __restrict__ start,
                                                           const int* __restrict__ order, T* __restrict__ out, size_t D) {
  const int m = blockIdx.y;
  const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * V;
  if (i >= D) return;
  const int e0 = start[m], e1 = start[m + 1];
  float a[V];
#pragma unroll
  for (int k = 0; k < V; k++) a[k] = 0.f;
  int e = e0;
  for (; e + 1 < e1; e += 2) {      // two maps in flight
    const VT v0 = __ldcs(reinterpret_cast<const VT*>(src + (size_t)order[e] * D + i));
    const VT v1 = __ldcs(reinterpret_cast<const VT*>(src + (size_t)order[e + 1] * D + i));
    acc_add(a, v0);
    acc_add(a, v1);
  }
  if (e < e1) acc_add(a, __ldcs(reinterpret_cast<const VT*>(src + (size_t)order[e] * D + i)));
  const float inv = 1.0f / (float)max(e1 - e0, 1);
  *reinterpret_cast<VT*>(out + (size_t)m * D + i) = acc_out(a, inv, (VT*)nullptr);
}

size_t plan_ints_bytes(size_t count) { return (count * sizeof(int) + 255) / 256 * 256; }

}  // namespace
}  // namespace droid

using namespace droid;

extern "C" size_t droid_scatter_mean_workspace_bytes(int n, int M) {
  if (n < 0 || M < 0) return 0;
  return 256 + plan_ints_bytes((size_t)M + 2) + plan_ints_bytes((size_t)n + 1) + 256;
}

// dtype: 0 = float32, 1 = float16.  src [n, D], ix [n] in [0, M), out [M, D]; D % 8 == 0 (float16) / % 4 == 0 (float32)
extern "C" int droid_scatter_mean(const void* src, const int64_t* ix, void* out, int dtype, int n, int M, long long D,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  if (n < 0 || M < 0 || D < 0 || (dtype != 0 && dtype != 1) || M > 65535) return DROID_ERR_INVALID_ARGUMENT;
  if (M == 0 || D == 0) return DROID_OK;
  if (!out || !workspace || (n > 0 && (!src || !ix))) return DROID_ERR_INVALID_ARGUMENT;
  const int V = dtype == 0 ? 4 : 8;
  if (D % V != 0 || ((uintptr_t)src & 15) != 0 || ((uintptr_t)out & 15) != 0) return DROID_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)workspace & 255) != 0) {
    const size_t adj = 256 - ((uintptr_t)workspace & 255);
    if (workspace_bytes < adj) return DROID_ERR_WORKSPACE_TOO_SMALL;
    workspace = (char*)workspace + adj;
    workspace_bytes -= adj;
  }
  if (workspace_bytes + 256 < droid_scatter_mean_workspace_bytes(n, M)) return DROID_ERR_WORKSPACE_TOO_SMALL;
  int* status = (int*)workspace;
  int* start = (int*)((char*)workspace + 256);
  int* order = (int*)((char*)start + plan_ints_bytes((size_t)M + 2));
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(status, 0, sizeof(int), st);
  DROID_LAUNCH scatter_plan_kernel<<<1, 1024, 0, st>>>(ix, n, M, start, order, status);
  const dim3 grid((unsigned)((D / V + 255) / 256), M);
  if (dtype == 0)
    DROID_LAUNCH scatter_mean_kernel<float, float4, 4><<<grid, 256, 0, st>>>((const float*)src, start, order, (float*)out, (size_t)D);
  else
    DROID_LAUNCH scatter_mean_kernel<__half, uint4, 8><<<grid, 256, 0, st>>>((const __half*)src, start, order, (__half*)out, (size_t)D);
  return check_launch();
}
