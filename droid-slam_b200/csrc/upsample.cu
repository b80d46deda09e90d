// Convex upsampling: cvx_upsample of droid_slam/droid_net.py:21-35, which DepthVideo.upsample
// (depth_video.py:123-127) applies to the inverse-depth maps after every update and which PyTorch runs as
// softmax + unfold + broadcast multiply + sum + permute (five passes over a [batch, 9, 8, 8, ht, wd] tensor).
// One pass here:  out[b, 8y+i, 8x+j, :] = sum_k softmax_k(mask[b, (k, i, j), y, x]) * data[b, y + k/3 - 1, x + k%3 - 1, :]
// with zeros outside the image.  The mask (576 logits per coarse pixel) is the traffic: 2304 B in, 256 B out per
// coarse pixel -- HBM-bound.  A lane owns one coarse pixel, so every mask load of a warp is one 128-byte line (all 72
// loads of a thread are issued before the first softmax); the outputs are staged in shared memory and leave as
// contiguous runs of the fine rows.
#include <cuda_fp16.h>

#include "common.cuh"

namespace droid {
namespace {

template <typename M>
__device__ __forceinline__ float mload(const M* p);
template <>
__device__ __forceinline__ float mload<float>(const float* p) { return __ldcs(p); }
template <>
__device__ __forceinline__ float mload<__half>(const __half* p) { return __half2float(__ldcs(p)); }

// softmax over the 9 logits; a half mask is rounded back to half like torch.softmax on a half tensor
template <typename M>
__device__ __forceinline__ void softmax9(float (&l)[9]) {
  float mx = l[0];
#pragma unroll
  for (int k = 1; k < 9; k++) mx = fmaxf(mx, l[k]);
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 9; k++) { l[k] = expf(l[k] - mx); s += l[k]; }
#pragma unroll
  const float inv = 1.0f / s;
#pragma unroll
  for (int k = 0; k < 9; k++) {
    l[k] = l[k] * inv;
    if (sizeof(M) == 2) l[k] = __half2float(__float2half_rn(l[k]));
  }
}

constexpr int kUpThreads = 128;

// dim == 1 (inverse depth).  CTA = 32 consecutive coarse pixels x 8 fine rows: warp i computes the outputs (i, 0..7)
// of the 32 pixels, i.e. a contiguous run of 256 floats of fine row 8y+i, stages them and writes them as such.
template <typename M>
__global__ void __launch_bounds__(256) cvx_upsample1_kernel(const float* __restrict__ data, const M* __restrict__ mask,
                                                            float* __restrict__ out, int ht, int wd) {
  __shared__ float stage[8][32 * 8 + 8];
  const int b = blockIdx.y, hw = ht * wd;
  const int lane = threadIdx.x & 31, i = threadIdx.x >> 5;
  const int p0 = blockIdx.x * 32;
  const int p = p0 + lane;
  const bool ok = p < hw;
  const int y = ok ? p / wd : 0, x = ok ? p - y * wd : 0;
  float nb[9];
#pragma unroll
  for (int k = 0; k < 9; k++) {
    const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
    nb[k] = (ok && yy >= 0 && yy < ht && xx >= 0 && xx < wd) ? __ldg(data + (size_t)b * hw + yy * wd + xx) : 0.f;
  }
  const M* mp = mask + (size_t)b * 576 * hw + p;
  float l[8][9];
#pragma unroll
  for (int j = 0; j < 8; j++)
#pragma unroll
    for (int k = 0; k < 9; k++) l[j][k] = ok ? mload<M>(mp + (size_t)(k * 64 + i * 8 + j) * hw) : 0.f;
#pragma unroll
  for (int j = 0; j < 8; j++) {
    softmax9<M>(l[j]);
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 9; k++) v += l[j][k] * nb[k];              // the order of torch.sum over dim 2
    stage[i][lane * 8 + j] = v;
  }
  __syncwarp();
  const size_t fw = (size_t)8 * wd;
  for (int e = lane; e < 256; e += 32) {
    const int q = p0 + (e >> 3);
    if (q >= hw) continue;
    const int qy = q / wd, qx = q - qy * wd;
    out[((size_t)b * 8 * ht + 8 * qy + i) * fw + 8 * qx + (e & 7)] = stage[i][e];
  }
}

// any dim: weights once per output position, neighbours through L1, direct stores
template <typename M>
__global__ void __launch_bounds__(kUpThreads) cvx_upsample_kernel(const float* __restrict__ data, const M* __restrict__ mask,
                                                                  float* __restrict__ out, int ht, int wd, int dim) {
  const int b = blockIdx.y, hw = ht * wd;
  const int p = blockIdx.x * kUpThreads + threadIdx.x;
  if (p >= hw) return;
  const int y = p / wd, x = p - y * wd;
  const M* mp = mask + (size_t)b * 576 * hw + p;
  const size_t fw = (size_t)8 * wd;
  for (int i = 0; i < 8; i++) {
    for (int j = 0; j < 8; j++) {
      float l[9];
#pragma unroll
      for (int k = 0; k < 9; k++) l[k] = mload<M>(mp + (size_t)(k * 64 + i * 8 + j) * hw);
      softmax9<M>(l);
      float* o = out + (((size_t)b * 8 * ht + 8 * y + i) * fw + 8 * x + j) * dim;
      for (int d = 0; d < dim; d++) {
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < 9; k++) {
          const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
          const float nbv = (yy >= 0 && yy < ht && xx >= 0 && xx < wd) ? __ldg(data + ((size_t)b * hw + yy * wd + xx) * dim + d) : 0.f;
          v += l[k] * nbv;
        }
        o[d] = v;
      }
    }
  }
}

}  // namespace
}  // namespace droid

using namespace droid;

// mask_dtype: 0 = float32, 1 = float16
extern "C" int droid_cvx_upsample(const float* data, const void* mask, float* out, int mask_dtype, int batch, int ht,
                                  int wd, int dim, void* stream) {
  if (batch < 0 || ht <= 0 || wd <= 0 || dim <= 0 || (mask_dtype != 0 && mask_dtype != 1)) return DROID_ERR_INVALID_ARGUMENT;
  if (batch == 0) return DROID_OK;
  if (!data || !mask || !out || batch > 65535) return DROID_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  const int hw = ht * wd;
  const dim3 grid((hw + kUpThreads - 1) / kUpThreads, batch);
  if (dim == 1) {
    const dim3 g1((hw + 31) / 32, batch);
    if (mask_dtype == 0) DROID_LAUNCH cvx_upsample1_kernel<float><<<g1, 256, 0, st>>>(data, (const float*)mask, out, ht, wd);
    else DROID_LAUNCH cvx_upsample1_kernel<__half><<<g1, 256, 0, st>>>(data, (const __half*)mask, out, ht, wd);
  } else {
    if (mask_dtype == 0) DROID_LAUNCH cvx_upsample_kernel<float><<<grid, kUpThreads, 0, st>>>(data, (const float*)mask, out, ht, wd, dim);
    else DROID_LAUNCH cvx_upsample_kernel<__half><<<grid, kUpThreads, 0, st>>>(data, (const __half*)mask, out, ht, wd, dim);
  }
  return check_launch();
}
