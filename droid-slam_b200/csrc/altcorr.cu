// altcorr_forward / altcorr_backward / indexed alt-correlation.
// Reference semantics: src/altcorr_kernel.cu:27-286 (1-warp CTAs, 32x32 shared tiles, 64 gathered
// tile loads per 32-channel chunk, global read-modify-write of the outputs).
//
// B200 design: one warp per (edge, sample, pixel).  The C channels are split over the 32 lanes as
// 16-byte vectors (C=128 -> one float4 per lane), so every tap is ONE fully coalesced 512-byte row read
// of the channel-last fmap2 (L1/L2 resident between neighbouring pixels); the (2r+2)^2 per-lane partial
// dot products are reduced with a transposing butterfly (62 shuffles for 64 taps instead of 320), the
// bilinear combination uses the reference's tap order, and outputs are staged through shared memory so
// each channel is written as one 32-byte sector per 8-pixel strip.
#include <cuda_fp16.h>
#include "common.cuh"

namespace droid {

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4(const __half* p) {
  uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
  const __half2 a = *reinterpret_cast<__half2*>(&r.x), b = *reinterpret_cast<__half2*>(&r.y);
  const float2 fa = __half22float2(a), fb = __half22float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}

// Transposing butterfly: every lane enters with NT per-lane partials v[0..NT) and leaves with the
// full sums of taps  (NT/32)*lane + {0..NT/32-1}  in v[0..NT/32).
template <int NT>
__device__ __forceinline__ void transpose_reduce(float* v, int lane) {
  int cnt = NT;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const int half = cnt >> 1;
    const bool up = lane & off;
#pragma unroll
    for (int i = 0; i < NT / 2; i++) {
      if (i < half) {
        const float send = up ? v[i] : v[i + half];
        const float keep = up ? v[i + half] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
    }
    cnt = half;
  }
}

// radius-3 forward.  TF = element type of the feature maps (float for the drop-in op, __half for the
// indexed pyramid op which converts on the fly exactly like `.float()` in modules/corr.py:120).
// CTA = 8 warps = a strip of 8 consecutive pixels of one (b, n).
template <typename TF, int CPL /*channels per lane, multiple of 4*/>
__global__ void __launch_bounds__(256) altcorr_fwd_r3_kernel(
    const TF* __restrict__ f1_base, const TF* __restrict__ f2_base, const int64_t* __restrict__ ii,
    const int64_t* __restrict__ jj, const float* __restrict__ coords, float* __restrict__ out, int N,
    int HW1, int H2, int W2, int C, const int* __restrict__ only_if, int nstrips, int B) {
  // Gated launch (`only_if` set; behind the tensor-core kernel of altcorr_tc.cu): a small 1-D grid whose CTAs walk over
  // all (b, n, strip) units -- when the flag is clear, which is the normal case, a few hundred CTAs exit instead of
  // tens of thousands (45 us per launch at 160 edges).  Plain launch: one unit per CTA, 3-D grid.
  if (only_if && *only_if == 0) return;      // the tensor-core kernel has done the work
  __shared__ float taps[8][64];
  __shared__ float stage[49][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long total = only_if ? (long long)nstrips * N * B : 1;
  for (long long unit = only_if ? blockIdx.x : 0; unit < total; unit += gridDim.x) {
  const int b = only_if ? (int)(unit / ((long long)nstrips * N)) : blockIdx.z;
  const int n = only_if ? (int)((unit / nstrips) % N) : blockIdx.y;
  const int pix0 = (only_if ? (int)(unit % nstrips) : blockIdx.x) * 8;
  const int pix = pix0 + warp;
  const bool pv = pix < HW1;
  const size_t i1 = ii ? (size_t)ii[b] : (size_t)b;
  const size_t i2 = jj ? (size_t)jj[b] : (size_t)b;
  float dx = 0.f, dy = 0.f;
  if (pv) {
    const float* cp = coords + (((size_t)b * N + n) * HW1 + pix) * 2;
    const float x0 = cp[0], y0 = cp[1];
    const float fx0 = floorf(x0), fy0 = floorf(y0);
    dx = x0 - fx0;
    dy = y0 - fy0;
    const int xs = window_origin(fx0, 3), ys = window_origin(fy0, 3);
    const TF* f1 = f1_base + (i1 * HW1 + pix) * (size_t)C + lane * CPL;
    const TF* f2 = f2_base + i2 * ((size_t)H2 * W2) * C + lane * CPL;
    float4 a[CPL / 4];
#pragma unroll
    for (int k = 0; k < CPL / 4; k++) a[k] = ld4(f1 + 4 * k);
    float v[64];
#pragma unroll
    for (int iy = 0; iy < 8; iy++) {
#pragma unroll
      for (int ix = 0; ix < 8; ix++) {
        const int h2 = ys + iy, w2 = xs + ix;
        float s = 0.f;
        if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
          const TF* p = f2 + ((size_t)h2 * W2 + w2) * C;
#pragma unroll
          for (int k = 0; k < CPL / 4; k++) {
            const float4 f = ld4(p + 4 * k);
            s = fmaf(a[k].x, f.x, s);
            s = fmaf(a[k].y, f.y, s);
            s = fmaf(a[k].z, f.z, s);
            s = fmaf(a[k].w, f.w, s);
          }
        }
        v[iy * 8 + ix] = s;
      }
    }
    transpose_reduce<64>(v, lane);
    taps[warp][2 * lane] = v[0];
    taps[warp][2 * lane + 1] = v[1];
  }
  __syncwarp();
  if (pv) {
    const float wse = (1 - dy) * (1 - dx), wsw = (1 - dy) * dx, wne = dy * (1 - dx), wnw = dy * dx;
    for (int o = lane; o < 49; o += 32) {
      const int ixo = o / 7, iyo = o - ixo * 7;  // channel = iy + 7*ix  (src/altcorr_kernel.cu:102-105)
      const float* t = &taps[warp][iyo * 8 + ixo];
      float acc = t[0] * wse;       // tap (iy  , ix  )
      acc = fmaf(t[1], wsw, acc);   // tap (iy  , ix+1)
      acc = fmaf(t[8], wne, acc);   // tap (iy+1, ix  )
      acc = fmaf(t[9], wnw, acc);   // tap (iy+1, ix+1)
      stage[o][warp] = acc;
    }
  }
  __syncthreads();
  float* ob = out + ((size_t)b * N + n) * 49 * HW1;
  for (int e = threadIdx.x; e < 49 * 8; e += 256) {
    const int c = e >> 3, p = e & 7;
    if (pix0 + p < HW1) ob[(size_t)c * HW1 + pix0 + p] = stage[c][p];
  }
  __syncthreads();      // `stage` / `taps` are reused by the next unit
  }
}

// Generic fallback (any radius, any C): one warp per pixel, lanes stride over channels.
__global__ void __launch_bounds__(256) altcorr_fwd_generic_kernel(
    const float* __restrict__ f1_base, const float* __restrict__ f2_base,
    const float* __restrict__ coords, float* __restrict__ out, int N, int HW1, int H2, int W2, int C,
    int r) {
  extern __shared__ float sm[];
  const int rd = 2 * r + 1, rt = rd + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* taps = sm + (size_t)warp * rt * rt;
  const int b = blockIdx.z, n = blockIdx.y;
  const int pix = blockIdx.x * 8 + warp;
  if (pix >= HW1) return;
  const float* cp = coords + (((size_t)b * N + n) * HW1 + pix) * 2;
  const float x0 = cp[0], y0 = cp[1];
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int xs = window_origin(fx0, r), ys = window_origin(fy0, r);
  const float* f1 = f1_base + ((size_t)b * HW1 + pix) * (size_t)C;
  const float* f2 = f2_base + (size_t)b * ((size_t)H2 * W2) * C;
  for (int iy = 0; iy < rt; iy++)
    for (int ix = 0; ix < rt; ix++) {
      const int h2 = ys + iy, w2 = xs + ix;
      float s = 0.f;
      if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
        const float* p = f2 + ((size_t)h2 * W2 + w2) * C;
        for (int c = lane; c < C; c += 32) s = fmaf(f1[c], p[c], s);
      }
      s = warp_sum(s);
      if (lane == 0) taps[iy * rt + ix] = s;
    }
  __syncwarp();
  const float wse = (1 - dy) * (1 - dx), wsw = (1 - dy) * dx, wne = dy * (1 - dx), wnw = dy * dx;
  float* ob = out + ((size_t)b * N + n) * rd * rd * HW1;
  for (int o = lane; o < rd * rd; o += 32) {
    const int ixo = o / rd, iyo = o - ixo * rd;
    const float* t = &taps[iyo * rt + ixo];
    float acc = t[0] * wse;
    acc = fmaf(t[1], wsw, acc);
    acc = fmaf(t[rt], wne, acc);
    acc = fmaf(t[rt + 1], wnw, acc);
    ob[(size_t)o * HW1 + pix] = acc;
  }
}

// Backward (f32 only, like src/altcorr_kernel.cu:345).  One warp per (b, n, pixel); lanes stride over
// channels.  fmap1_grad: each pixel is owned by one warp per sample n -> accumulated over n with
// atomics only when N > 1; fmap2_grad: atomics (as the reference).  Both grads must be zero on entry.
__global__ void __launch_bounds__(256) altcorr_bwd_kernel(
    const float* __restrict__ f1_base, const float* __restrict__ f2_base,
    const float* __restrict__ coords, const float* __restrict__ cg, float* __restrict__ f1g_base,
    float* __restrict__ f2g_base, int N, int HW1, int H2, int W2, int C, int r) {
  extern __shared__ float sm[];
  const int rd = 2 * r + 1, rt = rd + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* g = sm + (size_t)warp * rt * rt;
  const int b = blockIdx.z, n = blockIdx.y;
  const int pix = blockIdx.x * 8 + warp;
  if (pix >= HW1) return;
  const float* cp = coords + (((size_t)b * N + n) * HW1 + pix) * 2;
  const float x0 = cp[0], y0 = cp[1];
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int xs = window_origin(fx0, r), ys = window_origin(fy0, r);
  const float* gp = cg + ((size_t)b * N + n) * rd * rd * HW1 + pix;
  // per-tap gradient, reference order (:242-252): nw, ne, sw, se
  for (int t = lane; t < rt * rt; t += 32) {
    const int iy = t / rt, ix = t - iy * rt;
    float a = 0.f;
    if (iy > 0 && ix > 0) a += gp[(size_t)((iy - 1) + rd * (ix - 1)) * HW1] * dy * dx;
    if (iy > 0 && ix < rd) a += gp[(size_t)((iy - 1) + rd * ix) * HW1] * dy * (1 - dx);
    if (iy < rd && ix > 0) a += gp[(size_t)(iy + rd * (ix - 1)) * HW1] * (1 - dy) * dx;
    if (iy < rd && ix < rd) a += gp[(size_t)(iy + rd * ix) * HW1] * (1 - dy) * (1 - dx);
    g[t] = a;
  }
  __syncwarp();
  const float* f1 = f1_base + ((size_t)b * HW1 + pix) * (size_t)C;
  float* f1g = f1g_base + ((size_t)b * HW1 + pix) * (size_t)C;
  const size_t plane = (size_t)b * ((size_t)H2 * W2) * C;
  for (int c = lane; c < C; c += 32) {
    const float a1 = f1[c];
    float acc = 0.f;
    for (int iy = 0; iy < rt; iy++)
      for (int ix = 0; ix < rt; ix++) {
        const int h2 = ys + iy, w2 = xs + ix;
        if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
          const size_t off = plane + ((size_t)h2 * W2 + w2) * C + c;
          const float gt = g[iy * rt + ix];
          acc = fmaf(gt, f2_base[off], acc);
          atomicAdd(f2g_base + off, gt * a1);
        }
      }
    if (N > 1) atomicAdd(f1g + c, acc);
    else f1g[c] = acc;
  }
}

// altcorr_tc.cu
bool altcorr_tc_supported(int B, int N, int H1, int W1, int H2, int W2, int C, int radius, size_t frames1, size_t frames2);
int altcorr_tc_forward(const void* f1, const void* f2, const int64_t* ii, const int64_t* jj, const float* coords, float* out,
                       int* skip_if, size_t frames1, size_t frames2, int B, int N, int H1, int W1, int H2, int W2,
                       int C, int radius, cudaStream_t st, bool f1_is_f32);
int altcorr_tc_forward_levels(const void* f1, const void* const* f2, const int* H2, const int* W2, const float* cscale,
                              int nlev, const int64_t* ii, const int64_t* jj, const float* coords, float* out,
                              int* skip_if, size_t frames1, size_t frames2, int B, int N, int H1, int W1, int C,
                              int radius, cudaStream_t st, bool f1_is_f32);
int altcorr_tc_convert(const float* src, void* dst, int* flag, size_t n, cudaStream_t st);

static bool tc_enabled() {
  static const bool on = [] {
    const char* e = getenv("DROID_ALTCORR_TC");
    return !(e && e[0] == '0');
  }();
  return on;
}

// `only_if` (may be null): device flag; the CUDA-core kernel runs only when it is non-zero.
static int fwd_common(const void* f1, const void* f2, const int64_t* ii, const int64_t* jj,
                      const float* coords, float* out, int dtype, int B, int N, int H1, int W1,
                      int H2, int W2, int C, int radius, cudaStream_t st, const int* only_if = nullptr) {
  const int HW1 = H1 * W1;
  if (N > 65535 || B > 65535) return DROID_ERR_INVALID_ARGUMENT;
  dim3 grid((HW1 + 7) / 8, N, B);
  const int esz = dtype == DROID_DTYPE_F16 ? 2 : 4;
  const bool aligned = ((uintptr_t)f1 % (4 * esz) == 0) && ((uintptr_t)f2 % (4 * esz) == 0);
  if (radius == 3 && aligned && (C == 128 || C == 256 || (C == 384 && dtype == DROID_DTYPE_F32))) {
    const int nstrips = (HW1 + 7) / 8;
    const dim3 g = only_if ? dim3((unsigned)std::min<long long>((long long)nstrips * N * B, 148 * 8)) : grid;
#define ALT_LAUNCH(TF, CPL)                                                                     \
  DROID_LAUNCH altcorr_fwd_r3_kernel<TF, CPL><<<g, 256, 0, st>>>(reinterpret_cast<const TF*>(f1),          \
                                                        reinterpret_cast<const TF*>(f2), ii, jj, \
                                                        coords, out, N, HW1, H2, W2, C, only_if, nstrips, B)
    if (dtype == DROID_DTYPE_F32) {
      if (C == 128) ALT_LAUNCH(float, 4);
      else if (C == 256) ALT_LAUNCH(float, 8);
      else if (C == 384) ALT_LAUNCH(float, 12);
      else goto generic;
    } else if (dtype == DROID_DTYPE_F16) {
      if (C == 128) ALT_LAUNCH(__half, 4);
      else if (C == 256) ALT_LAUNCH(__half, 8);
      else return DROID_ERR_INVALID_ARGUMENT;
    } else {
      return DROID_ERR_UNSUPPORTED_DTYPE;
    }
#undef ALT_LAUNCH
    return check_launch();
  }
generic:
  if (dtype != DROID_DTYPE_F32 || ii || jj || only_if) return DROID_ERR_UNSUPPORTED_DTYPE;
  {
    const int rt = 2 * radius + 2;
    size_t smem = (size_t)8 * rt * rt * sizeof(float);
    if (smem > 48 * 1024) return DROID_ERR_INVALID_ARGUMENT;
    DROID_LAUNCH altcorr_fwd_generic_kernel<<<grid, 256, smem, st>>>(reinterpret_cast<const float*>(f1),
                                                        reinterpret_cast<const float*>(f2), coords,
                                                        out, N, HW1, H2, W2, C, radius);
  }
  return check_launch();
}

}  // namespace droid

using namespace droid;

extern "C" int droid_altcorr_forward(const void* fmap1, const void* fmap2, const float* coords,
                                     void* out, int dtype, int B, int N, int H1, int W1, int H2,
                                     int W2, int C, int radius, void* stream) {
  if (B < 0 || N < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0 || radius < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0 || N == 0) return DROID_OK;
  if (!fmap1 || !fmap2 || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  if (dtype != DROID_DTYPE_F32) return DROID_ERR_UNSUPPORTED_DTYPE;
  return fwd_common(fmap1, fmap2, nullptr, nullptr, coords, reinterpret_cast<float*>(out), dtype, B,
                    N, H1, W1, H2, W2, C, radius, (cudaStream_t)stream);
}

// Workspace of droid_altcorr_forward_ws: an fp16 copy of the target map + the exactness flag; 0 = shape not covered by the
// tensor-core kernel (the call then behaves exactly like droid_altcorr_forward).
extern "C" size_t droid_altcorr_forward_workspace_bytes(int B, int N, int H1, int W1, int H2, int W2, int C, int radius) {
  if (B <= 0 || N <= 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0) return 0;
  if (!tc_enabled() || !altcorr_tc_supported(B, N, H1, W1, H2, W2, C, radius, (size_t)B, (size_t)B)) return 0;
  const size_t n2 = (size_t)B * H2 * W2 * C;
  return 256 + ((n2 * 2 + 255) & ~(size_t)255) + 256;
}

extern "C" int droid_altcorr_forward_ws(const void* fmap1, const void* fmap2, const float* coords, void* out,
                                        int dtype, int B, int N, int H1, int W1, int H2, int W2, int C, int radius,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  if (B < 0 || N < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0 || radius < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0 || N == 0) return DROID_OK;
  if (!fmap1 || !fmap2 || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  if (dtype != DROID_DTYPE_F32) return DROID_ERR_UNSUPPORTED_DTYPE;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t need = droid_altcorr_forward_workspace_bytes(B, N, H1, W1, H2, W2, C, radius);
  const bool aligned = (((uintptr_t)fmap1 | (uintptr_t)fmap2 | (uintptr_t)workspace) & 15) == 0;
  if (need && workspace && workspace_bytes >= need && aligned) {
    const size_t n2 = (size_t)B * H2 * W2 * C;
    unsigned char* w = reinterpret_cast<unsigned char*>(workspace);
    int* flag = reinterpret_cast<int*>(w);
    void* f2h = w + 256;
    // the target map is copied to fp16 here (every CTA of an edge re-reads it); the source map is converted tile by
    // tile inside the kernel.  Both raise the flag on a value that is not an fp16 number.
    int rc = altcorr_tc_convert(reinterpret_cast<const float*>(fmap2), f2h, flag, n2, st);
    if (rc != DROID_OK) return rc;
    rc = altcorr_tc_forward(fmap1, f2h, nullptr, nullptr, coords, reinterpret_cast<float*>(out), flag, (size_t)B, (size_t)B,
                            B, N, H1, W1, H2, W2, C, radius, st, true);
    if (rc == DROID_OK)      // behind it, gated on the flag: the CUDA-core kernel for inputs that are not fp16 values
      return fwd_common(fmap1, fmap2, nullptr, nullptr, coords, reinterpret_cast<float*>(out), dtype, B, N, H1, W1, H2,
                        W2, C, radius, st, flag);
    if (rc != DROID_ERR_NOT_IMPLEMENTED) return rc;
  }
  return fwd_common(fmap1, fmap2, nullptr, nullptr, coords, reinterpret_cast<float*>(out), dtype, B, N, H1, W1, H2, W2,
                    C, radius, st);
}

extern "C" int droid_altcorr_forward_indexed(const void* pyr0, const void* pyrL, const int64_t* ii,
                                             const int64_t* jj, const float* coords, float* out,
                                             int dtype, int F, int B, int N, int H1, int W1, int H2,
                                             int W2, int C, int radius, void* stream) {
  if (F <= 0 || B < 0 || N < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0 || radius != 3)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0 || N == 0) return DROID_OK;
  if (!pyr0 || !pyrL || !ii || !jj || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  if (C != 128 && C != 256) return DROID_ERR_INVALID_ARGUMENT;
  if (dtype == DROID_DTYPE_F16 && tc_enabled()) {
    const int rc = altcorr_tc_forward(pyr0, pyrL, ii, jj, coords, out, nullptr, (size_t)F, (size_t)F, B, N, H1, W1, H2,
                                      W2, C, radius, (cudaStream_t)stream, false);
    if (rc != DROID_ERR_NOT_IMPLEMENTED) return rc;
  }
  return fwd_common(pyr0, pyrL, ii, jj, coords, out, dtype, B, N, H1, W1, H2, W2, C, radius,
                    (cudaStream_t)stream);
}

extern "C" int droid_altcorr_pyramid_indexed(const void* const* pyr_host, int num_levels, const int64_t* ii,
                                            const int64_t* jj, const float* coords, float* out, int dtype, int F,
                                            int B, int N, int H1, int W1, int C, int radius, void* stream) {
  if (F <= 0 || B < 0 || N < 0 || H1 <= 0 || W1 <= 0 || C <= 0 || radius != 3 || num_levels < 1 || num_levels > 4 ||
      !pyr_host)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0 || N == 0) return DROID_OK;
  if (!ii || !jj || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  int H2[4], W2[4];
  float cs[4];
  for (int l = 0; l < num_levels; l++) {
    if (!pyr_host[l]) return DROID_ERR_INVALID_ARGUMENT;
    H2[l] = H1 >> l;
    W2[l] = W1 >> l;
    cs[l] = 1.0f / (float)(1 << l);
    if (H2[l] <= 0 || W2[l] <= 0) return DROID_ERR_INVALID_ARGUMENT;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DROID_DTYPE_F16 && tc_enabled()) {
    const int rc = altcorr_tc_forward_levels(pyr_host[0], pyr_host, H2, W2, cs, num_levels, ii, jj, coords, out, nullptr,
                                             (size_t)F, (size_t)F, B, N, H1, W1, C, radius, st, false);
    if (rc != DROID_ERR_NOT_IMPLEMENTED) return rc;
  }
  return DROID_ERR_NOT_IMPLEMENTED;   // the Python host then loops over droid_altcorr_forward_indexed
}

extern "C" int droid_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords,
                                      const float* corr_grad, float* fmap1_grad, float* fmap2_grad,
                                      float* coords_grad, int B, int N, int H1, int W1, int H2,
                                      int W2, int C, int radius, void* stream) {
  if (B < 0 || N < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0 || radius < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0 || N == 0) return DROID_OK;
  if (!fmap1 || !fmap2 || !coords || !corr_grad || !fmap1_grad || !fmap2_grad || !coords_grad)
    return DROID_ERR_INVALID_ARGUMENT;
  if (N > 65535 || B > 65535) return DROID_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t HW1 = (size_t)H1 * W1;
  cudaMemsetAsync(fmap1_grad, 0, (size_t)B * HW1 * C * sizeof(float), st);
  cudaMemsetAsync(fmap2_grad, 0, (size_t)B * H2 * W2 * C * sizeof(float), st);
  cudaMemsetAsync(coords_grad, 0, (size_t)B * N * HW1 * 2 * sizeof(float), st);
  const int rt = 2 * radius + 2;
  size_t smem = (size_t)8 * rt * rt * sizeof(float);
  if (smem > 48 * 1024) return DROID_ERR_INVALID_ARGUMENT;
  dim3 grid(((int)HW1 + 7) / 8, N, B);
  DROID_LAUNCH altcorr_bwd_kernel<<<grid, 256, smem, st>>>(fmap1, fmap2, coords, corr_grad, fmap1_grad,
                                              fmap2_grad, N, (int)HW1, H2, W2, C, radius);
  return check_launch();
}
