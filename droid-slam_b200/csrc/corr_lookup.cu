// corr_index_forward / corr_index_backward / fused 4-level pyramid lookup.
// Reference semantics: src/correlation_kernels.cu:19-124 (one thread per pixel scanning the
// (2r+2)^2 integer window and read-modify-writing the 4 neighbouring outputs in global memory).
//
// B200 design (gather-bound, HBM/L2): 8 lanes cooperate on one pixel, one lane per window row.
// A lane fetches its row's 8 taps with two (f16) / three (f32) ALIGNED 16-byte loads and realigns them
// in registers, gets the row below with warp shuffles, and produces the 7 outputs of its row with the
// reference's own operation order and per-step rounding (for f16 every product and partial sum is
// rounded to half exactly like the reference's `corr[...] += s * scalar_t(w)` on c10::Half).  Results
// are transposed through shared memory so the [rd,rd,H1,W1] output is written in full 32-pixel lines.
#include <cuda_fp16.h>
#include "common.cuh"

namespace droid {

template <typename T>
struct Arith;
template <>
struct Arith<float> {
  static __device__ __forceinline__ float zero() { return 0.f; }
  static __device__ __forceinline__ float wcast(float w) { return w; }
  // corr += s * w  (nvcc contracts this to one FFMA in the reference build as well)
  static __device__ __forceinline__ float mad(float acc, float s, float w) { return fmaf(s, w, acc); }
};
template <>
struct Arith<double> {
  static __device__ __forceinline__ double zero() { return 0.0; }
  static __device__ __forceinline__ double wcast(float w) { return (double)w; }
  static __device__ __forceinline__ double mad(double acc, double s, double w) { return fma(s, w, acc); }
};
template <>
struct Arith<__half> {
  static __device__ __forceinline__ __half zero() { return __float2half_rn(0.f); }
  static __device__ __forceinline__ __half wcast(float w) { return __float2half_rn(w); }
  // c10::Half: product in float (exact) rounded to half, sum in float rounded to half.
  static __device__ __forceinline__ __half mad(__half acc, __half s, __half w) {
    __half p = __float2half_rn(__half2float(s) * __half2float(w));
    return __float2half_rn(__half2float(acc) + __half2float(p));
  }
};

// ---- row fetch: 8 taps starting at element xs of a row of W2 elements ---------------------------
template <typename T>
__device__ __forceinline__ void load_row_scalar(const T* __restrict__ row, int xs, int W2, T* t) {
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int x = xs + k;
    t[k] = (x >= 0 && x < W2) ? row[x] : Arith<T>::zero();
  }
}

__device__ __forceinline__ uint4 ldg16(const void* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

// f16: taps live in two aligned 8-element chunks; realign with funnel shifts.
__device__ __forceinline__ void load_row_vec(const __half* __restrict__ row, int xs, int W2,
                                             __half* t) {
  const int a0 = (xs >> 3) << 3;  // floor to a multiple of 8 (arithmetic shift handles xs < 0)
  const int o = xs - a0;          // 0..7
  uint32_t w[8];
  uint4 c0 = make_uint4(0, 0, 0, 0), c1 = make_uint4(0, 0, 0, 0);
  if (a0 >= 0 && a0 < W2) c0 = ldg16(row + a0);
  if (a0 + 8 >= 0 && a0 + 8 < W2) c1 = ldg16(row + a0 + 8);
  w[0] = c0.x; w[1] = c0.y; w[2] = c0.z; w[3] = c0.w;
  w[4] = c1.x; w[5] = c1.y; w[6] = c1.z; w[7] = c1.w;
  const int idx = o >> 1;
  const uint32_t sh = (o & 1) * 16;
  uint32_t s[5];
#pragma unroll
  for (int k = 0; k < 5; k++) {
    uint32_t v = w[k];
    v = idx == 1 ? w[k + 1] : v;
    v = idx == 2 ? w[k + 2] : v;
    v = idx == 3 ? (k + 3 < 8 ? w[k + 3] : 0u) : v;
    s[k] = v;
  }
  uint32_t r[4];
#pragma unroll
  for (int k = 0; k < 4; k++) r[k] = __funnelshift_r(s[k], s[k + 1], sh);
  const __half2* h = reinterpret_cast<const __half2*>(r);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    t[2 * k] = __low2half(h[k]);
    t[2 * k + 1] = __high2half(h[k]);
  }
}

// f32: taps live in three aligned 4-element chunks.
__device__ __forceinline__ void load_row_vec(const float* __restrict__ row, int xs, int W2, float* t) {
  const int a0 = (xs >> 2) << 2;
  const int o = xs - a0;  // 0..3
  float w[12];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    uint4 v = make_uint4(0, 0, 0, 0);
    const int a = a0 + 4 * c;
    if (a >= 0 && a < W2) v = ldg16(row + a);
    w[4 * c + 0] = __uint_as_float(v.x);
    w[4 * c + 1] = __uint_as_float(v.y);
    w[4 * c + 2] = __uint_as_float(v.z);
    w[4 * c + 3] = __uint_as_float(v.w);
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    float v = w[k];
    v = o == 1 ? w[k + 1] : v;
    v = o == 2 ? w[k + 2] : v;
    v = o == 3 ? w[k + 3] : v;
    t[k] = v;
  }
}
__device__ __forceinline__ void load_row_vec(const double* __restrict__ row, int xs, int W2,
                                             double* t) {
  load_row_scalar(row, xs, W2, t);
}

template <typename T>
__device__ __forceinline__ T shfl_down1(T v);
template <>
__device__ __forceinline__ float shfl_down1(float v) { return __shfl_down_sync(0xffffffffu, v, 1); }
template <>
__device__ __forceinline__ double shfl_down1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }
template <>
__device__ __forceinline__ __half shfl_down1(__half v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// One level of the radius-3 lookup for the 32-pixel tile of this CTA.  256 threads: 32 pixels x 8 rows.
//   vol   : [B,HW1,H2,W2] of this level;  scale : 1/2^level applied to the level-0 coords
//   out   : channel c = i*7 + j of this level goes to out[(b*CH + ch0 + c)*HW1 + pix]
// what one thread (pixel p = tid >> 3, window row = tid & 7) holds of one level between the load and the arithmetic
template <typename T>
struct RowTaps {
  T t[8];        // taps (xs .. xs+7, ys), zero outside the plane
  unsigned m;    // bit k: tap (xs+k, ys) is inside the plane
  float dx, dy;  // bilinear fractions
};

// load phase: x0, y0 = the pixel's coordinates at this level (already scaled); pv = pixel inside the image
template <typename T>
__device__ __forceinline__ void lookup_load_r3(const T* __restrict__ vol, int b, int pix, bool pv, float x0, float y0,
                                               int HW1, int H2, int W2, bool vec, RowTaps<T>& R) {
  const int row = threadIdx.x & 7;
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  R.dx = x0 - fx0; R.dy = y0 - fy0;
  const int xs = window_origin(fx0, 3);
  const int ys = window_origin(fy0, 3) + row;
  const bool rv = pv && ys >= 0 && ys < H2;
  if (rv) {
    const T* rowp = vol + ((size_t)b * HW1 + pix) * ((size_t)H2 * W2) + (size_t)ys * W2;
    if (vec) load_row_vec(rowp, xs, W2, R.t);
    else load_row_scalar(rowp, xs, W2, R.t);
  } else {
#pragma unroll
    for (int k = 0; k < 8; k++) R.t[k] = Arith<T>::zero();
  }
  R.m = 0;
  if (rv) {
#pragma unroll
    for (int k = 0; k < 8; k++) R.m |= ((xs + k >= 0 && xs + k < W2) ? 1u : 0u) << k;
  }
}

// arithmetic phase of one level: 49 channels x 32 pixels of this CTA's tile into the staging rows
template <typename T>
__device__ __forceinline__ void lookup_finish_r3(const RowTaps<T>& R, T (*stage)[33]) {
  const int tid = threadIdx.x;
  const int p = tid >> 3, row = tid & 7;
  const T* t = R.t;
  const unsigned m = R.m;
  const float dx = R.dx, dy = R.dy;
  if constexpr (sizeof(T) == 2) {
    // f16: native packed arithmetic.  HMUL2 rounds the (exact) product once and HADD2 the (exact)
    // sum once, which is bit-identical to c10::Half's float-op-then-round; out-of-plane taps were loaded as
    // zeros and adding +0 leaves the partial sum unchanged, so no per-tap predicates are needed.
    (void)m;
    const __half2 E0 = __halves2half2(t[0], t[1]), E1 = __halves2half2(t[2], t[3]);
    const __half2 E2 = __halves2half2(t[4], t[5]), E3 = __halves2half2(t[6], t[7]);
    const unsigned e0 = *reinterpret_cast<const unsigned*>(&E0), e1 = *reinterpret_cast<const unsigned*>(&E1);
    const unsigned e2 = *reinterpret_cast<const unsigned*>(&E2), e3 = *reinterpret_cast<const unsigned*>(&E3);
    const unsigned f0 = __shfl_down_sync(0xffffffffu, e0, 1), f1 = __shfl_down_sync(0xffffffffu, e1, 1);
    const unsigned f2 = __shfl_down_sync(0xffffffffu, e2, 1), f3 = __shfl_down_sync(0xffffffffu, e3, 1);
    auto H2 = [](unsigned u) { return *reinterpret_cast<const __half2*>(&u); };
    // odd-aligned pairs (t[2p+1], t[2p+2])
    const unsigned o0 = __byte_perm(e0, e1, 0x5432), o1 = __byte_perm(e1, e2, 0x5432);
    const unsigned o2 = __byte_perm(e2, e3, 0x5432), o3 = __byte_perm(e3, 0u, 0x5432);
    const unsigned g0 = __byte_perm(f0, f1, 0x5432), g1 = __byte_perm(f1, f2, 0x5432);
    const unsigned g2 = __byte_perm(f2, f3, 0x5432), g3 = __byte_perm(f3, 0u, 0x5432);
    const __half2 W00 = __float2half2_rn((1.0f - dx) * (1.0f - dy)), W01 = __float2half2_rn((1.0f - dx) * dy);
    const __half2 W10 = __float2half2_rn(dx * (1.0f - dy)), W11 = __float2half2_rn(dx * dy);
    auto quad = [&](unsigned ea, unsigned fa, unsigned oa, unsigned ga) {
      __half2 acc = __hmul2_rn(H2(ea), W00);
      acc = __hadd2_rn(acc, __hmul2_rn(H2(fa), W01));
      acc = __hadd2_rn(acc, __hmul2_rn(H2(oa), W10));
      acc = __hadd2_rn(acc, __hmul2_rn(H2(ga), W11));
      return acc;
    };
    if (row < 7) {
      const __half2 r0 = quad(e0, f0, o0, g0), r1 = quad(e1, f1, o1, g1);
      const __half2 r2 = quad(e2, f2, o2, g2), r3 = quad(e3, f3, o3, g3);
      stage[0 * 7 + row][p] = __low2half(r0); stage[1 * 7 + row][p] = __high2half(r0);
      stage[2 * 7 + row][p] = __low2half(r1); stage[3 * 7 + row][p] = __high2half(r1);
      stage[4 * 7 + row][p] = __low2half(r2); stage[5 * 7 + row][p] = __high2half(r2);
      stage[6 * 7 + row][p] = __low2half(r3);
    }
  } else {
  T t2[8];
#pragma unroll
  for (int k = 0; k < 8; k++) t2[k] = shfl_down1<T>(t[k]);
  const unsigned m2 = __shfl_down_sync(0xffffffffu, m, 1);
  const T w00 = Arith<T>::wcast((1.0f - dx) * (1.0f - dy));
  const T w01 = Arith<T>::wcast((1.0f - dx) * dy);
  const T w10 = Arith<T>::wcast(dx * (1.0f - dy));
  const T w11 = Arith<T>::wcast(dx * dy);
  if (row < 7) {
#pragma unroll
    for (int a = 0; a < 7; a++) {
      T acc = Arith<T>::zero();
      if (m & (1u << a)) acc = Arith<T>::mad(acc, t[a], w00);
      if (m2 & (1u << a)) acc = Arith<T>::mad(acc, t2[a], w01);
      if (m & (2u << a)) acc = Arith<T>::mad(acc, t[a + 1], w10);
      if (m2 & (2u << a)) acc = Arith<T>::mad(acc, t2[a + 1], w11);
      stage[a * 7 + row][p] = acc;
    }
  }
  }
}

// store phase: nch staged channels of the 32-pixel tile -> out[(b*CH + ch0 + c)*HW1 + pix], 64/128-byte rows per warp
template <typename T>
__device__ __forceinline__ void lookup_store_r3(T* __restrict__ out, int b, int pix0, int HW1, int CH, int ch0, int nch,
                                                T (*stage)[33]) {
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (pix0 + lane < HW1) {
    for (int c = warp; c < nch; c += 8)
      out[((size_t)b * CH + ch0 + c) * HW1 + pix0 + lane] = stage[c][lane];
  }
  __syncthreads();
}

// One level of the radius-3 lookup for the 32-pixel tile of this CTA.  256 threads: 32 pixels x 8 rows.
//   vol   : [B,HW1,H2,W2] of this level;  scale : 1/2^level applied to the level-0 coords
//   out   : channel c = i*7 + j of this level goes to out[(b*CH + ch0 + c)*HW1 + pix]
template <typename T, bool VEC>
__device__ __forceinline__ void lookup_tile_r3(const T* __restrict__ vol,
                                               const float* __restrict__ coords, T* __restrict__ out,
                                               int b, int pix0, int HW1, int H2, int W2, float scale,
                                               int CH, int ch0, T (*stage)[33]) {
  const int pix = pix0 + (threadIdx.x >> 3);
  const bool pv = pix < HW1;
  float x0 = 0.f, y0 = 0.f;
  if (pv) {
    x0 = coords[((size_t)b * 2 + 0) * HW1 + pix] * scale;
    y0 = coords[((size_t)b * 2 + 1) * HW1 + pix] * scale;
  }
  RowTaps<T> R;
  lookup_load_r3<T>(vol, b, pix, pv, x0, y0, HW1, H2, W2, VEC, R);
  lookup_finish_r3<T>(R, stage);
  lookup_store_r3<T>(out, b, pix0, HW1, CH, ch0, 49, stage);
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(256) corr_index_fwd_r3_kernel(const T* __restrict__ vol,
                                                                const float* __restrict__ coords,
                                                                T* __restrict__ out, int HW1, int H2,
                                                                int W2) {
  __shared__ T stage[49][33];
  lookup_tile_r3<T, VEC>(vol, coords, out, blockIdx.y, blockIdx.x * 32, HW1, H2, W2, 1.0f, 49, 0,
                         stage);
}

struct PyrArgs {
  const void* vol[4];
  int vec[4];
};

template <typename T>
__global__ void __launch_bounds__(256) corr_index_pyr_r3_kernel(PyrArgs a, int L,
                                                                const float* __restrict__ coords,
                                                                T* __restrict__ out, int HW1, int H2,
                                                                int W2) {
  __shared__ T stage[4 * 49][33];
  // The four levels read four unrelated planes: all their loads are issued before any arithmetic, so that a CTA pays
  // the DRAM latency once (coordinates -> rows) instead of once per level.
  const int b = blockIdx.y, pix0 = blockIdx.x * 32;
  const int pix = pix0 + (threadIdx.x >> 3);
  const bool pv = pix < HW1;
  float x0 = 0.f, y0 = 0.f;
  if (pv) {
    x0 = coords[((size_t)b * 2 + 0) * HW1 + pix];
    y0 = coords[((size_t)b * 2 + 1) * HW1 + pix];
  }
  RowTaps<T> R[4];
  float scale = 1.0f;
#pragma unroll
  for (int l = 0; l < 4; l++) {
    if (l < L)
      lookup_load_r3<T>(reinterpret_cast<const T*>(a.vol[l]), b, pix, pv, x0 * scale, y0 * scale, HW1, H2 >> l, W2 >> l,
                        a.vec[l] != 0, R[l]);
    scale *= 0.5f;
  }
#pragma unroll
  for (int l = 0; l < 4; l++)
    if (l < L) lookup_finish_r3<T>(R[l], stage + 49 * l);
  lookup_store_r3<T>(out, b, pix0, HW1, 49 * L, 0, 49 * L, stage);
}

// Generic radius (any r >= 0): one thread per pixel and output row, no staging.  Used only when
// radius != 3; DROID itself always uses 3 (droid_slam/modules/corr.py:24).
template <typename T>
__global__ void __launch_bounds__(256) corr_index_fwd_generic_kernel(
    const T* __restrict__ vol, const float* __restrict__ coords, T* __restrict__ out, int HW1,
    int H2, int W2, int r) {
  const int pix = blockIdx.x * 256 + threadIdx.x;
  const int b = blockIdx.y;
  if (pix >= HW1) return;
  const int rd = 2 * r + 1;
  const float x0 = coords[((size_t)b * 2 + 0) * HW1 + pix];
  const float y0 = coords[((size_t)b * 2 + 1) * HW1 + pix];
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int xs = window_origin(fx0, r), ys = window_origin(fy0, r);
  const T w00 = Arith<T>::wcast((1.0f - dx) * (1.0f - dy));
  const T w01 = Arith<T>::wcast((1.0f - dx) * dy);
  const T w10 = Arith<T>::wcast(dx * (1.0f - dy));
  const T w11 = Arith<T>::wcast(dx * dy);
  const T* plane = vol + ((size_t)b * HW1 + pix) * ((size_t)H2 * W2);
  for (int a = 0; a < rd; a++) {
    for (int bb = 0; bb < rd; bb++) {
      T acc = Arith<T>::zero();
      const int xa = xs + a, yb = ys + bb;
      const bool x_0 = xa >= 0 && xa < W2, x_1 = xa + 1 >= 0 && xa + 1 < W2;
      const bool y_0 = yb >= 0 && yb < H2, y_1 = yb + 1 >= 0 && yb + 1 < H2;
      if (x_0 && y_0) acc = Arith<T>::mad(acc, plane[(size_t)yb * W2 + xa], w00);
      if (x_0 && y_1) acc = Arith<T>::mad(acc, plane[(size_t)(yb + 1) * W2 + xa], w01);
      if (x_1 && y_0) acc = Arith<T>::mad(acc, plane[(size_t)yb * W2 + xa + 1], w10);
      if (x_1 && y_1) acc = Arith<T>::mad(acc, plane[(size_t)(yb + 1) * W2 + xa + 1], w11);
      out[((size_t)b * rd * rd + a * rd + bb) * HW1 + pix] = acc;
    }
  }
}

// ---- backward -----------------------------------------------------------------------------------
// volume_grad[b,pix,y1,x1] = sum of <=4 corr_grad entries (reference order, :107-120) inside the
// (2r+2)^2 window, 0 elsewhere.  One warp per pixel plane; the rd*rd grads of the pixel are staged in
// shared memory, the plane is written once with coalesced stores (no separate memset pass).
template <typename T>
__global__ void __launch_bounds__(256) corr_index_bwd_kernel(const float* __restrict__ coords,
                                                             const T* __restrict__ cg,
                                                             T* __restrict__ vg, int HW1, int H2,
                                                             int W2, int r) {
  extern __shared__ unsigned char smem_raw[];
  const int rd = 2 * r + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T* g = reinterpret_cast<T*>(smem_raw) + (size_t)warp * rd * rd;
  const int b = blockIdx.y;
  const int pix = blockIdx.x * 8 + warp;
  if (pix >= HW1) return;
  for (int c = lane; c < rd * rd; c += 32) g[c] = cg[((size_t)b * rd * rd + c) * HW1 + pix];
  __syncwarp();
  const float x0 = coords[((size_t)b * 2 + 0) * HW1 + pix];
  const float y0 = coords[((size_t)b * 2 + 1) * HW1 + pix];
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int xs = window_origin(fx0, r), ys = window_origin(fy0, r);
  const T w00 = Arith<T>::wcast((1.0f - dx) * (1.0f - dy));
  const T w01 = Arith<T>::wcast((1.0f - dx) * dy);
  const T w10 = Arith<T>::wcast(dx * (1.0f - dy));
  const T w11 = Arith<T>::wcast(dx * dy);
  T* plane = vg + ((size_t)b * HW1 + pix) * ((size_t)H2 * W2);
  const int n = H2 * W2;
  for (int e = lane; e < n; e += 32) {
    const int y1 = e / W2, x1 = e - y1 * W2;
    const int i = x1 - xs, j = y1 - ys;
    T acc = Arith<T>::zero();
    if (i >= 0 && i <= rd && j >= 0 && j <= rd) {
      if (i > 0 && j > 0) acc = Arith<T>::mad(acc, g[(i - 1) * rd + (j - 1)], w11);
      if (i > 0 && j < rd) acc = Arith<T>::mad(acc, g[(i - 1) * rd + j], w10);
      if (i < rd && j > 0) acc = Arith<T>::mad(acc, g[i * rd + (j - 1)], w01);
      if (i < rd && j < rd) acc = Arith<T>::mad(acc, g[i * rd + j], w00);
    }
    plane[e] = acc;
  }
}

template <typename T>
static int launch_fwd(const void* volume, const float* coords, void* out, int B, int H1, int W1,
                      int H2, int W2, int radius, cudaStream_t st) {
  const int HW1 = H1 * W1;
  const T* vol = reinterpret_cast<const T*>(volume);
  T* o = reinterpret_cast<T*>(out);
  if (radius == 3) {
    const int epc = 16 / (int)sizeof(T);
    const bool vec = sizeof(T) <= 4 && (W2 % epc == 0) && ((uintptr_t)volume % 16 == 0) &&
                     (((size_t)H2 * W2) % epc == 0);
    dim3 grid((HW1 + 31) / 32, B);
    if (vec) DROID_LAUNCH corr_index_fwd_r3_kernel<T, true><<<grid, 256, 0, st>>>(vol, coords, o, HW1, H2, W2);
    else DROID_LAUNCH corr_index_fwd_r3_kernel<T, false><<<grid, 256, 0, st>>>(vol, coords, o, HW1, H2, W2);
  } else {
    dim3 grid((HW1 + 255) / 256, B);
    DROID_LAUNCH corr_index_fwd_generic_kernel<T><<<grid, 256, 0, st>>>(vol, coords, o, HW1, H2, W2, radius);
  }
  return check_launch();
}

template <typename T>
static int launch_bwd(const float* coords, const void* cg, void* vg, int B, int H1, int W1, int H2,
                      int W2, int radius, cudaStream_t st) {
  const int HW1 = H1 * W1;
  const int rd = 2 * radius + 1;
  dim3 grid((HW1 + 7) / 8, B);
  size_t smem = (size_t)8 * rd * rd * sizeof(T);
  if (smem > 48 * 1024) return DROID_ERR_INVALID_ARGUMENT;
  DROID_LAUNCH corr_index_bwd_kernel<T><<<grid, 256, smem, st>>>(coords, reinterpret_cast<const T*>(cg),
                                                    reinterpret_cast<T*>(vg), HW1, H2, W2, radius);
  return check_launch();
}

}  // namespace droid

using namespace droid;

extern "C" int droid_corr_index_forward(const void* volume, const float* coords, void* out,
                                        int dtype, int B, int H1, int W1, int H2, int W2, int radius,
                                        void* stream) {
  if (B < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || radius < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0) return DROID_OK;
  if (B > 65535) return DROID_ERR_INVALID_ARGUMENT;
  if (!volume || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case DROID_DTYPE_F32: return launch_fwd<float>(volume, coords, out, B, H1, W1, H2, W2, radius, st);
    case DROID_DTYPE_F16: return launch_fwd<__half>(volume, coords, out, B, H1, W1, H2, W2, radius, st);
    case DROID_DTYPE_F64: return launch_fwd<double>(volume, coords, out, B, H1, W1, H2, W2, radius, st);
  }
  return DROID_ERR_UNSUPPORTED_DTYPE;
}

extern "C" int droid_corr_index_backward(const float* coords, const void* corr_grad,
                                         void* volume_grad, int dtype, int B, int H1, int W1, int H2,
                                         int W2, int radius, void* stream) {
  if (B < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || radius < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0) return DROID_OK;
  if (B > 65535) return DROID_ERR_INVALID_ARGUMENT;
  if (!coords || !corr_grad || !volume_grad) return DROID_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case DROID_DTYPE_F32: return launch_bwd<float>(coords, corr_grad, volume_grad, B, H1, W1, H2, W2, radius, st);
    case DROID_DTYPE_F16: return launch_bwd<__half>(coords, corr_grad, volume_grad, B, H1, W1, H2, W2, radius, st);
    case DROID_DTYPE_F64: return launch_bwd<double>(coords, corr_grad, volume_grad, B, H1, W1, H2, W2, radius, st);
  }
  return DROID_ERR_UNSUPPORTED_DTYPE;
}

extern "C" int droid_corr_index_pyramid(const void* const* volumes_host, int num_levels,
                                        const float* coords, void* out, int dtype, int B, int H1,
                                        int W1, int H2, int W2, int radius, void* stream) {
  if (B < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0) return DROID_ERR_INVALID_ARGUMENT;
  if (radius != 3 || num_levels < 1 || num_levels > 4) return DROID_ERR_INVALID_ARGUMENT;
  if ((H2 >> (num_levels - 1)) <= 0 || (W2 >> (num_levels - 1)) <= 0) return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0) return DROID_OK;
  if (B > 65535 || !volumes_host || !coords || !out) return DROID_ERR_INVALID_ARGUMENT;
  PyrArgs a;
  const int esz = dtype == DROID_DTYPE_F16 ? 2 : 4;
  const int epc = 16 / esz;
  for (int l = 0; l < 4; l++) {
    a.vol[l] = l < num_levels ? volumes_host[l] : nullptr;
    a.vec[l] = 0;
    if (l < num_levels) {
      if (!a.vol[l]) return DROID_ERR_INVALID_ARGUMENT;
      const int h2 = H2 >> l, w2 = W2 >> l;
      a.vec[l] = (w2 % epc == 0) && ((uintptr_t)a.vol[l] % 16 == 0) && (((size_t)h2 * w2) % epc == 0);
    }
  }
  const int HW1 = H1 * W1;
  dim3 grid((HW1 + 31) / 32, B);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == DROID_DTYPE_F16)
    DROID_LAUNCH corr_index_pyr_r3_kernel<__half><<<grid, 256, 0, st>>>(a, num_levels, coords,
                                                           reinterpret_cast<__half*>(out), HW1, H2, W2);
  else if (dtype == DROID_DTYPE_F32)
    DROID_LAUNCH corr_index_pyr_r3_kernel<float><<<grid, 256, 0, st>>>(a, num_levels, coords,
                                                          reinterpret_cast<float*>(out), HW1, H2, W2);
  else
    return DROID_ERR_UNSUPPORTED_DTYPE;
  return check_launch();
}
