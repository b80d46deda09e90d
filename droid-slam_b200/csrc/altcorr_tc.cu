// On-the-fly ("alt") correlation on the tensor cores (tcgen05 / TMEM / TMA), sm_100a.
// Same result as altcorr_fwd_r3_kernel (altcorr.cu), i.e. the reference's altcorr_forward
// (src/altcorr_kernel.cu:27-149) for radius 3 and 128 channels, for fp16 feature maps -- the only case the
// reference's callers produce: AltCorrBlock keeps an fp16 pyramid and hands `.float()` copies to the op
// (droid_slam/modules/corr.py:91-120).
//
// Why a GEMM: every source pixel needs 64 dot products of 128 channels against an 8x8 window of the target frame,
// and the windows of neighbouring pixels overlap almost entirely.  The CUDA-core kernel re-reads a 256 / 512-byte row
// per tap (27 instructions per tap, 6.6 TFLOP/s).  Here a CTA takes 128 source pixels (one A tile, loaded once) and
// sweeps the target level in chunks of NC = 64 or 128 pixels (whole image rows): D = A B^T (128 x NC, K = 128,
// kind::f16, fp32 accumulation in TMEM) holds the correlation of every source pixel of the tile with every target
// pixel of the chunk, and each source pixel picks the taps of its own window out of its accumulator row.  Only the
// chunks some pixel of the CTA needs are computed: for a smooth flow field the window rows of two image rows (~9-12 of
// 40 rows), for arbitrary coordinates the whole frame -- the result never depends on it.  Up to four pyramid levels
// run in one launch (the chunk stream continues across levels; level l + 1 is multiplied while level l is sampled).
// fp16 products are exact in fp32 and K = 128 is 8 chained MMAs, so the values agree with the fp32 FMA chain of the
// CUDA-core kernel to ~1e-7 of the largest value (the reference's own bar for the lookups is 1e-5).
//
// Warp roles (192 threads; NC = 64: 100 KB of shared memory, 256 TMEM columns, two CTAs per SM):
//   warps 0-3  samplers: thread = TMEM lane = source pixel.  Per chunk: accumulator row -> registers (tcgen05.ld) ->
//              the thread's own row of a shared staging buffer -> the taps of the window rows that lie in the chunk,
//              into registers with static indices; chunks no lane of the warp needs are only acknowledged.  After the
//              last chunk of a level: bilinear weights, 49 coalesced stores per pixel;
//   warp 4     TMA producer: the A tile at kernel entry, then the B chunks through a 2-stage ring (two 64-channel
//              K slabs per stage, 128-byte swizzle);
//   warp 5     MMA issuer: 8 tcgen05.mma per chunk into accumulator (chunk & 3); TMEM = 4 x NC columns.
//
// The fp32 entry point (the drop-in signature) copies the TARGET map to fp16 in a pre-pass and converts the SOURCE map
// tile by tile inside this kernel (template flag A32), both with a round-trip check; if a value is not an fp16 number
// (a caller that feeds real fp32 features) a device-side flag makes the remaining CTAs exit and the CUDA-core kernel,
// launched behind it, redo the call -- no host round trip.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstring>

#include "common.cuh"

namespace droid {
namespace atc {

static constexpr int kC = 128;
static constexpr int kTileM = 128;                   // source pixels per A tile = TMEM lanes
static constexpr int kTiles = 1;                     // A tiles per CTA
static constexpr int kSlots = 4;                     // accumulators per tile
static constexpr int kASlab = kTileM * 128;          // 64 channels (128 B) x 128 rows
static constexpr int kATile = 2 * kASlab;            // 32 KB
static constexpr int kBStages = 2;
static constexpr int kSamplers = 4 * kTiles;         // sampler warps
static constexpr int kThreads = 32 * (kSamplers + 2);
static constexpr int kAOff = 0;
static constexpr int kBOff = kTiles * kATile;
// Everything that depends on the accumulator width NC = target pixels per chunk: 64 (image rows of up to 64 pixels:
// 100 KB of shared memory and 256 TMEM columns per CTA, two CTAs per SM) or 128 (rows of up to 128 pixels, e.g. the
// 73-pixel rows of a 16:9 video at 1/8 resolution: 164 KB and 512 columns, one CTA per SM).
template <int NC>
struct Cfg {
  static constexpr int kN = NC;
  static constexpr int kBSlab = NC * 128;                            // 64 channels x NC rows
  static constexpr int kBStage = 2 * kBSlab;
  static constexpr int kRowLd = NC + 4;                              // floats per staging row (pad: conflict-free 16-byte stores)
  static constexpr int kStageWarp = 32 * kRowLd * 4;
  static constexpr int kStOff = kBOff + kBStages * kBStage;
  static constexpr int kBarOff = kStOff + kSamplers * kStageWarp;
  static constexpr int kSmem = kBarOff + 512 + 1024;
  static constexpr int kTmemCols = kTiles * kSlots * NC;             // 256 / 512: a power of two
  static constexpr int kCtasPerSm = NC == 64 && kTiles == 1 ? 2 : 1;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!done);
}
// 2-D tiled load (channel, row) -> shared memory, completion on an mbarrier
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c, int row) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c), "r"(row) : "memory");
}
// shared-memory matrix descriptor, K-major 16-bit operands, SWIZZLE_128B: rows of 128 bytes (64 channels), atoms of
// 8 rows (1024 B = SBO).  One MMA consumes K = 16 channels = 32 bytes of every row.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;                  // LBO: unused for swizzled K-major operands
  d |= (uint64_t)(1024 >> 4) << 32;        // SBO
  d |= (uint64_t)1 << 46;                  // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                  // SWIZZLE_128B
  return d;
}
// instruction descriptor, kind::f16: D = f32, A = B = f16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t umma_idesc(int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

static constexpr int kMaxLevels = 4;

// Everything the kernel needs about the target side: one tensor map and one geometry per pyramid level.
struct Params {
  CUtensorMap a;                    // f1 rows: [frames1 * HW1][128] fp16
  CUtensorMap b[kMaxLevels];        // f2 rows of level l: [frames2 * H2 * W2][128] fp16 (channel-last)
  int H2[kMaxLevels], W2[kMaxLevels];
  int Rh[kMaxLevels];               // whole image rows per chunk = 64 / W2
  float cscale[kMaxLevels];         // coordinates of level l = coords * cscale (2^-l: exact, = coords / 2^l of corr.py:116)
  int nlev;
};

// The lookup window of one pixel at one level: integer origin, the valid taps of a window row, the chunks it touches.
struct Window {
  int xs, ys, c_lo, c_hi;           // c_lo > c_hi: no tap inside the image
  unsigned colmask;
  float dx, dy;
};
__device__ __forceinline__ Window make_window(float x, float y, int H2, int W2, int Rh, bool live) {
  Window w;
  const float fx0 = floorf(x), fy0 = floorf(y);
  w.dx = x - fx0;
  w.dy = y - fy0;
  w.xs = window_origin(fx0, 3);
  w.ys = window_origin(fy0, 3);
  w.colmask = 0;
#pragma unroll
  for (int ix = 0; ix < 8; ix++) w.colmask |= ((unsigned)(w.xs + ix) < (unsigned)W2) ? (1u << ix) : 0u;
  const int r_lo = max(w.ys, 0), r_hi = min(w.ys + 7, H2 - 1);
  w.c_lo = 1;
  w.c_hi = 0;
  if (live && w.colmask && r_lo <= r_hi) {
    w.c_lo = r_lo / Rh;
    w.c_hi = r_hi / Rh;
  }
  return w;
}

// grid = (ceil(HW1 / 256), N, B).  out [B, N, nlev * 49, HW1].  `skip_if` (may be null): a device flag; non-zero =
// leave everything to the CUDA-core kernel launched behind this one.
// A32: the source map arrives as fp32 (`f1_32`: [B][HW1][128], the drop-in signature) and the sampler warps convert
// their A tile to fp16 themselves -- with the same round-trip check as the pre-pass, raising `flag` -- instead of a
// TMA load from an fp16 copy: the copy of f1 (3/4 of the conversion traffic over a pyramid) is never made.
template <int NC, bool A32>
__global__ void __launch_bounds__(kThreads, Cfg<NC>::kCtasPerSm)
altcorr_tc_kernel(const __grid_constant__ Params P, const int64_t* __restrict__ ii, const int64_t* __restrict__ jj,
                  const float* __restrict__ coords, float* __restrict__ out, int* __restrict__ flag, int N,
                  int HW1, const float* __restrict__ f1_32) {
  if (flag && *reinterpret_cast<volatile int*>(flag)) return;
  using C = Cfg<NC>;
  constexpr int kN = C::kN, kBSlab = C::kBSlab, kBStage = C::kBStage, kRowLd = C::kRowLd, kStageWarp = C::kStageWarp;
  constexpr int kStOff = C::kStOff, kBarOff = C::kBarOff, kTmemCols = C::kTmemCols;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBarOff);
  uint64_t* a_full = bars;                               // [1]
  uint64_t* b_full = bars + 1;                           // [kBStages]
  uint64_t* b_empty = b_full + kBStages;                 // [kBStages]
  uint64_t* acc_full = b_empty + kBStages;               // [kTiles * kSlots]
  uint64_t* acc_empty = acc_full + kTiles * kSlots;      // [kTiles * kSlots]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kTiles * kSlots);
  int* crange = reinterpret_cast<int*>(tmem_slot + 2);   // [2 * nlev]: first / last chunk of a level any pixel of the CTA needs

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, n = blockIdx.y;
  const int pix0 = blockIdx.x * (kTiles * kTileM);
  const int nlev = P.nlev;

  // every sampler thread: its pixel; the coordinate load is in flight while the barriers / TMEM are set up
  const int tile = warp >> 2, q = warp & 3;
  const int pix = pix0 + tile * kTileM + q * 32 + lane;
  const bool live = warp < kSamplers && pix < HW1;
  float x0 = 0.f, y0 = 0.f;
  if (live) {
    const float2 xy = __ldg(reinterpret_cast<const float2*>(coords + (((size_t)b * N + n) * HW1 + pix) * 2));
    x0 = xy.x;
    y0 = xy.y;
  }

  // The producer thread sets up the barriers and starts the A tiles at once: they do not depend on the chunk ranges
  // computed below, and their latency (frame index -> TMA -> HBM) otherwise sits on every CTA's critical path.
  int fr2 = 0;
  if (threadIdx.x == kSamplers * 32) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&P.a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&P.b[0]) : "memory");
    mbar_init(a_full, A32 ? kSamplers : 1);
    for (int i = 0; i < kBStages; i++) {
      mbar_init(b_full + i, 1);
      mbar_init(b_empty + i, 1);
    }
    for (int i = 0; i < kTiles * kSlots; i++) {
      mbar_init(acc_full + i, 1);
      mbar_init(acc_empty + i, 4);                       // one arrival per sampler warp of the tile
    }
    for (int l = 0; l < kMaxLevels; l++) {
      crange[2 * l] = 1 << 30;
      crange[2 * l + 1] = -1;
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fr2 = jj ? (int)jj[b] : b;
    if constexpr (!A32) {
      const int fr1 = ii ? (int)ii[b] : b;
      mbar_expect_tx(a_full, kTiles * kATile);
      for (int t = 0; t < kTiles; t++)
        for (int ks = 0; ks < 2; ks++)
          tma_load_2d(smem + kAOff + t * kATile + ks * kASlab, &P.a, a_full, ks * 64, fr1 * HW1 + pix0 + t * kTileM);
    }
  }
  if (warp == kSamplers) {
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();

  if constexpr (A32) {
    // fp32 A tile -> fp16, K-major SWIZZLE_128B: a warp converts 32 rows, one row (512 bytes) per step, lane l the
    // channels 4l .. 4l + 3 = half of the 16-byte unit (l & 15) >> 1 of K slab l >> 4
    if (warp < kSamplers) {
      const float* src = f1_32 + ((size_t)(ii ? (size_t)ii[b] : (size_t)b) * HW1 + pix0 + tile * kTileM) * kC + 4 * lane;
      unsigned char* dstA = smem + kAOff + tile * kATile + (lane >> 4) * kASlab + ((lane & 1) << 3);
      const int unit = (lane & 15) >> 1;
      bool bad = false;
#pragma unroll 8
      for (int i = 0; i < 32; i++) {
        const int r = q * 32 + i;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (pix0 + tile * kTileM + r < HW1) v = __ldg(reinterpret_cast<const float4*>(src + (size_t)r * kC));
        const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
        const float2 f0 = __half22float2(h0), f1v = __half22float2(h1);
        bad |= !(f0.x == v.x && f0.y == v.y && f1v.x == v.z && f1v.y == v.w);
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t*>(&h0);
        o.y = *reinterpret_cast<const uint32_t*>(&h1);
        *reinterpret_cast<uint2*>(dstA + r * 128 + ((unit ^ (r & 7)) << 4)) = o;
      }
      if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(flag, 1);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(a_full);
    }
  }
  // ---- per level, the chunks the pixel's window touches ----
  if (warp < kSamplers) {
    for (int l = 0; l < nlev; l++) {
      const Window w = make_window(x0 * P.cscale[l], y0 * P.cscale[l], P.H2[l], P.W2[l], P.Rh[l], live);
      int lo = w.c_lo <= w.c_hi ? w.c_lo : (1 << 30), hi = w.c_lo <= w.c_hi ? w.c_hi : -1;
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
      }
      if (lane == 0 && hi >= 0) {
        atomicMin(crange + 2 * l, lo);
        atomicMax(crange + 2 * l + 1, hi);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  int total = 0;                                          // chunks of all levels; cmax < cmin: nothing to correlate
  for (int l = 0; l < nlev; l++) total += max(crange[2 * l + 1] - crange[2 * l] + 1, 0);

  if (warp == kSamplers) {
    // ---------------- TMA producer ----------------
    if (!A32 && lane == 0 && total == 0) mbar_wait(a_full, 0);   // nothing to correlate: the A tiles must still have landed
    if (lane == 0 && total > 0) {
      int it = 0;
      for (int l = 0; l < nlev; l++) {
        const int row0 = fr2 * P.H2[l] * P.W2[l], step = P.Rh[l] * P.W2[l];
        for (int c = crange[2 * l]; c <= crange[2 * l + 1]; c++, it++) {
          const int s = it % kBStages;
          if (it >= kBStages) mbar_wait(b_empty + s, ((it / kBStages) - 1) & 1);
          mbar_expect_tx(b_full + s, kBStage);
          for (int ks = 0; ks < 2; ks++)
            tma_load_2d(smem + kBOff + s * kBStage + ks * kBSlab, &P.b[l], b_full + s, ks * 64, row0 + c * step);
        }
      }
    }
    __syncwarp();
  } else if (warp == kSamplers + 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0 && total > 0) {
      const uint32_t idesc = umma_idesc(kN);
      mbar_wait(a_full, 0);
      for (int it = 0; it < total; it++) {
        const int s = it % kBStages, slot = it & (kSlots - 1);
        mbar_wait(b_full + s, (it / kBStages) & 1);
        const uint32_t bs = smem_u32(smem + kBOff + s * kBStage);
        for (int t = 0; t < kTiles; t++) {
          if (it >= kSlots) mbar_wait(acc_empty + t * kSlots + slot, ((it / kSlots) - 1) & 1);
          tc_fence_after();
          const uint32_t as = smem_u32(smem + kAOff + t * kATile);
#pragma unroll
          for (int k = 0; k < kC / 16; k++) {          // k >> 2 = K slab, (k & 3) * 32 bytes inside the 128-byte row
            const uint64_t ad = umma_desc(as + (k >> 2) * kASlab + (k & 3) * 32);
            const uint64_t bd = umma_desc(bs + (k >> 2) * kBSlab + (k & 3) * 32);
            umma_f16(tmem + (t * kSlots + slot) * kN, ad, bd, idesc, k > 0);
          }
          umma_commit(acc_full + t * kSlots + slot);
        }
        umma_commit(b_empty + s);
      }
    }
    __syncwarp();
  } else {
    // ---------------- samplers ----------------
    float* mine = reinterpret_cast<float*>(smem + kStOff + warp * kStageWarp) + lane * kRowLd;
    int it = 0;
    for (int l = 0; l < nlev; l++) {
      const int H2 = P.H2[l], W2 = P.W2[l], Rh = P.Rh[l];
      const Window w = make_window(x0 * P.cscale[l], y0 * P.cscale[l], H2, W2, Rh, live);
      float taps[64];
#pragma unroll
      for (int i = 0; i < 64; i++) taps[i] = 0.f;
      for (int c = crange[2 * l]; c <= crange[2 * l + 1]; c++, it++) {
        const int slot = it & (kSlots - 1);
        const int h0 = c * Rh;
        const bool need = c >= w.c_lo && c <= w.c_hi;
        mbar_wait(acc_full + tile * kSlots + slot, (it / kSlots) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, need)) {
          const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + (tile * kSlots + slot) * kN;
#pragma unroll
          for (int h = 0; h < kN / 64; h++) {          // 64 columns at a time: two loads in flight, one wait
            uint32_t r[64];
            tmem_ld32(taddr + 64 * h, r);
            tmem_ld32(taddr + 64 * h + 32, r + 32);
            tmem_ld_wait();
            uint4* dst = reinterpret_cast<uint4*>(mine + 64 * h);
#pragma unroll
            for (int j = 0; j < 16; j++) dst[j] = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty + tile * kSlots + slot);
        if (need) {
          // window rows iy with first <= iy < first + cnt lie in this chunk (and inside the image)
          const int first = max(max(h0, 0) - w.ys, 0);
          const int cnt = max(min(min(h0 + Rh, H2) - w.ys, 8) - first, 0);
          const float* row0 = mine + (w.ys - h0) * W2 + w.xs;
#pragma unroll
          for (int iy = 0; iy < 8; iy++) {
            if ((unsigned)(iy - first) < (unsigned)cnt) {
              const float* row = row0 + iy * W2;
#pragma unroll
              for (int ix = 0; ix < 8; ix++)
                if (w.colmask & (1u << ix)) taps[iy * 8 + ix] = row[ix];
            }
          }
        }
      }
      if (live) {
        const float dx = w.dx, dy = w.dy;
        const float wse = (1 - dy) * (1 - dx), wsw = (1 - dy) * dx, wne = dy * (1 - dx), wnw = dy * dx;
        float* ob = out + (((size_t)b * N + n) * nlev + l) * 49 * HW1 + pix;
#pragma unroll
        for (int ixo = 0; ixo < 7; ixo++) {
#pragma unroll
          for (int iyo = 0; iyo < 7; iyo++) {      // channel = iy + 7 * ix (src/altcorr_kernel.cu:102-105)
            const int t = iyo * 8 + ixo;
            float acc = taps[t] * wse;
            acc = fmaf(taps[t + 1], wsw, acc);
            acc = fmaf(taps[t + 8], wne, acc);
            acc = fmaf(taps[t + 9], wnw, acc);
            __stcs(ob + (size_t)(iyo + 7 * ixo) * HW1, acc);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kSamplers) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
  }
}

// fp32 -> fp16 with a round-trip check: *flag |= 1 if a value is not exactly representable (NaN included).
__global__ void __launch_bounds__(256) to_half_checked_kernel(const float4* __restrict__ src, uint2* __restrict__ dst,
                                                               size_t n4, int* __restrict__ flag) {
  bool bad = false;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldcs(src + i);
    const __half2 a = __floats2half2_rn(v.x, v.y), c = __floats2half2_rn(v.z, v.w);
    const float2 fa = __half22float2(a), fc = __half22float2(c);
    bad |= !(fa.x == v.x && fa.y == v.y && fc.x == v.z && fc.y == v.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&a);
    o.y = *reinterpret_cast<const uint32_t*>(&c);
    dst[i] = o;
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows][128] fp16 as (channel, row); box = 64 channels (128 bytes) x box_rows, 128-byte swizzle, zero fill outside
static bool make_map(CUtensorMap* m, const void* base, size_t rows, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {(cuuint64_t)kC, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)kC * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace atc

// True if the tensor-core kernel covers this shape (fp16 maps, 128 channels, radius 3, image rows of <= 128 pixels).
bool altcorr_tc_supported(int B, int N, int H1, int W1, int H2, int W2, int C, int radius, size_t frames1, size_t frames2) {
  if (C != atc::kC || radius != 3 || W2 > 128 || W2 <= 0 || H2 <= 0) return false;
  if (B > 65535 || N > 65535) return false;
  const size_t r1 = frames1 * (size_t)H1 * W1, r2 = frames2 * (size_t)H2 * W2;
  return r1 < (1u << 31) && r2 < (1u << 31);
}

// Enqueue the tensor-core forward over `nlev` pyramid levels in ONE launch.  f1: fp16 channel-last maps of `frames1`
// frames; f2[l]: level l of `frames2` frames, H2[l] x W2[l]; coordinates are multiplied by cscale[l] for level l;
// ii / jj may be null (frame = b); out [B, N, nlev * 49, H1, W1].  Returns DROID_ERR_NOT_IMPLEMENTED if a shape is not
// covered (nothing enqueued then).
int altcorr_tc_forward_levels(const void* f1, const void* const* f2, const int* H2, const int* W2, const float* cscale,
                              int nlev, const int64_t* ii, const int64_t* jj, const float* coords, float* out,
                              int* skip_if, size_t frames1, size_t frames2, int B, int N, int H1, int W1, int C,
                              int radius, cudaStream_t st, bool f1_is_f32) {
  if (f1_is_f32 && !skip_if) return DROID_ERR_INVALID_ARGUMENT;      // the in-kernel conversion reports through the flag
  if (nlev < 1 || nlev > atc::kMaxLevels) return DROID_ERR_NOT_IMPLEMENTED;
  if (((uintptr_t)f1 & 15) || ((uintptr_t)coords & 7)) return DROID_ERR_NOT_IMPLEMENTED;
  const int HW1 = H1 * W1;
  atc::Params P;
  memset(&P, 0, sizeof(P));
  P.nlev = nlev;
  if (!f1_is_f32 && !atc::make_map(&P.a, f1, frames1 * (size_t)HW1, atc::kTileM)) return DROID_ERR_NOT_IMPLEMENTED;
  int wmax = 0;
  for (int l = 0; l < nlev; l++) wmax = W2[l] > wmax ? W2[l] : wmax;
  const int nc = wmax <= 64 ? 64 : 128;                 // accumulator width: the widest level decides
  for (int l = 0; l < nlev; l++) {
    if (!altcorr_tc_supported(B, N, H1, W1, H2[l], W2[l], C, radius, frames1, frames2)) return DROID_ERR_NOT_IMPLEMENTED;
    if (((uintptr_t)f2[l] & 15) || !atc::make_map(&P.b[l], f2[l], frames2 * (size_t)H2[l] * W2[l], nc))
      return DROID_ERR_NOT_IMPLEMENTED;
    P.H2[l] = H2[l];
    P.W2[l] = W2[l];
    P.Rh[l] = nc / W2[l];
    P.cscale[l] = cscale[l];
  }
  DevCtx& dc = dev_ctx();
  if (dc.first(kOnceAltcorrTc)) {
    cudaFuncSetAttribute(atc::altcorr_tc_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::Cfg<64>::kSmem);
    cudaFuncSetAttribute(atc::altcorr_tc_kernel<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::Cfg<128>::kSmem);
    cudaFuncSetAttribute(atc::altcorr_tc_kernel<64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::Cfg<64>::kSmem);
    cudaFuncSetAttribute(atc::altcorr_tc_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::Cfg<128>::kSmem);
  }
  dim3 grid((HW1 + atc::kTiles * atc::kTileM - 1) / (atc::kTiles * atc::kTileM), N, B);
  const float* f32 = f1_is_f32 ? reinterpret_cast<const float*>(f1) : nullptr;
#define ATC_LAUNCH(NC, A32) \
  DROID_LAUNCH atc::altcorr_tc_kernel<NC, A32><<<grid, atc::kThreads, atc::Cfg<NC>::kSmem, st>>>(P, ii, jj, coords, out, skip_if, N, HW1, f32)
  if (nc == 64) {
    if (f1_is_f32) ATC_LAUNCH(64, true); else ATC_LAUNCH(64, false);
  } else {
    if (f1_is_f32) ATC_LAUNCH(128, true); else ATC_LAUNCH(128, false);
  }
#undef ATC_LAUNCH
  return check_launch();
}

int altcorr_tc_forward(const void* f1, const void* f2, const int64_t* ii, const int64_t* jj, const float* coords, float* out,
                       int* skip_if, size_t frames1, size_t frames2, int B, int N, int H1, int W1, int H2, int W2,
                       int C, int radius, cudaStream_t st, bool f1_is_f32) {
  const float one = 1.f;
  return altcorr_tc_forward_levels(f1, &f2, &H2, &W2, &one, 1, ii, jj, coords, out, skip_if, frames1, frames2, B, N, H1, W1,
                                   C, radius, st, f1_is_f32);
}

// fp32 -> fp16 copy of a map into `dst`, *flag = 0 first, then 1 if any value is not exactly a half.
int altcorr_tc_convert(const float* src, void* dst, int* flag, size_t n, cudaStream_t st) {
  cudaMemsetAsync(flag, 0, sizeof(int), st);
  const size_t a = n / 4;
  DROID_LAUNCH atc::to_half_checked_kernel<<<(unsigned)std::min<size_t>((a + 255) / 256, 148 * 16), 256, 0, st>>>(
      reinterpret_cast<const float4*>(src), reinterpret_cast<uint2*>(dst), a, flag);
  return check_launch();
}

}  // namespace droid
