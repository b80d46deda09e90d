// All-pairs correlation volume + 4-level average-pool pyramid (CorrBlock build) for sm_100a:
// tcgen05.mma with the accumulator in TMEM, operands brought in by TMA, pooling fused into the epilogue.
// Replaces droid_slam/modules/corr.py:63-71 (torch.matmul of the two /4-scaled fp16 feature maps) and
// corr.py:35-38 (3x F.avg_pool2d(2,2)), i.e. one cuBLAS HGEMM + three pooling kernels + four HBM round trips.
//
//   corr[b, m, n] = sum_c (f1[b,c,m]/4) * (f2[b,c,n]/4)      m = pixel of frame 1, n = (h2,w2) pixel of frame 2
//
// Layout facts that shape the kernel:
//  * both operands are stored [C][HW] (pixel index contiguous, channel strided) => "MN-major" UMMA operands;
//    a TMA box of 64 pixels x 128 channels with the 128-byte swizzle is exactly 16 canonical 8x128B swizzle
//    atoms stacked along K, so no repacking pass is needed;
//  * K = C = 128 fits in shared memory whole: no K pipeline; the B operand (8 image rows of frame 2 = 8*W
//    pixels, 128 KB at W = 64) is loaded ONCE per CTA and reused for every 128-pixel tile of frame 1, only the
//    32 KB A tiles stream (double-buffered);
//  * the accumulator tile is 128 (m) x 8*W (n) fp32 = the whole TMEM (512 columns at W = 64): lane = m, column =
//    (h2 - 8*tile, w2).  An 8-row N tile holds complete 2x2, 4x4 and 8x8 pooling windows, so each epilogue
//    thread (one TMEM lane = one pixel m) pools its own row without any cross-thread exchange;
//  * the kernel is HBM-WRITE bound, not tensor bound: 2.4 GFLOP vs 26.6 MB of fp16 output per edge.
//    The two TMEM halves (4 image rows each) are drained and refilled independently so the MMAs of the next tile
//    overlap the stores of the current one.
//
// Warp roles: W == 64 (288 threads): warps 0-3 / 4-7 = the two epilogue groups (TMEM lane quarter = warp & 3, accumulator
// half = warp >> 2), warp 8 = TMA producer + MMA issuer; W < 64 (160 threads): warps 0-3 = epilogue, warp 4 = producer.
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"

namespace droid {
namespace cb {

static constexpr int kC = 128;           // channels (K)
static constexpr int kTileM = 128;       // frame-1 pixels per tile (TMEM lanes)
static constexpr int kBoxPix = 64;       // pixels per TMA box (128 bytes of fp16 = swizzle span)
static constexpr int kBoxBytes = kBoxPix * kC * 2;   // 16 KB
static constexpr int kABytes = 2 * kBoxBytes;        // 128 pixels
static constexpr int kMaxNBoxes = 8;                 // 8 rows x 64 columns
static constexpr int kThreads = 160;                 // 4 epilogue warps + producer (W < 64)
static constexpr int kThreads64 = 288;               // 8 epilogue warps + producer (W == 64)

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
// TMA: 3-D tiled load (pixel, channel, batch) -> shared memory, completion on an mbarrier
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}
// UMMA shared-memory matrix descriptor: MN-major, 128-byte swizzle (layout type 2), descriptor version 1.
//   LBO = byte distance between consecutive 64-pixel blocks (one TMA box), SBO = distance between 8-channel groups.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);            // start address        bits [0,14)
  d |= (uint64_t)(kBoxBytes >> 4) << 16;               // leading byte offset  bits [16,30)
  d |= (uint64_t)(1024 >> 4) << 32;                    // stride byte offset   bits [32,46)
  d |= (uint64_t)1 << 46;                              // version = 1 (Blackwell)
  d |= (uint64_t)2 << 61;                              // SWIZZLE_128B
  return d;
}
// instruction descriptor, kind::f16: D = f32, A = B = f16, both MN-major, M = 128, N = n
__device__ __forceinline__ uint32_t umma_idesc(int n) {
  return (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// TMA store of one staged level-0 row block: shared [128 pixels][64 columns] (128B-swizzled) -> lvl0[b][m..][h2][:]
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
// one named barrier per epilogue group (id 1 / 2), 128 threads each
__device__ __forceinline__ void epi_barrier_g(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }
__device__ __forceinline__ float2 unpack_h2(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }
__device__ __forceinline__ uint32_t pool4_h2(uint32_t top_a, uint32_t bot_a, uint32_t top_b, uint32_t bot_b);

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// 2x2 mean of fp16 values exactly like F.avg_pool2d on half: float accumulation in the order
// (h0,w0),(h0,w1),(h1,w0),(h1,w1), divide by 4, round to half.
__device__ __forceinline__ float pool4(float a, float b, float c, float d) {
  return __half2float(__float2half_rn((((a + b) + c) + d) * 0.25f));
}
__device__ __forceinline__ float rnd_h(float x) { return __half2float(__float2half_rn(x)); }
// two horizontally adjacent 2x2 windows: (top_a, bot_a) and (top_b, bot_b) hold (w, w+1) pairs of two rows
__device__ __forceinline__ uint32_t pool4_h2(uint32_t top_a, uint32_t bot_a, uint32_t top_b, uint32_t bot_b) {
  const float2 ta = unpack_h2(top_a), ba = unpack_h2(bot_a), tb = unpack_h2(top_b), bb = unpack_h2(bot_b);
  return pack_h2(pool4(ta.x, ta.y, ba.x, ba.y), pool4(tb.x, tb.y, bb.x, bb.y));
}

}  // namespace cb
}  // namespace droid

namespace droid {
namespace cb {

struct Smem {
  // operand buffers first (1024-byte aligned for the 128-byte swizzle)
  static constexpr int kBOff = 0;
  static constexpr int kAOff = kMaxNBoxes * kBoxBytes;          // 128 KB
  static constexpr int kStageOff = kAOff + kABytes;             // +32 KB: ONE A tile (see the producer)
  static constexpr int kStageBytes = kTileM * 64 * 2;           // 128 pixels x 64 columns fp16 = one level-0 row block
  static constexpr int kXchOff = kStageOff + 2 * kStageBytes;   // two staging buffers (W == 64: one per epilogue group)
  static constexpr int kXchBytes = kTileM * 8 * 4;              // level-2 values of the upper half, one tile
  static constexpr int kBarOff = kXchOff + 2 * kXchBytes;
  static constexpr int kBytes = kBarOff + 256 + 1024;
};


template <int W>
__global__ void __launch_bounds__(W == 64 ? kThreads64 : kThreads, 1)
corr_build_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmOut, __half* __restrict__ lvl0, __half* __restrict__ lvl1, __half* __restrict__ lvl2,
                  __half* __restrict__ lvl3, int H, int HW) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // the 128-byte swizzle needs 1024-byte aligned atoms; align by hand (1 KB of slack is allocated)
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  unsigned char* Bs = smem + Smem::kBOff;
  unsigned char* As = smem + Smem::kAOff;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Smem::kBarOff);
  uint64_t* b_full = bars + 0;
  uint64_t* a_full = bars + 1;     // the A tile of this iteration has landed
  uint64_t* a_empty = bars + 2;    // ... and its MMAs are done
  uint64_t* acc_full = bars + 3;   // [2]
  uint64_t* acc_empty = bars + 5;  // [2]
  uint64_t* x_full = bars + 7;     // [2]  W == 64: level-2 values of the upper half handed to the lower-half group
  uint64_t* x_empty = bars + 9;    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);
  constexpr int kProducer = W == 64 ? 8 : 4;      // warp index of the TMA / MMA warp = number of epilogue warps

  constexpr int NHALF = 4 * W;               // accumulator columns per half (4 image rows)
  constexpr int NBOX = 8 * W / kBoxPix;      // TMA boxes of the B tile
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nt = blockIdx.x, b = blockIdx.y;
  const int T = (HW + kTileM - 1) / kTileM;  // frame-1 tiles

  if (threadIdx.x == 0) {
    mbar_init(b_full, 1);
    mbar_init(a_full, 1);
    mbar_init(a_empty, 1);
    for (int i = 0; i < 2; i++) {
      mbar_init(acc_full + i, 1);
      mbar_init(acc_empty + i, 128);
      mbar_init(x_full + i, 128);
      mbar_init(x_empty + i, 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kProducer) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == kProducer) {
    if (lane == 0) {
      // ---------------- TMA producer + MMA issuer (one thread) ----------------
      // ONE A buffer: the MMAs of a tile take ~1 us and the tile after it cannot start before the epilogue has freed
      // the accumulator (several us later), so the next A tile is fetched as soon as this one's MMAs are done and is
      // there long before it is needed; the 32 KB saved pay for the second epilogue group's staging.
      mbar_expect_tx(b_full, NBOX * kBoxBytes);
      for (int i = 0; i < NBOX; i++) tma_load_3d(Bs + i * kBoxBytes, &tmB, b_full, nt * 8 * W + i * kBoxPix, 0, b);
      mbar_expect_tx(a_full, kABytes);
      for (int j = 0; j < 2; j++) tma_load_3d(As + j * kBoxBytes, &tmA, a_full, j * kBoxPix, 0, b);
      const uint32_t idesc = umma_idesc(NHALF);
      mbar_wait(b_full, 0);
      for (int t = 0; t < T; t++) {
        mbar_wait(a_full, t & 1);
        const uint64_t adesc = umma_desc(smem_u32(As));
        for (int h = 0; h < 2; h++) {
          if (t > 0) mbar_wait(acc_empty + h, (t - 1) & 1);
          tc_fence_after();
          const uint64_t bdesc = umma_desc(smem_u32(Bs + h * (NHALF / kBoxPix) * kBoxBytes));
#pragma unroll
          for (int k = 0; k < kC / 16; k++)   // 16 channels = two 8-channel swizzle atoms = 2048 bytes per step
            umma_f16(tmem + h * NHALF, adesc + (uint64_t)(k * 2048 >> 4), bdesc + (uint64_t)(k * 2048 >> 4), idesc, k > 0);
          umma_commit(acc_full + h);
        }
        umma_commit(a_empty);
        if (t + 1 < T) {
          mbar_wait(a_empty, t & 1);
          mbar_expect_tx(a_full, kABytes);
          for (int j = 0; j < 2; j++)
            tma_load_3d(As + j * kBoxBytes, &tmA, a_full, (t + 1) * kTileM + j * kBoxPix, 0, b);
        }
      }
    }
    __syncwarp();
  } else {
    // ---------------- epilogue: one TMEM lane (= one frame-1 pixel) per thread ----------------
    const float scale = 0.0625f;                 // (1/4) * (1/4), exact
    const int H2 = H / 2, W2 = W / 2, H4 = H / 4, W4 = W / 4, H8 = H / 8, W8 = W / 8;
    if constexpr (W == 64) {
      // Level 0 (75 % of the bytes) leaves through shared memory and TMA: every thread writes its pixel's 128-byte
      // row of the staged [128 x 64] block (128B-swizzled: conflict-free), one elected thread issues the bulk store.
      // TWO epilogue groups of four warps: group g owns accumulator half g (image rows 4g .. 4g+3 of the N tile), its
      // own staging buffer and named barrier, so both halves are drained at the same time and every scheduler has two
      // epilogue warps to interleave (with one group the kernel was the latency chain of that group: ~1540 cycles per
      // row, DRAM 37 % busy).  Levels 0-2 are complete inside a half; level 3 pools across the two, so group 0 hands
      // its eight level-2 values per pixel to group 1 through shared memory (double-buffered, mbarrier hand-shake).
      const int g = warp >> 2, q = warp & 3, h = g;
      unsigned char* buf = smem + Smem::kStageOff + g * Smem::kStageBytes;
      uint32_t* xch = reinterpret_cast<uint32_t*>(smem + Smem::kXchOff);
      const int row_in_tile = q * 32 + lane;
      const bool leader = q == 0 && lane == 0;
      for (int t = 0; t < T; t++) {
        const int m = t * kTileM + row_in_tile;
        const bool live = m < HW;
        const size_t pix = (size_t)b * HW + (live ? m : 0);
        uint32_t l2[8];
        mbar_wait(acc_full + h, t & 1);
        tc_fence_after();
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + h * NHALF;
        uint32_t prev[32], p1keep[16];
#pragma unroll
        for (int r = 0; r < 4; r++) {
          uint32_t cur[32];
          {
            float v[64];
#pragma unroll
            for (int j = 0; j < 4; j++) tmem_ld16(taddr + r * 64 + j * 16, v + j * 16);
            tmem_ld_wait();
            if (r == 3) {                         // this half of TMEM is drained: let the next tile's MMAs start
              tc_fence_before();
              mbar_arrive(acc_empty + h);
            }
#pragma unroll
            for (int i = 0; i < 32; i++) cur[i] = pack_h2(v[2 * i] * scale, v[2 * i + 1] * scale);
          }
          if (leader) bulk_wait_read<0>();        // the store of the previous row has read the staging buffer
          epi_barrier_g(g);
          uint4* rowp = reinterpret_cast<uint4*>(buf + row_in_tile * 128);
#pragma unroll
          for (int j = 0; j < 8; j++)
            rowp[j ^ (row_in_tile & 7)] = make_uint4(cur[4 * j], cur[4 * j + 1], cur[4 * j + 2], cur[4 * j + 3]);
          fence_async_smem();
          epi_barrier_g(g);
          if (leader) {
            tma_store_4d(&tmOut, buf, 0, nt * 8 + h * 4 + r, t * kTileM, b);
            bulk_commit();
          }
          if (r & 1) {
            uint32_t p1[16];
#pragma unroll
            for (int i = 0; i < 16; i++) p1[i] = pool4_h2(prev[2 * i], cur[2 * i], prev[2 * i + 1], cur[2 * i + 1]);
            if (live) {
              uint4* dst = reinterpret_cast<uint4*>(lvl1 + (pix * H2 + (nt * 4 + h * 2 + (r >> 1))) * W2);
#pragma unroll
              for (int j = 0; j < 4; j++) dst[j] = make_uint4(p1[4 * j], p1[4 * j + 1], p1[4 * j + 2], p1[4 * j + 3]);
            }
            if (r == 1) {
#pragma unroll
              for (int i = 0; i < 16; i++) p1keep[i] = p1[i];
            } else {
#pragma unroll
              for (int i = 0; i < 8; i++) l2[i] = pool4_h2(p1keep[2 * i], p1[2 * i], p1keep[2 * i + 1], p1[2 * i + 1]);
              if (live) {
                uint4* dst = reinterpret_cast<uint4*>(lvl2 + (pix * H4 + (nt * 2 + h)) * W4);
                dst[0] = make_uint4(l2[0], l2[1], l2[2], l2[3]);
                dst[1] = make_uint4(l2[4], l2[5], l2[6], l2[7]);
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i++) prev[i] = cur[i];
          }
        }
        // level 3: one 8x8 window = upper half (group 0) over lower half (group 1)
        const int xs = t & 1;
        uint4* xp = reinterpret_cast<uint4*>(xch + xs * (Smem::kXchBytes / 4) + row_in_tile * 8);
        if (g == 0) {
          if (t >= 2) mbar_wait(x_empty + xs, ((t >> 1) - 1) & 1);
          xp[0] = make_uint4(l2[0], l2[1], l2[2], l2[3]);
          xp[1] = make_uint4(l2[4], l2[5], l2[6], l2[7]);
          mbar_arrive(x_full + xs);
        } else {
          mbar_wait(x_full + xs, (t >> 1) & 1);
          const uint4 u0 = xp[0], u1 = xp[1];
          mbar_arrive(x_empty + xs);
          if (live) {
            uint4* dst = reinterpret_cast<uint4*>(lvl3 + (pix * H8 + nt) * W8);
            dst[0] = make_uint4(pool4_h2(u0.x, l2[0], u0.y, l2[1]), pool4_h2(u0.z, l2[2], u0.w, l2[3]),
                                pool4_h2(u1.x, l2[4], u1.y, l2[5]), pool4_h2(u1.z, l2[6], u1.w, l2[7]));
          }
        }
      }
      if (leader) bulk_wait_read<0>();
    } else
    for (int t = 0; t < T; t++) {
      const int m = t * kTileM + warp * 32 + lane;
      const bool live = m < HW;
      const size_t pix = (size_t)b * HW + (live ? m : 0);
      float l2row[2][W / 4];                     // level-2 values of the two halves (already fp16-rounded)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        mbar_wait(acc_full + h, t & 1);
        tc_fence_after();
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + h * NHALF;
#pragma unroll
        for (int w0 = 0; w0 < W; w0 += 16) {
          float v[4][16];
#pragma unroll
          for (int r = 0; r < 4; r++) tmem_ld16(taddr + r * W + w0, v[r]);
          tmem_ld_wait();
#pragma unroll
          for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 16; i++) v[r][i] = rnd_h(v[r][i] * scale);
            if (live) {
              uint4* dst = reinterpret_cast<uint4*>(lvl0 + (pix * H + (nt * 8 + h * 4 + r)) * W + w0);
              dst[0] = make_uint4(pack_h2(v[r][0], v[r][1]), pack_h2(v[r][2], v[r][3]), pack_h2(v[r][4], v[r][5]),
                                  pack_h2(v[r][6], v[r][7]));
              dst[1] = make_uint4(pack_h2(v[r][8], v[r][9]), pack_h2(v[r][10], v[r][11]), pack_h2(v[r][12], v[r][13]),
                                  pack_h2(v[r][14], v[r][15]));
            }
          }
          float p1[2][8];
#pragma unroll
          for (int r = 0; r < 2; r++)
#pragma unroll
            for (int i = 0; i < 8; i++)
              p1[r][i] = pool4(v[2 * r][2 * i], v[2 * r][2 * i + 1], v[2 * r + 1][2 * i], v[2 * r + 1][2 * i + 1]);
          if (live) {
#pragma unroll
            for (int r = 0; r < 2; r++) {
              uint4* dst = reinterpret_cast<uint4*>(lvl1 + (pix * H2 + (nt * 4 + h * 2 + r)) * W2 + w0 / 2);
              dst[0] = make_uint4(pack_h2(p1[r][0], p1[r][1]), pack_h2(p1[r][2], p1[r][3]), pack_h2(p1[r][4], p1[r][5]),
                                  pack_h2(p1[r][6], p1[r][7]));
            }
          }
          float p2[4];
#pragma unroll
          for (int i = 0; i < 4; i++) {
            p2[i] = pool4(p1[0][2 * i], p1[0][2 * i + 1], p1[1][2 * i], p1[1][2 * i + 1]);
            l2row[h][w0 / 4 + i] = p2[i];
          }
          if (live) {
            uint2* dst = reinterpret_cast<uint2*>(lvl2 + (pix * H4 + (nt * 2 + h)) * W4 + w0 / 4);
            dst[0] = make_uint2(pack_h2(p2[0], p2[1]), pack_h2(p2[2], p2[3]));
          }
        }
        tc_fence_before();
        mbar_arrive(acc_empty + h);
      }
      if (live) {
        uint32_t* dst = reinterpret_cast<uint32_t*>(lvl3 + (pix * H8 + nt) * W8);
#pragma unroll
        for (int i = 0; i < W8; i += 2)
          dst[i / 2] = pack_h2(pool4(l2row[0][2 * i], l2row[0][2 * i + 1], l2row[1][2 * i], l2row[1][2 * i + 1]),
                               pool4(l2row[0][2 * i + 2], l2row[0][2 * i + 3], l2row[1][2 * i + 2], l2row[1][2 * i + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kProducer) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
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

// [B][C][HW] fp16 viewed as a 3-D tensor (pixel, channel, batch); box = 64 pixels x 128 channels, 128B swizzle
static bool make_map(CUtensorMap* m, const void* base, int B, int HW) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t dims[3] = {(cuuint64_t)HW, (cuuint64_t)kC, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)HW * 2, (cuuint64_t)HW * kC * 2};
  cuuint32_t box[3] = {kBoxPix, kC, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// lvl0 [B][HW][H][W] fp16 as (w, h2, m, b); box = one image row of 128 consecutive pixels m, 128B swizzle
static bool make_out_map(CUtensorMap* m, void* base, int B, int HW, int H, int W) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t dims[4] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)HW, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)W * 2, (cuuint64_t)H * W * 2, (cuuint64_t)HW * H * W * 2};
  cuuint32_t box[4] = {(cuuint32_t)W, 1, kTileM, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            W == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int W>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, void* l0, void* l1, void* l2, void* l3, int B, int H,
                  cudaStream_t st) {
  // per device and per instantiation; a host-side table lookup, cheaper to repeat than to track
  cudaFuncSetAttribute(corr_build_kernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::kBytes);
  ProfScope ps(DROID_PROF_CORR_GEMM, st);
  DROID_LAUNCH corr_build_kernel<W><<<dim3(H / 8, B), W == 64 ? kThreads64 : kThreads, Smem::kBytes, st>>>(
      ta, tb, to, (__half*)l0, (__half*)l1, (__half*)l2, (__half*)l3, H, H * W);
  return check_launch();
}

}  // namespace cb
}  // namespace droid

using namespace droid;

extern "C" size_t droid_corr_build_workspace_bytes(int B, int C, int H, int W) {
  (void)B; (void)C; (void)H; (void)W;
  return 256;  // no scratch needed: the GEMM reads the fmaps through TMA and writes the pyramid directly
}

extern "C" int droid_corr_build(const void* fmap1, const void* fmap2, void* lvl0, void* lvl1, void* lvl2,
                                void* lvl3, int B, int C, int H, int W, void* workspace,
                                size_t workspace_bytes, void* stream) {
  (void)workspace; (void)workspace_bytes;
  if (B < 0 || H <= 0 || W <= 0) return DROID_ERR_INVALID_ARGUMENT;
  if (B == 0) return DROID_OK;
  if (!fmap1 || !fmap2 || !lvl0 || !lvl1 || !lvl2 || !lvl3) return DROID_ERR_INVALID_ARGUMENT;
  // supported shapes: 128 channels (DroidNet fnet), H % 8 == 0, W in {16,32,48,64}
  if (C != cb::kC || (H % 8) != 0 || (W % 16) != 0 || W > 64) return DROID_ERR_NOT_IMPLEMENTED;
  if (((uintptr_t)fmap1 & 15) || ((uintptr_t)fmap2 & 15) || B > 65535) return DROID_ERR_INVALID_ARGUMENT;
  CUtensorMap ta, tb, to;
  if (!cb::make_map(&ta, fmap1, B, H * W) || !cb::make_map(&tb, fmap2, B, H * W) ||
      !cb::make_out_map(&to, lvl0, B, H * W, H, W))
    return DROID_ERR_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  switch (W) {
    case 16: return cb::launch<16>(ta, tb, to, lvl0, lvl1, lvl2, lvl3, B, H, st);
    case 32: return cb::launch<32>(ta, tb, to, lvl0, lvl1, lvl2, lvl3, B, H, st);
    case 48: return cb::launch<48>(ta, tb, to, lvl0, lvl1, lvl2, lvl3, B, H, st);
    default: return cb::launch<64>(ta, tb, to, lvl0, lvl1, lvl2, lvl3, B, H, st);
  }
}
