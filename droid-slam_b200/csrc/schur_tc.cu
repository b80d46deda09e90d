// Schur complement of the per-pixel inverse depths on the tensor cores (tcgen05 / TMEM), sm_100a.
// Replaces, for source frames with at most kSchurTcMaxDeg = 31 outgoing edges, the CUDA-core Gram product of
// ba_schur_kernel (ba.cu) and with it the reference's EEt6x6 / Ev6x1 kernels (src/droid_kernels.cu:1001-1091).
//
// For one source frame k with targets j_1..j_deg the Schur terms are
//     S[a,b] = sum_p E_a[p] Q[p] E_b[p]^T          (6x6 blocks, a,b in {k, j_1..j_deg})
//     v_S[a] = sum_p E_a[p] Q[p] w[p]
// The source pose's rows are a LINEAR image of the edge rows, E_k[p] = sum_e N_e E_{j_e}[p] with the per-edge matrix
// N_e = -Ad(G_e)^T (ba_lin.cuh), so they never enter the per-pixel work: with F[6e+a][p] = sqrt(Q[p]) E_{j_e}[p][a]
// and one extra row F[6 deg][p] = sqrt(Q[p]) w[p],
//     G = F F^T  (M x M, M = 6 deg + 1 <= 128, K = number of pixels)
// holds every edge-edge block and v_S of the targets, and the source pose's blocks follow once per frame from G:
//     S[k,j_e'] = sum_e N_e G[e,e'],   S[k,k] = sum_{e,e'} N_e G[e,e'] N_e'^T,   v_S[k] = sum_e N_e G[e,w].
// Frames with 22..31 edges: the edges are cut into three groups and one CTA takes each PAIR of groups (<= 21 edges);
// a pair of rows is "owned" by exactly one of the three CTAs (cross-group pairs by the CTA of that pair of groups,
// pairs inside a group by the CTA that lists the group first), and a CTA adds only what it owns.
// G is a GEMM with K = 3072 .. 5760, which belongs on the tensor pipe.  fp32 accuracy needs two things:
//  * the 3xTF32 split  F = F_hi + F_lo (both rounded to tf32),  G = F_hi F_hi^T + F_hi F_lo^T + F_lo F_hi^T;
//  * SHORT accumulation chains: the TMEM accumulator truncates, so the error of a sum grows linearly with the number
//    of chained MMAs (measured: 1.8e-6 relative over 1152 MMAs, 2.6e-7 over 24).  Every chunk of 32 pixels (12 MMAs)
//    therefore starts a fresh accumulator, and the chunks are summed with ordinary fp32 adds in registers.
//
// Warp roles (512 threads, one CTA per SM, CTA = (pixel slab, source frame, pair of edge groups)):
//   warps 0-7   builders: stream the packed E_j of every (pixel, edge) (left behind by ba_accum_kernel) into
//               registers one pass ahead, unpack, scale by sqrt(Q), split into tf32 hi + lo and store straight into
//               the K-major UMMA operand tiles (SWIZZLE_128B, one 16-byte store per row and array; no fp32 staging);
//   warp 13     one thread issues the 12 tcgen05.mma of a chunk into TMEM accumulator (chunk & 3);
//   warps 8-11, 14-15  drainers: TMEM -> registers, fp32 accumulation of half a row of G per thread;
//   epilogue    (all warps) G goes to shared memory, the source pose's rows are derived from it, and everything is
//               subtracted from the dense reduced system with fp64 atomics, lanes along a row (same block placement
//               rules as flush_block in ba.cu).
// Four operand stages and four TMEM accumulators: building chunk c+1, the MMAs of chunk c and the drain of chunk c-1
// overlap.
#include <cstdint>

#include "ba_lin.cuh"
#include "common.cuh"

namespace droid {
namespace stc {

static constexpr int kThreads = 512;             // 8 builder warps, 6 drain warps, 1 MMA warp, 1 spare (128 registers)
static constexpr int kT = 32;                    // pixels per chunk = one 128-byte row of a K-major operand tile
static constexpr int kHalf = 128 * kT * 4;       // one operand array (F_hi or F_lo) of one stage: 128 rows x 128 B
static constexpr int kStage = 2 * kHalf;         // 32 KB
static constexpr int kStages = 4;                // = chunks of one build pass (128 pixels)
static constexpr int kChain = 1;                 // chunks summed in one TMEM accumulator before it is drained (2 was
                                                 // measured: 24-MMA chains, 1.5 % faster, 1.4x the rounding error)
static constexpr int kTB = kStages * kT;
static constexpr int kELd = 132;                 // row stride (floats) of G in the epilogue
static constexpr int kGBytes = 128 * kELd * 4;
static constexpr int kTsOff = 72 * 1024;         // epilogue scratch behind G, inside the (then idle) operand stages
static constexpr int kMetaOff = kStages * kStage;
static constexpr int kSmem = kMetaOff + 8192 + 1024;
static_assert(kGBytes <= kTsOff && kTsOff + 5 * 6 * 128 * 4 <= kMetaOff, "epilogue scratch must fit in the stages");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
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
// same, for the warps that only wait for other warps' work (MMA issuer, drainers): back off between polls so that the
// polling does not take issue slots from the builder warps on the same scheduler
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(32);
  }
}
// shared-memory matrix descriptor, K-major 32-bit operands, SWIZZLE_128B (layout type 2): a tile is [128 rows][32
// pixels] with 128-byte rows, atoms of 8 rows (1024 B, SBO) whose 16-byte chunks are XOR-ed with (row & 7).  One MMA
// consumes K = 8 pixels = 32 bytes of every row: the k-th MMA of a chunk starts 32 k bytes into the tile.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;                  // LBO: unused for swizzled K-major operands
  d |= (uint64_t)(1024 >> 4) << 32;        // SBO
  d |= (uint64_t)1 << 46;                  // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                  // SWIZZLE_128B
  return d;
}
// instruction descriptor, kind::tf32: D = f32, A = B = tf32, both K-major, M = N = 128
__device__ __forceinline__ uint32_t umma_idesc() {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
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
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// round-to-nearest (ties away) to tf32 = what cvt.rna.tf32.f32 does, with two full-rate integer instructions
// (the conversion instruction runs at a fraction of the ALU rate and made the pack pass cost as much as the MMAs)
__device__ __forceinline__ uint32_t tf32_rna(float x) { return (__float_as_uint(x) + 0x1000u) & 0xFFFFE000u; }

template <bool kVec>
__global__ void __launch_bounds__(kThreads, 1) ba_schur_tc_kernel(
    const float* __restrict__ ejc, const int64_t* __restrict__ jj, const float* __restrict__ gij, int* __restrict__ status,
    const int* __restrict__ kx, const int* __restrict__ eptr, const int* __restrict__ eidx,
    const float* __restrict__ Qi, const float* __restrict__ Wi, double* __restrict__ A,
    double* __restrict__ bvec, int HW, int wd, int slab, int t0, int P, int n_eta, int frame_lo,
    int frame_hi, const int* __restrict__ ppos) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int m = blockIdx.y;
  const int Mk = status[1];
  if (m >= Mk || m >= n_eta) return;
  const int k = kx[m];
  if (k < frame_lo || k >= frame_hi) return;
  const int p0 = blockIdx.x * slab;
  const int p1 = min(HW, p0 + slab);
  if (p0 >= p1) return;
  const int e0 = eptr[m], e1 = eptr[m + 1];
  const int deg = e1 - e0;
  if (deg == 0 || deg > kSchurTcMaxDeg) return;     // nothing to do / handled by ba_schur_kernel
  // edge groups: one group (this CTA takes all edges) or three (this CTA takes groups z and z+1 mod 3)
  const int z = blockIdx.z;
  const int ng = deg <= kSchurTcMaxU ? 1 : 3;
  if (z >= ng) return;
  const int gbase = deg / 3, grem = deg - 3 * gbase;
  const int g1 = z, g2 = (z + 1) % 3;
  const int start1 = ng == 1 ? 0 : g1 * gbase + min(g1, grem), start2 = g2 * gbase + min(g2, grem);
  const int nfirst = ng == 1 ? deg : gbase + (g1 < grem ? 1 : 0);
  const int nu = ng == 1 ? deg : nfirst + gbase + (g2 < grem ? 1 : 0);      // edges of this CTA (<= kSchurTcMaxU)
  const int R = 6 * nu;                  // rows of E; row R of F carries sqrt(Q) w

  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* Gs = reinterpret_cast<float*>(smem);                        // epilogue: G [128][kELd] (stages idle by then)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kMetaOff);
  uint64_t* full = bars;        // [4]  operands of a chunk in place      (64 builder arrivals: the 2 warps of a stage)
  uint64_t* empty = bars + 4;   // [4]  operands consumed                 (tcgen05.commit)
  uint64_t* gdone = bars + 8;   // [4]  accumulator of a chunk complete   (tcgen05.commit)
  uint64_t* tfree = bars + 12;  // [4]  accumulator drained               (192 drainer arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  int* s_e = reinterpret_cast<int*>(bars + 18);
  int* s_jx = s_e + 24;
  int* pose = s_jx + 24;
  float* s_g = reinterpret_cast<float*>(bars + 18 + 64);   // 16-byte aligned: [nu][kEdgeStride]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < kStages; i++) {
      mbar_init(full + i, 64);
      mbar_init(empty + i, 1);
      mbar_init(gdone + i, 1);
      mbar_init(tfree + i, 192);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 13) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // operand rows > R are never written and must read as zero
  for (int i = tid; i < kStages * kStage / 16; i += kThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  auto edge_of = [&](int i) { return eidx[e0 + (i < nfirst ? start1 + i : start2 + (i - nfirst))]; };
  for (int i = tid; i < nu; i += kThreads) {
    const int e = edge_of(i);
    s_e[i] = e;
    s_jx[i] = (int)jj[e];
  }
  for (int i = tid; i < nu * kEdgeStride; i += kThreads)
    s_g[i] = gij[(size_t)edge_of(i / kEdgeStride) * kEdgeStride + i % kEdgeStride];
  for (int d = tid; d <= nu; d += kThreads) {       // pose[0] = source pose, pose[1 + i] = target of edge i
    int pz = (d == 0) ? k - t0 : (int)jj[edge_of(d - 1)] - t0;
    if (pz < 0 || pz >= P) pz = -1;
    else if (ppos != nullptr) pz = ppos[pz];
    pose[d] = pz;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int nchunk = (p1 - p0 + kT - 1) / kT;

  if (warp == 13) {
    if (lane == 0) {
      // ---------------- MMA issuer ----------------
      const uint32_t idesc = umma_idesc();
      for (int c = 0; c < nchunk; c++) {
        const int s = c & (kStages - 1);          // operand stage
        const int g = c / kChain, a = g & 3;      // accumulator group, TMEM accumulator (4 x 128 columns)
        const bool first = (c % kChain) == 0, last = (c % kChain) == kChain - 1 || c == nchunk - 1;
        // (no back-off here: with builders and tensor pipe balanced, every hand-off of the issuer is on the critical path)
        mbar_wait(full + s, (c >> 2) & 1);
        if (first && g >= 4) mbar_wait(tfree + a, ((g >> 2) - 1) & 1);     // accumulator drained
        tc_fence_after();
        const uint32_t hi = smem_u32(smem + s * kStage), lo = hi + kHalf;
        const uint32_t d_t = tmem + a * 128;
#pragma unroll
        for (int kk = 0; kk < kT / 8; kk++) {
          const uint64_t f_hi = umma_desc(hi + kk * 32), f_lo = umma_desc(lo + kk * 32);
          umma_tf32(d_t, f_hi, f_hi, idesc, (kk > 0 || !first) ? 1u : 0u);
          umma_tf32(d_t, f_hi, f_lo, idesc, 1u);
          umma_tf32(d_t, f_lo, f_hi, idesc, 1u);
        }
        umma_commit(empty + s);
        if (last) umma_commit(gdone + a);
      }
    }
  } else if (warp < 8) {
    // ---------------- F builders ----------------
    // A build pass covers kTB = 128 pixels = 4 chunks = the 4 operand stages.  Warp w builds chunk (w & 3) of every
    // pass; inside the warp a lane owns 4 consecutive pixels (one 16-byte unit of a K-major row: kq = lane & 7) of the
    // edges i = slice, slice + 8, slice + 16 with slice = 4 (w >> 2) + (lane >> 3).  Per (4 pixels, edge) the packed
    // E_j (ba_lin.cuh: 5 floats per pixel-edge, left behind by ba_accum_kernel) arrives as five 16-byte loads, is
    // unpacked, scaled by sqrt(Q), split into tf32 hi + lo and stored straight into the swizzled operand tile: one
    // 16-byte store per row and array, a quarter-warp covers one whole 128-byte row (conflict-free).  The loads of a
    // pass are issued one pass ahead into registers.
    const int sub = warp & 3, kq = lane & 7, slice = 4 * (warp >> 2) + (lane >> 3);
    unsigned char* stage = smem + sub * kStage;
    const int npass = (p1 - p0 + kTB - 1) / kTB;
    constexpr int kMaxI = (kSchurTcMaxU + 7) / 8;     // edges per thread
    float4 r_c[kMaxI][kEjComp];
    float4 n_q = make_float4(0.f, 0.f, 0.f, 0.f), n_w = n_q;
    // grids whose pixel count is not a multiple of 4 (41x73, ...): planes start at odd offsets and the last group of a
    // frame is partial, so the 16-byte loads become four masked scalar ones (pixels >= p1 read as zero -> zero rows)
    auto ld4 = [&](const float* src, int p, bool stream) {
      float4 v;
      if constexpr (kVec) {
        v = stream ? __ldcs(reinterpret_cast<const float4*>(src)) : __ldg(reinterpret_cast<const float4*>(src));
      } else {
        v.x = p + 0 < p1 ? __ldg(src + 0) : 0.f;
        v.y = p + 1 < p1 ? __ldg(src + 1) : 0.f;
        v.z = p + 2 < p1 ? __ldg(src + 2) : 0.f;
        v.w = p + 3 < p1 ? __ldg(src + 3) : 0.f;
      }
      return v;
    };
    auto fetch = [&](int it, int p, bool ok) {
      const int i = slice + 8 * it;
      if (ok && i < nu) {
        const float* src = ejc + (size_t)s_e[i] * kEjComp * HW + p;
#pragma unroll
        for (int q = 0; q < kEjComp; q++) r_c[it][q] = ld4(src + (size_t)q * HW, p, true);
        if (p + kTB < p1) {      // and the pass after that on its way into L2
#pragma unroll
          for (int q = 0; q < kEjComp; q++) asm volatile("prefetch.global.L2 [%0];" ::"l"(src + (size_t)q * HW + kTB));
        }
      }
    };
    auto fetch_pix = [&](int p, bool ok) {
      n_q = ok ? ld4(Qi + (size_t)m * HW + p, p, false) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (slice == 7) n_w = ok ? ld4(Wi + (size_t)m * HW + p, p, false) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    // 16-byte store of 4 pixels of one row: hi and lo parts at byte offset `off` of the stage's two arrays
    auto put_row = [&](int off, float x0, float x1, float x2, float x3) {
      uint4 h, l;
      h.x = tf32_rna(x0); h.y = tf32_rna(x1); h.z = tf32_rna(x2); h.w = tf32_rna(x3);
      l.x = tf32_rna(x0 - __uint_as_float(h.x)); l.y = tf32_rna(x1 - __uint_as_float(h.y));
      l.z = tf32_rna(x2 - __uint_as_float(h.z)); l.w = tf32_rna(x3 - __uint_as_float(h.w));
      *reinterpret_cast<uint4*>(stage + off) = h;
      *reinterpret_cast<uint4*>(stage + kHalf + off) = l;
    };
    // swizzled offset of row r: 8-row atom, row inside the atom, 16-byte unit XOR-ed with it
    auto row_off = [&](int r) { const int r7 = r & 7; return ((r >> 3) << 10) + (r7 << 7) + ((kq ^ r7) << 4); };
    // rows of this lane's edges: r = 6 (slice + 8 it) + n = (6 slice + n) + 48 it, and 48 rows are 6 whole atoms
    int off_n[6];
#pragma unroll
    for (int n = 0; n < 6; n++) off_n[n] = row_off(6 * slice + n);
    const int off_w = row_off(R);
#pragma unroll
    for (int it = 0; it < kMaxI; it++)
#pragma unroll
      for (int q = 0; q < kEjComp; q++) r_c[it][q] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int pq = p0 + 32 * sub + 4 * kq;             // this lane's 4 pixels in pass 0
    fetch_pix(pq, pq < p1);
#pragma unroll
    for (int it = 0; it < kMaxI; it++) fetch(it, pq, pq < p1);
    for (int ps = 0; ps < npass; ps++) {
      const int c = ps * kStages + sub;
      if (c >= nchunk) break;
      const int p = pq + ps * kTB;
      const bool valid = p < p1;
      const int pn = p + kTB;
      const bool valid_n = pn < p1;
      const float sq0 = valid ? sqrtf(n_q.x) : 0.f, sq1 = valid ? sqrtf(n_q.y) : 0.f;
      const float sq2 = valid ? sqrtf(n_q.z) : 0.f, sq3 = valid ? sqrtf(n_q.w) : 0.f;
      const float4 wv = n_w;
      fetch_pix(pn, valid_n);
      if (ps > 0) mbar_wait(empty + sub, (ps - 1) & 1);          // the MMAs of the previous pass have read this stage
#pragma unroll
      for (int it = 0; it < kMaxI; it++) {
        const int i = slice + 8 * it;
        if (i < nu) {
          float4 pk[kEjComp];
#pragma unroll
          for (int q = 0; q < kEjComp; q++) pk[q] = r_c[it][q];
          fetch(it, pn, valid_n);
          float e0v[6], e1v[6], e2v[6], e3v[6];
          {
            const float c0[5] = {pk[0].x, pk[1].x, pk[2].x, pk[3].x, pk[4].x};
            const float c1[5] = {pk[0].y, pk[1].y, pk[2].y, pk[3].y, pk[4].y};
            const float c2[5] = {pk[0].z, pk[1].z, pk[2].z, pk[3].z, pk[4].z};
            const float c3[5] = {pk[0].w, pk[1].w, pk[2].w, pk[3].w, pk[4].w};
            ej_unpack(c0, e0v); ej_unpack(c1, e1v); ej_unpack(c2, e2v); ej_unpack(c3, e3v);
          }
#pragma unroll
          for (int n = 0; n < 6; n++) put_row(off_n[n] + it * 6144, sq0 * e0v[n], sq1 * e1v[n], sq2 * e2v[n], sq3 * e3v[n]);
        }
      }
      if (slice == 7) put_row(off_w, sq0 * wv.x, sq1 * wv.y, sq2 * wv.z, sq3 * wv.w);
      fence_async_smem();
      mbar_arrive(full + sub);
    }
  } else if (warp != 12) {
    // ---------------- drainers: half a row of G per thread (TMEM lane quarter = warp % 4) ----------------
    // warps 8-11 take columns 0-63 of quarters 0-3, warps 14-15 columns 64-127 of quarters 2-3 (columns <= row only)
    const int qd = warp & 3, hh = warp >> 3 & (warp >> 2) & 1;
    const int r = 32 * qd + lane;
    float acc[64];
#pragma unroll
    for (int j = 0; j < 64; j++) acc[j] = 0.f;
    const int ngroup = (nchunk + kChain - 1) / kChain;
    for (int g = 0; g < ngroup; g++) {
      const int s = g & 3;
      mbar_wait_relaxed(gdone + s, (g >> 2) & 1);
      tc_fence_after();
      {
        const uint32_t taddr = tmem + s * 128 + 64 * hh + ((uint32_t)(32 * qd) << 16);
#pragma unroll
        for (int part = 0; part < 2; part++) {
          float v[32];
          tmem_ld32(taddr + 32 * part, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j++) acc[32 * part + j] += v[j];
        }
      }
      tc_fence_before();
      mbar_arrive(tfree + s);
    }
    // ---- hand the row to the epilogue: G[r][64 hh .. 64 hh + 63] -> shared memory (the operand stages are idle: this
    // thread has just waited for the commit behind the last chunk's MMAs, which covers every MMA before it)
    float4* grow = reinterpret_cast<float4*>(Gs + r * kELd + 64 * hh);
#pragma unroll
    for (int j = 0; j < 16; j++) grow[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
  }
  tc_fence_before();
  __syncthreads();
  // ---------------- epilogue (all warps): G in shared memory (lower triangle valid), operand stages free ----------------
  {
    float* Ts = reinterpret_cast<float*>(smem + kTsOff);   // [4][6][128] partial sums of T = sum_e N_e G[e, :]
    float* Tf = Ts + 4 * 6 * 128;                     // [6][128]
    const int n6 = 6 * P;
    const int ps = pose[0];
    auto own = [&](int a, int b) { return a < nfirst || b < nfirst; };      // edge indices inside this CTA's list
    auto place = [&](int pa, int n, int pb, int mm, bool same_rows, double val) {
      // element (n, mm) of the 6x6 block E_a Q E_b^T -> lower triangle of H - S (flush_block in ba.cu)
      if (pa > pb) atomicAdd(&A[(size_t)(6 * pa + n) * n6 + 6 * pb + mm], val);
      else if (pa < pb) atomicAdd(&A[(size_t)(6 * pb + mm) * n6 + 6 * pa + n], val);
      else if (same_rows) atomicAdd(&A[(size_t)(6 * pa + n) * n6 + 6 * pa + mm], val);          // caller passes mm <= n
      else atomicAdd(&A[(size_t)(6 * pa + max(n, mm)) * n6 + 6 * pa + min(n, mm)], n == mm ? 2.0 * val : val);
    };
    // (1) partial T: thread = (column c, quarter q of the edges)
    {
      const int c = tid & 127, q = tid >> 7;
      float t[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (c <= R) {
        const int cb = c / 6;
        for (int e = q; e < nu; e += 4) {
          if (!(c < R ? own(e, cb) : e < nfirst)) continue;
          float gv[6];
#pragma unroll
          for (int a = 0; a < 6; a++) {
            const int r = 6 * e + a;
            gv[a] = r >= c ? Gs[r * kELd + c] : Gs[c * kELd + r];
          }
          const float* N = s_g + kEdgeStride * e + 8;
#pragma unroll
          for (int n = 0; n < 6; n++)
#pragma unroll
            for (int a = 0; a < 6; a++) t[n] = fmaf(N[n * 6 + a], gv[a], t[n]);
        }
      }
#pragma unroll
      for (int n = 0; n < 6; n++) Ts[(q * 6 + n) * 128 + c] = t[n];
    }
    // (2) edge-edge blocks and v_S of the targets: one row of G per warp step, lanes along the row
    for (int r = warp; r <= R; r += kThreads / 32) {
      const int rb = r / 6, n = r - 6 * rb;
      const int pa = r < R ? pose[1 + rb] : 0;
      for (int c = lane; c <= r && c < R; c += 32) {
        const int cb = c / 6, mm = c - 6 * cb;
        const int pb = pose[1 + cb];
        if (pb < 0) continue;
        const double val = -(double)Gs[r * kELd + c];
        if (r == R) { if (cb < nfirst) atomicAdd(&bvec[6 * pb + mm], val); continue; }       // v_S
        if (pa < 0 || !own(rb, cb)) continue;
        place(pa, n, pb, mm, rb == cb, val);
      }
    }
    __syncthreads();
    // (3) the source pose against every target, and its v_S
    for (int idx = tid; idx < 6 * 128; idx += kThreads) {
      const int n = idx >> 7, c = idx & 127;
      const float v = (Ts[n * 128 + c] + Ts[(6 + n) * 128 + c]) + (Ts[(12 + n) * 128 + c] + Ts[(18 + n) * 128 + c]);
      Tf[n * 128 + c] = v;
      if (c > R || ps < 0) continue;
      if (c == R) { atomicAdd(&bvec[6 * ps + n], -(double)v); continue; }
      const int cb = c / 6, mm = c - 6 * cb;
      const int pb = pose[1 + cb];
      if (pb < 0) continue;
      place(ps, n, pb, mm, false, -(double)v);
    }
    __syncthreads();
    // (4) the source pose's own block: S[k,k] = sum_c T[:, c] N_{e(c)}[:, a(c)]^T  (symmetric; lower triangle)
    if (tid < 36 && ps >= 0) {
      const int n = tid / 6, mm = tid - 6 * n;
      if (mm <= n) {
        float sacc = 0.f;
        for (int c = 0; c < R; c++) {
          const int e = c / 6, a = c - 6 * e;
          sacc = fmaf(Tf[n * 128 + c], s_g[kEdgeStride * e + 8 + mm * 6 + a], sacc);
        }
        atomicAdd(&A[(size_t)(6 * ps + n) * n6 + 6 * ps + mm], -(double)sacc);
      }
    }
  }
  if (warp == 13) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

}  // namespace stc

int launch_schur_tc(const float* ejc, const int64_t* jj, const float* gij, int* status, const int* kx, const int* eptr, const int* eidx,
                    const float* Qi, const float* Wi, double* A, double* bvec, int HW, int wd, int slab, int S, int t0,
                    int P, int n_eta, int frame_lo, int frame_hi, const int* ppos, cudaStream_t st) {
  if (dev_ctx().first(kOnceSchurTc)) {
    cudaFuncSetAttribute(stc::ba_schur_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, stc::kSmem);
    cudaFuncSetAttribute(stc::ba_schur_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, stc::kSmem);
  }
  if ((HW & 3) == 0) {
    DROID_LAUNCH stc::ba_schur_tc_kernel<true><<<dim3(S, n_eta, 3), stc::kThreads, stc::kSmem, st>>>(
        ejc, jj, gij, status, kx, eptr, eidx, Qi, Wi, A, bvec, HW, wd, slab, t0, P, n_eta, frame_lo, frame_hi, ppos);
  } else {     // pixel count not a multiple of 4: masked scalar loads in the builders
    DROID_LAUNCH stc::ba_schur_tc_kernel<false><<<dim3(S, n_eta, 3), stc::kThreads, stc::kSmem, st>>>(
        ejc, jj, gij, status, kx, eptr, eidx, Qi, Wi, A, bvec, HW, wd, slab, t0, P, n_eta, frame_lo, frame_hi, ppos);
  }
  return check_launch();
}

}  // namespace droid
