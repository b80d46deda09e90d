// Device-side dense Cholesky solve in fp64 for the reduced camera system of droid_backends.ba.
// Replaces the host Eigen::SimplicialLLT<SparseMatrix<double>> factor+solve of the reference
// (src/droid_kernels.cu:1192-1213: D2H of all blocks, one CPU thread, H2D of dx).
//
// Storage: A is [n+1][n] row-major fp64; rows 0..n-1 hold the SPD matrix (LOWER triangle is read and
// overwritten by L), row n holds the right-hand side b.  Treating b as one more matrix row makes the
// forward substitution L y = b fall out of the right-looking factorisation for free (the panel's
// triangular solve and the trailing update are applied to row n like to any other row).  A separate
// block-backward sweep then solves L^T x = y in place, so x ends up in row n.
//   n <= kSmallMax : one CTA does everything in shared memory (frontend / trajectory-filler windows);
//   otherwise      : blocked right-looking factorisation, NB = 64, two launches per panel
//                    (potrf+trsm, then the trailing SYRK on all SMs), one launch per backward block.
// Failure (non-positive or NaN pivot) sets *info != 0; the caller zeroes dx like the reference
// (src/droid_kernels.cu:1207-1210).  No host synchronisation anywhere.
#include "common.cuh"
#include <cstdio>
#include <cstring>
#include <cstdlib>

namespace droid {

static constexpr int NB = 64;
static constexpr int kSmallMax = kCholSmallMax;  // (n+1)*(n+1)*8 bytes <= ~197 KB of shared memory
__host__ __device__ inline int chol_small_ld(int n) {  // >= n, == 2 (mod 4)
  return ((n + 1) & ~3) + 2 >= n ? (((n + 1) & ~3) + 2) : (((n + 1) & ~3) + 6);
}
static size_t chol_small_smem(int n) {
  return ((size_t)(n + 1) * chol_small_ld(n) + 8 * 8 + 160 + 160) * sizeof(double);
}
static constexpr int kSmallSmemMax = (157 * 158 + 64 + 320) * 8;

// ------------------------------------------------------------------------------------------------
// small systems: whole solve in one CTA
// ------------------------------------------------------------------------------------------------
// One CTA, one matrix row per thread (row n = rhs), S [n+1][ld] in shared memory with ld == 2 (mod 4) so that a
// thread's own row can be read two columns at a time (LDS.128) without bank conflicts.  Left-looking in blocks of
// KB = 8 columns:
//   (i)   every row subtracts the contribution of all finished columns from its 8 entries of the block
//         (a[k] -= sum_{j<c0} L[r][j] L[c0+k][j]; own row vectorised, the 8 diagonal rows are broadcasts);
//   (ii)  the 8 diagonal rows publish their entries (the updated 8x8 diagonal block);
//   (iii) EVERY thread factorises that 8x8 block redundantly in registers and eliminates its own row in the same
//         pass - no communication, no barrier inside the block; the serial chain is 8 x (reciprocal + 2 DFMA);
//   (iv)  L entries go back to S.
// Two barriers per 8 columns.  Treating the rhs as row n solves L y = b on the way.  Backward sweep L^T x = y:
// thread t keeps y_t in a register; 32 columns at a time are solved inside one warp with shuffles, then all
// earlier threads subtract that block's contribution (row c of L is contiguous in shared memory).
// A is NOT written back on this path (only x, into row n); the blocked path below leaves L in A.
static constexpr int KB = 8;

__device__ __forceinline__ double shfl_f64(double v, int src) {
  return __hiloint2double(__shfl_sync(0xffffffffu, __double2hiint(v), src),
                          __shfl_sync(0xffffffffu, __double2loint(v), src));
}
__device__ __forceinline__ double shfl_xor_f64(double v, int m) {
  return __hiloint2double(__shfl_xor_sync(0xffffffffu, __double2hiint(v), m),
                          __shfl_xor_sync(0xffffffffu, __double2loint(v), m));
}

// damp != 0: the Levenberg-Marquardt damping of ba (diag += ep + lm * diag, src/droid_kernels.cu:1197) is applied to the
// copy in shared memory and *info is written in either case (0 / 1), so that ba needs neither a damping kernel nor a
// memset in front of the solve of a frontend-sized system.
__global__ void __launch_bounds__(160) chol_small_kernel(double* __restrict__ A, int n,
                                                         int* __restrict__ info, int damp, float lm, float ep) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int R = n + 1;
  const int ld = chol_small_ld(n);
  double* S = reinterpret_cast<double*>(smem_raw);
  double* Dsh = S + (size_t)R * ld;  // [KB][KB]
  double* rsd = Dsh + KB * KB;       // [160]
  double* xs = rsd + 160;            // [160]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  __shared__ int fail;
  if (tid == 0) fail = 0;
  for (int r = warp; r < R; r += nwarps) {
    const int lim = (r == n) ? n : r + 1;
    const double* src = A + (size_t)r * n;
    const unsigned dst = (unsigned)__cvta_generic_to_shared(S + r * ld);
    for (int c = lane; c < lim; c += 32)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst + c * 8), "l"(src + c));
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (damp) {
    for (int r = tid; r < n; r += blockDim.x) {
      double* d = S + r * ld + r;
      *d += (double)ep + (double)lm * (*d);
    }
    __syncthreads();
  }

  const bool is_rhs = tid + 0 == n;  // re-evaluated per block below (row index changes with c0)
  (void)is_rhs;
  for (int c0 = 0; c0 < n; c0 += KB) {
    const int kb = min(KB, n - c0);
    const int r = c0 + tid;
    const bool active = r < R;
    double a[KB];
    if (active) {
      const double* own = S + r * ld;
#pragma unroll
      for (int k = 0; k < KB; k++) a[k] = (k < kb && (c0 + k <= r || r == n)) ? own[c0 + k] : 0.0;
      // (i) left-looking update with the finished columns 0..c0-1
      const double2* own2 = reinterpret_cast<const double2*>(own);
#pragma unroll 2
      for (int j = 0; j < c0; j += 2) {
        const double2 o = own2[j >> 1];
#pragma unroll
        for (int k = 0; k < KB; k++) {
          if (k < kb) {
            const double2 d = *reinterpret_cast<const double2*>(S + (c0 + k) * ld + j);
            a[k] = fma(-o.x, d.x, a[k]);
            a[k] = fma(-o.y, d.y, a[k]);
          }
        }
      }
      // (ii) the diagonal rows publish the updated diagonal block
      if (tid < kb) {
#pragma unroll
        for (int k = 0; k < KB; k++) Dsh[tid * KB + k] = a[k];
      }
    }
    __syncthreads();
    if (active) {
      // (iii) redundant factorisation of the 8x8 block + elimination of the own row
      double D[KB][KB];
#pragma unroll
      for (int q = 0; q < KB; q++)
#pragma unroll
        for (int k = 0; k <= q; k++) D[q][k] = (q < kb) ? Dsh[q * KB + k] : (q == k ? 1.0 : 0.0);
      double rs[KB];
      bool bad = false;
#pragma unroll
      for (int C = 0; C < KB; C++) {
        const double d = D[C][C];
        bad |= !(d > 0.0);
        const double inv = 1.0 / d;
        const double m = a[C] * inv;
#pragma unroll
        for (int q = C + 1; q < KB; q++) {
          const double mq = D[q][C] * inv;
#pragma unroll
          for (int k = C + 1; k <= q; k++) D[q][k] = fma(-mq, D[k][C], D[q][k]);
          a[q] = fma(-m, D[q][C], a[q]);
        }
        rs[C] = sqrt(inv);
      }
      // (iv) L entries of this block
      double* own = S + r * ld;
#pragma unroll
      for (int k = 0; k < KB; k++)
        if (k < kb && (c0 + k <= r || r == n)) own[c0 + k] = a[k] * rs[k];
      if (tid < kb) {
        double mine = 0.0;
#pragma unroll
        for (int k = 0; k < KB; k++) mine = (k == tid) ? rs[k] : mine;
        rsd[c0 + tid] = mine;
      }
      if (tid == 0 && bad) fail = 1;
    }
    __syncthreads();
  }

  // backward sweep
  double y = (tid < n) ? S[n * ld + tid] : 0.0;
  const double myrs = (tid < n) ? rsd[tid] : 0.0;
  double x = 0.0;
  for (int w = (n - 1) >> 5; w >= 0; w--) {
    const int b0 = w << 5;
    if (warp == w) {
      double Lr[32];
#pragma unroll
      for (int c = 0; c < 32; c++) Lr[c] = (b0 + c < n && lane < c) ? S[(b0 + c) * ld + b0 + lane] : 0.0;
#pragma unroll
      for (int c = 31; c >= 0; c--) {
        const double xc = shfl_f64(y * myrs, c);
        if (lane == c) x = xc;
        y = fma(-Lr[c], xc, y);
      }
      xs[b0 + lane] = x;
    }
    __syncthreads();
    if (warp < w) {
      const int cend = min(32, n - b0);
      double y0 = 0.0, y1 = 0.0;
#pragma unroll 4
      for (int c = 0; c + 1 < cend; c += 2) {
        y0 = fma(S[(b0 + c) * ld + tid], xs[b0 + c], y0);
        y1 = fma(S[(b0 + c + 1) * ld + tid], xs[b0 + c + 1], y1);
      }
      if (cend & 1) y0 = fma(S[(b0 + cend - 1) * ld + tid], xs[b0 + cend - 1], y0);
      y -= y0 + y1;
    }
  }
  if (tid < n) A[(size_t)n * n + tid] = x;
  if (tid == 0 && (fail || damp)) *info = fail ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// blocked path.  Outer panels of PB = 128 columns (depth of the trailing update), each factorised as two
// sub-panels of NB = 64 columns by chol_panel_kernel; then one trailing SYRK on all SMs.
//
// chol_panel_kernel<PRT>: every CTA holds the 64 rows of the diagonal block (threads 0..63, redundantly in
// every CTA) plus PRT trailing rows (threads 64..), ONE MATRIX ROW PER THREAD IN REGISTERS (a[64]).  All rows
// run the same right-looking elimination in lockstep:
//     for c:  m = a[c] / d_c ;  a[q] -= m * col_c[q]  (q > c)
// where col_c[q] = (un-normalised) entry (q, c) of the diagonal block, published through shared memory by
// the diagonal rows one column ahead, with the reciprocal pivot computed by the publishing thread only.
// One __syncthreads per column, no triangular-solve phase: the trailing rows finish together with the
// 64x64 factorisation.  L = a[c] / sqrt(d_c) at the end.
//   * PRT = 64: the per-column cost of a CTA is the dependent pivot chain PLUS the fp64 issue of its rows' bulk
//     update (rows x remaining columns / 64 lanes per clock); with 256 rows per CTA the bulk was as long as the
//     chain (343 cycles per pivot), with 128 it hides behind it.  The panel kernels run a handful of CTAs
//     (the rows inside the envelope), so three times as many CTAs cost nothing.
//   * prev = 64 (second half of an outer panel): the rows first apply the rank-64 update of the first half to
//     their 64 entries, a[c] -= sum_k L[r][j0-64+k] L[j0+c][j0-64+k], from two staged blocks -- this replaces a
//     separate strip-SYRK launch (its own staging, C-tile round trip and two launch boundaries per 128 columns).
//   * linv != nullptr: one extra CTA eliminates the rows of the 64x64 identity, which yields I L_KK^-T: the rows
//     of the inverse transposed diagonal block.  The backward sweep then solves a block with a 64x64
//     matrix-vector product instead of a 64-step dependent substitution.
// The rhs (row n of A) is just one more trailing row, which is how L y = b gets solved for free.
// ------------------------------------------------------------------------------------------------
static constexpr int PB = 128;
static constexpr int PRT = 64;                 // trailing rows per panel CTA
static constexpr int kPanelThreads = 64 + PRT;
#ifdef PANEL_TS
__device__ long long g_ts[16];
#define PTS(i) if (blockIdx.x == 0 && threadIdx.x == 0) g_ts[i] = clock64();
#else
#define PTS(i)
#endif
static constexpr int kRld = 66;   // staging row stride: even (16-byte rows for cp.async / LDS.128), 66*8 B spreads the banks
static constexpr int kPanelSmem1 = (kPanelThreads * kRld + 2 * 128 + 80 + 80 + 64) * (int)sizeof(double);
static constexpr int kPanelSmem2 = kPanelSmem1 + kPanelThreads * kRld * (int)sizeof(double);   // + the previous half

// 1 / d without the IEEE fix-up of the compiler's division: MUFU seed (2^-23) and one third-order step,
// e = 1 - d x0, x1 = x0 (1 + e + e^2): 4 dependent operations instead of 7, result within ~1 ulp.
__device__ __forceinline__ double rcp_fast(double d) {
  double x0;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x0) : "d"(d));
  const double e = fma(-d, x0, 1.0);
  const double t = fma(e, e, e);
  return fma(x0, t, x0);
}

// One column of the lockstep elimination.  The serial chain of the whole factorisation is
//   r_c (reciprocal pivot) -> pivot d_{c+1} -> r_{c+1},
// and fp64 latency on this part is ~40 cycles per dependent operation, so the chain is kept as short as possible:
// the row that owns the next pivot forms it as d_{c+1} = a[c+1] - a[c]^2 r_c with a[c]^2 computed one step early
// (off the chain), i.e. ONE fma after r_c arrives, then the 4-operation reciprocal.  It is written without a branch
// (all diagonal rows compute it, one stores it) so that the compiler can interleave it with the bulk update.
template <int C>
struct ElimStep {
  static __device__ __forceinline__ void run(double (&a)[64], double& qq, double* col, double* inv, double* dv,
                                             int tid) {
    __syncthreads();
    const double* cc = col + (C & 1) * 128;               // cc[q] = entry (q, C) of the diagonal block
    const double r = inv[C];
    if constexpr (C + 1 < 64) {
      if (tid < 64) {
        const double d = fma(-qq, r, a[C + 1]);           // the pivot d_{C+1} if this is row C+1
        const double rn = rcp_fast(d);
        if (tid == C + 1) { inv[C + 1] = rn; dv[C + 1] = d; }
      }
    }
    const double m = a[C] * r;
    if constexpr (C + 1 < 64) {
      a[C + 1] = fma(-m, cc[C + 1], a[C + 1]);
      qq = a[C + 1] * a[C + 1];
      if (tid < 64 && tid >= C + 1) col[((C + 1) & 1) * 128 + tid] = a[C + 1];   // diagonal rows publish column C+1
    }
    constexpr int Q0 = C + 2;
    constexpr int QA = (Q0 & 1) ? Q0 + 1 : Q0;
    if constexpr ((Q0 & 1) && Q0 < 64) a[Q0] = fma(-m, cc[Q0], a[Q0]);
#pragma unroll
    for (int q = QA; q < 64; q += 2) {
      const double2 w = *reinterpret_cast<const double2*>(cc + q);
      a[q] = fma(-m, w.x, a[q]);
      a[q + 1] = fma(-m, w.y, a[q + 1]);
    }
    ElimStep<C + 1>::run(a, qq, col, inv, dv, tid);
  }
};
template <>
struct ElimStep<64> {
  static __device__ __forceinline__ void run(double (&)[64], double&, double*, double*, double*, int) {}
};

// Chain mode (droid_chol_solve_chains_f64): several column ranges ("chains") that are structurally decoupled from
// each other are factorised concurrently; they only meet in the separator rows [sep, n].  A launch that works for
// chain c must not touch the rows / columns of the later chains -- the hole [hole_lo, hole_hi) -- and, except for
// chain 0, accumulates its update of the separator x separator block into a private scratch copy (row stride lds)
// that is added to A once all chains are done.
struct ChainArgs {
  int hole_lo, hole_hi;   // rows / columns left alone (hole_lo >= hole_hi: none)
  int sep;                // first separator row
  double* scr;            // scratch for C[r][c], r >= sep and c >= sep (nullptr: update A in place)
  int lds;
  const int* lo64r;       // lo64 of this range (INT_MAX = the 64-row block has no entry in the range's columns), or nullptr
};

// do the rows [r_lo, r_hi] have an entry in the range's columns LEFT OF col_end (the range whose lo64 row is lo64r)?
// Rows whose envelope starts further right are still structurally zero in the panel being eliminated -- e.g. the rows
// of the separator a chain meets at its END, during all but the chain's last panels.  (Row n, the rhs: always.)
__device__ __forceinline__ bool rows_touched(const int* __restrict__ lo64r, int n, int r_lo, int r_hi, int col_end) {
  if (lo64r == nullptr || r_hi >= n) return true;
  for (int b = r_lo >> 6; b <= r_hi >> 6; b++)
    if (lo64r[b] < col_end) return true;
  return false;
}

static constexpr int kMaxChains = 8;                  // ranges factorised concurrently (one pair of streams each)
static constexpr int kMaxRanges = 2 * kMaxChains;     // chains + separators of all levels
struct Ranges {            // range c = [b[c], b[c+1]), c = 0..k, in elimination-tree layout: all ranges of level 0 (the
  int k;                   // chains), then level 1, ...; a range is coupled only to ranges of HIGHER levels
  int b[kMaxRanges + 2];
};

__global__ void __launch_bounds__(kPanelThreads, 1) chol_panel_kernel(double* __restrict__ A, int n, int j0, int jb,
                                                                      int prev, int* __restrict__ info,
                                                                      const int* __restrict__ hi64,
                                                                      int* __restrict__ arrive, int vec, int hole_lo,
                                                                      int hole_hi, double* __restrict__ linv,
                                                                      const int* __restrict__ lo64r) {
  constexpr int NT = kPanelThreads;
  constexpr int NW = NT / 32;                        // warps = rows staged per step
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* R = reinterpret_cast<double*>(smem_raw);  // [NT][kRld] row staging
  double* col = R + NT * kRld;                       // [2][128]  published column (double-buffered, padded)
  double* inv = col + 256;                           // [80]      1 / pivot
  double* dv = inv + 80;                             // [80]      pivots
  double* rs = dv + 80;                              // [64]      1 / sqrt(pivot)
  double* X = rs + 64;                               // [NT][kRld] the rows' entries in the previous 64 columns (prev only)
  const int tid = threadIdx.x;
  // the last CTA of a launch with linv eliminates the identity instead of matrix rows
  const bool ident = linv != nullptr && blockIdx.x == gridDim.x - 1;
  const int trow0 = j0 + jb + blockIdx.x * PRT;      // first trailing row of this CTA
  // envelope: rows past hi64[] have no entry in these columns (and get none); the CTA holding the rhs row stays
  if (!ident && hi64 != nullptr && blockIdx.x != 0 &&
      ((trow0 > hi64[(j0 + jb - 1) >> 6] && trow0 + PRT <= n) || (trow0 >= hole_lo && trow0 + PRT <= hole_hi) ||
       !rows_touched(lo64r, n, trow0, trow0 + PRT - 1, j0 + jb))) {
    if (threadIdx.x == 0) atomicAdd(&arrive[j0 >> 6], 1);
    return;
  }
  auto row_of = [&](int t) { return t < 64 ? j0 + t : trow0 + t - 64; };
  auto row_ok = [&](int t) { return t < 64 ? t < jb : (!ident && (trow0 + t - 64) <= n); };

  PTS(0)
  // stage: a warp-instruction moves 512 contiguous bytes of one row (16 B per lane) straight into shared memory
  const int sr = tid >> 5, sc = (tid & 31) * 2;      // row within a group of NW, column pair
  const bool vecp = vec && (jb & 1) == 0;            // 16-byte accesses also for a partial (even) sub-panel
  if (vecp) {
#pragma unroll 8
    for (int q = 0; q < NT / NW; q++) {
      const int t = q * NW + sr;
      const unsigned dst = (unsigned)__cvta_generic_to_shared(R + t * kRld + sc);
      if (row_ok(t) && sc >= jb) {      // columns behind a partial sub-panel: zero; the previous half is whole
        *reinterpret_cast<double2*>(R + t * kRld + sc) = make_double2(0.0, 0.0);
        if (prev)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(X + t * kRld + sc)), "l"(A + (size_t)row_of(t) * n + j0 + sc - prev));
      } else if (row_ok(t)) {
        const double* src = A + (size_t)row_of(t) * n + j0 + sc;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
        if (prev)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(X + t * kRld + sc)), "l"(src - prev));
      } else {
        *reinterpret_cast<double2*>(R + t * kRld + sc) = make_double2(0.0, 0.0);
        if (prev) *reinterpret_cast<double2*>(X + t * kRld + sc) = make_double2(0.0, 0.0);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  } else {
    for (int q = 0; q < NT / NW; q++) {
      const int t = q * NW + sr;
      double2 v = make_double2(0.0, 0.0);
      if (row_ok(t)) {
        const double* src = A + (size_t)row_of(t) * n + j0 + sc;
        if (sc < jb) v.x = src[0];
        if (sc + 1 < jb) v.y = src[1];
      }
      *reinterpret_cast<double2*>(R + t * kRld + sc) = v;
    }
  }
  if (tid < 128) { col[tid] = 0.0; col[128 + tid] = 0.0; }
  if (tid < 80) inv[tid] = 0.0;
  __syncthreads();
  // every CTA has now read the un-factorised diagonal block; CTA 0 overwrites it at the end, but only after all
  // CTAs of this launch have checked in here (a CTA may be scheduled late when the SMs are busy with the SYRK)
  if (tid == 0 && blockIdx.x != 0) atomicAdd(&arrive[j0 >> 6], 1);
  double a[64];
#pragma unroll
  for (int c = 0; c < 64; c += 2) {
    const double2 v = *reinterpret_cast<const double2*>(R + tid * kRld + c);
    a[c] = v.x; a[c + 1] = v.y;
  }
  if (prev) {
    // a[c] -= sum_k L[r][k] L[j0 + c][k] over the previous 64 columns: own row from X[tid], the multiplier rows are
    // the diagonal rows' entries X[0..63] (broadcast reads); 8 k at a time, 64 independent accumulators
    const double* xr = X + tid * kRld;
#pragma unroll 1
    for (int k0 = 0; k0 < 64; k0 += 8) {
      double x[8];
#pragma unroll
      for (int u = 0; u < 8; u += 2) {
        const double2 v = *reinterpret_cast<const double2*>(xr + k0 + u);
        x[u] = v.x; x[u + 1] = v.y;
      }
#pragma unroll
      for (int c = 0; c < 64; c++) {
        const double* d = X + c * kRld + k0;
        double acc = a[c];
#pragma unroll
        for (int u = 0; u < 8; u += 2) {
          const double2 w = *reinterpret_cast<const double2*>(d + u);
          acc = fma(-x[u], w.x, acc);
          acc = fma(-x[u + 1], w.y, acc);
        }
        a[c] = acc;
      }
    }
  }
  if (tid < 64) {   // diagonal block: lower triangle only, partial block padded with I
#pragma unroll
    for (int c = 0; c < 64; c++) {
      if (c > tid) a[c] = 0.0;
      if (tid >= jb && c == tid) a[c] = 1.0;
    }
    col[tid] = a[0];
    if (tid == 0) { inv[0] = rcp_fast(a[0]); dv[0] = a[0]; }
  } else if (ident) {
#pragma unroll
    for (int c = 0; c < 64; c++) a[c] = (c == tid - 64) ? 1.0 : 0.0;
  }
  double qq = a[0] * a[0];
  PTS(1)
  ElimStep<0>::run(a, qq, col, inv, dv, tid);
  __syncthreads();
  PTS(2)
  if (tid < 64) {
    const double d = dv[tid];
    if (!(d > 0.0) && blockIdx.x == 0) *info = 1;
    rs[tid] = 1.0 / sqrt(d);
  }
  __syncthreads();
  if (ident) {      // rows of I L^-T: row i = tid - 64 of the inverse transposed diagonal block (the jb real rows only:
                    // the rows behind a partial block belong to the next range)
    if (tid >= 64 && tid - 64 < jb) {
      double* dst = linv + (size_t)(j0 + tid - 64) * 64;
#pragma unroll
      for (int c = 0; c < 64; c += 2)
        *reinterpret_cast<double2*>(dst + c) = make_double2(a[c] * rs[c], a[c + 1] * rs[c + 1]);
    }
    return;
  }
#pragma unroll
  for (int c = 0; c < 64; c += 2)
    *reinterpret_cast<double2*>(R + tid * kRld + c) = make_double2(a[c] * rs[c], a[c + 1] * rs[c + 1]);
  PTS(3)
  if (blockIdx.x == 0 && tid == 0) {
    volatile int* cnt = arrive + (j0 >> 6);
    while (*cnt < (int)gridDim.x - 1) {}
  }
  __syncthreads();
  if (vecp) {
#pragma unroll 8
    for (int q = 0; q < NT / NW; q++) {
      const int t = q * NW + sr;
      if (!row_ok(t) || sc >= jb) continue;
      if (t < 64 && (blockIdx.x != 0 || sc > t)) continue;   // the diagonal block is written once, lower part
      const double2 v = *reinterpret_cast<const double2*>(R + t * kRld + sc);
      double* dst = A + (size_t)row_of(t) * n + j0 + sc;
      if (t < 64 && sc + 1 > t) dst[0] = v.x;
      else *reinterpret_cast<double2*>(dst) = v;
    }
  } else {
    for (int q = 0; q < NT / NW; q++) {
      const int t = q * NW + sr;
      if (!row_ok(t)) continue;
      double* dst = A + (size_t)row_of(t) * n + j0 + sc;
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const int c = sc + u;
        if (c >= jb) continue;
        if (t < 64 && (blockIdx.x != 0 || c > t)) continue;
        dst[u] = R[t * kRld + c];
      }
    }
  }
  PTS(4)
}

// ------------------------------------------------------------------------------------------------
// trailing update  C[bi][bj] -= X_bi X_bj^T over k in [J, J+kb), for all TM x TM tiles bi >= bj of
// rows/cols r0 = J+kb .. (row n included as a row, never as a column).  One CTA (256 threads) per tile,
// (TM/16)^2 fp64 accumulators per thread, operands staged k-major in shared memory in chunks of 16 k,
// double-buffered through registers so that global loads overlap the DFMA stream.
// ------------------------------------------------------------------------------------------------
template <int TM, bool VEC>
__global__ void __launch_bounds__(256, TM == 128 ? 1 : 2)
    chol_syrk_kernel(double* __restrict__ A, int n, int J, int kb, int strip, const int* __restrict__ hi64,
                     const ChainArgs ca, int tile_na, int tile_skip) {
  constexpr int F = TM / 64;
  constexpr int KC = 16;
  constexpr int LD = TM + 2;
  constexpr int LPT = TM / 32;  // double2 loads per thread per operand per chunk
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);  // [2 buffers][2 operands][KC][LD]
  const int t = blockIdx.x;
  int bi, bj;
  if (strip > 0) {   // only the first `strip` columns (ceil(strip/TM) block columns)
    const int nbj = (strip + TM - 1) / TM;
    bi = t / nbj; bj = t - bi * nbj;
    if (bj > bi) return;
  } else if (strip < 0) {   // everything except the first -strip block columns (look-ahead split)
    bi = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
    while ((bi + 1) * (bi + 2) / 2 <= t) bi++;
    while (bi * (bi + 1) / 2 > t) bi--;
    bj = t - bi * (bi + 1) / 2 - strip;
    bi -= strip;
  } else {
    bi = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
    while ((bi + 1) * (bi + 2) / 2 <= t) bi++;
    while (bi * (bi + 1) / 2 > t) bi--;
    bj = t - bi * (bi + 1) / 2;
  }
  // chain mode: the launch enumerates only the tile rows / columns that are not entirely inside the hole
  if (bi >= tile_na) bi += tile_skip;
  if (bj >= tile_na) bj += tile_skip;
  const int r0 = J + kb;
  const int i0 = r0 + bi * TM, c0 = r0 + bj * TM;
  const int rows = min(TM, n + 1 - i0);
  const int cols = min(min(TM, n - c0), strip > 0 ? strip - bj * TM : TM);
  if (cols <= 0 || rows <= 0) return;
  if (hi64 != nullptr) {   // envelope: X_bi or X_bj structurally zero in these kb columns
    const int hi = hi64[(J + kb - 1) >> 6];
    if (c0 > hi || (i0 > hi && i0 + rows <= n)) return;
  }
  if ((i0 >= ca.hole_lo && i0 + rows <= ca.hole_hi) || (c0 >= ca.hole_lo && c0 + cols <= ca.hole_hi)) return;
  // rows or columns (as rows of the factor) without an entry in this range's columns: nothing to subtract
  if (!rows_touched(ca.lo64r, n, i0, i0 + rows - 1, J + kb) || !rows_touched(ca.lo64r, n, c0, c0 + cols - 1, J + kb)) return;
  auto in_hole = [&](int g) { return g >= ca.hole_lo && g < ca.hole_hi; };
  // C element (gr, gc), gc even: in A, or in the chain's scratch copy of the separator block
  auto c_ptr = [&](int gr, int gc) -> double* {
    if (VEC && ca.scr != nullptr && gr >= ca.sep && gc >= ca.sep)
      return ca.scr + (size_t)(gr - ca.sep) * ca.lds + (gc - ca.sep);
    return A + (size_t)gr * n + gc;
  };
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  double acc[F * 4][F * 4];
#pragma unroll
  for (int a = 0; a < F * 4; a++)
#pragma unroll
    for (int b = 0; b < F * 4; b++) acc[a][b] = 0.0;
  double2 ra[LPT], rb[LPT];
  const int nch = (kb + KC - 1) / KC;
  auto gload = [&](int ch) {
#pragma unroll
    for (int q = 0; q < LPT; q++) {
      const int idx = tid + 256 * q;
      const int r = idx >> 3, k = ch * KC + 2 * (idx & 7);
      double2 va = make_double2(0.0, 0.0), vb = make_double2(0.0, 0.0);
      if (k < kb) {
        if (r < rows) {
          const double* p = A + (size_t)(i0 + r) * n + J + k;
          if (VEC) va = *reinterpret_cast<const double2*>(p);
          else { va.x = p[0]; if (k + 1 < kb) va.y = p[1]; }
        }
        if (r < cols) {
          const double* p = A + (size_t)(c0 + r) * n + J + k;
          if (VEC) vb = *reinterpret_cast<const double2*>(p);
          else { vb.x = p[0]; if (k + 1 < kb) vb.y = p[1]; }
        }
      }
      ra[q] = va; rb[q] = vb;
    }
  };
  auto sstore = [&](int buf) {
    double* As = sm + (size_t)buf * 2 * KC * LD;
    double* Bs = As + KC * LD;
#pragma unroll
    for (int q = 0; q < LPT; q++) {
      const int idx = tid + 256 * q;
      const int r = idx >> 3, k = 2 * (idx & 7);
      As[k * LD + r] = ra[q].x; As[(k + 1) * LD + r] = ra[q].y;
      Bs[k * LD + r] = rb[q].x; Bs[(k + 1) * LD + r] = rb[q].y;
    }
  };

  gload(0);
  // The C tile is fetched into shared memory with cp.async while the DFMA loop runs: the epilogue's
  // read-modify-write would otherwise expose 8 serial round trips to DRAM per tile (33 % of the stall
  // samples in profiles/r1_chol_ncu_summary.txt).
  double* Cs = sm + 2 * 2 * KC * LD;                 // [TM][TM]
  if (VEC) {
#pragma unroll 4
    for (int q = 0; q < TM * TM / 2 / 256; q++) {
      const int idx = tid + 256 * q;
      const int r = idx / (TM / 2), c = 2 * (idx % (TM / 2));
      if (r < rows && c < cols) {
        const unsigned dst = (unsigned)__cvta_generic_to_shared(Cs + r * TM + c);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(c_ptr(i0 + r, c0 + c)));
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }

  sstore(0);
  __syncthreads();
  for (int ch = 0; ch < nch; ch++) {
    if (ch + 1 < nch) gload(ch + 1);
    const double* As = sm + (size_t)(ch & 1) * 2 * KC * LD;
    const double* Bs = As + KC * LD;
#pragma unroll
    for (int k = 0; k < KC; k++) {
      double av[F * 4], bv[F * 4];
#pragma unroll
      for (int f = 0; f < F; f++) {
        // element index e = f*4 + h*2 + {0,1}  <->  row/col  f*64 + h*32 + 2*ty + {0,1}: the 16 lanes of a
        // half-warp read 16 consecutive 16-byte words (two conflict-free wavefronts per LDS.128)
        const double2 a01 = *reinterpret_cast<const double2*>(As + k * LD + f * 64 + ty * 2);
        const double2 a23 = *reinterpret_cast<const double2*>(As + k * LD + f * 64 + 32 + ty * 2);
        const double2 b01 = *reinterpret_cast<const double2*>(Bs + k * LD + f * 64 + tx * 2);
        const double2 b23 = *reinterpret_cast<const double2*>(Bs + k * LD + f * 64 + 32 + tx * 2);
        av[f * 4 + 0] = a01.x; av[f * 4 + 1] = a01.y; av[f * 4 + 2] = a23.x; av[f * 4 + 3] = a23.y;
        bv[f * 4 + 0] = b01.x; bv[f * 4 + 1] = b01.y; bv[f * 4 + 2] = b23.x; bv[f * 4 + 3] = b23.y;
      }
#pragma unroll
      for (int a = 0; a < F * 4; a++)
#pragma unroll
        for (int b = 0; b < F * 4; b++) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
    }
    if (ch + 1 < nch) sstore((ch + 1) & 1);
    __syncthreads();
  }
  if (VEC) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
  }
  // C -= acc.  The loads of one row are issued together (a load after a store to the same array could not
  // be hoisted by the compiler, which would serialise 64 round trips to L2).
#pragma unroll
  for (int a = 0; a < F * 4; a++) {
    const int r = (a >> 2) * 64 + ((a >> 1) & 1) * 32 + ty * 2 + (a & 1);
    if (r >= rows || in_hole(i0 + r)) continue;
    double* dst = A + (size_t)(i0 + r) * n + c0;
    const int cmax = min(cols - 1, i0 + r - c0);  // lower triangle only
    double cv[F * 4];
    if (VEC) {
#pragma unroll
      for (int b = 0; b < F * 4; b += 2) {
        const int c = (b >> 2) * 64 + ((b >> 1) & 1) * 32 + tx * 2;
        if (c <= cmax) {
          const double2 v = *reinterpret_cast<const double2*>(Cs + r * TM + c);
          cv[b] = v.x; cv[b + 1] = v.y;
        }
      }
#pragma unroll
      for (int b = 0; b < F * 4; b += 2) {
        const int c = (b >> 2) * 64 + ((b >> 1) & 1) * 32 + tx * 2;
        if (c > cmax || in_hole(c0 + c)) continue;
        // Ranges of one level run concurrently and may enumerate the same tile of a higher-level block; structurally
        // only one of them has a non-zero update for a given element (the other's operand rows are exactly zero), so
        // an element whose update is exactly 0 is left alone instead of being written back unchanged (a lost update
        // for the range that really changes it).
        if (acc[a][b] == 0.0 && acc[a][b + 1] == 0.0) continue;
        double* d2 = c_ptr(i0 + r, c0 + c);
        if (c + 1 <= cmax) {
          *reinterpret_cast<double2*>(d2) = make_double2(cv[b] - acc[a][b], cv[b + 1] - acc[a][b + 1]);
        } else {
          d2[0] = cv[b] - acc[a][b];
        }
      }
    } else {
#pragma unroll
      for (int b = 0; b < F * 4; b++) {
        const int c = (b >> 2) * 64 + ((b >> 1) & 1) * 32 + tx * 2 + (b & 1);
        if (c <= cmax) cv[b] = dst[c];
      }
#pragma unroll
      for (int b = 0; b < F * 4; b++) {
        const int c = (b >> 2) * 64 + ((b >> 1) & 1) * 32 + tx * 2 + (b & 1);
        if (c <= cmax && acc[a][b] != 0.0) dst[c] = cv[b] - acc[a][b];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// backward sweep, one launch per 64-block from the last to the first:
//   x_j = L_jj^-T y_j  (every CTA redundantly, warp 0), then  y[i] -= sum_r L[j0+r][i] x_j[r], i < j0.
// y lives in row n of A and is overwritten by x.  CTA = 64 columns i x 4 row groups of 16.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chol_bwd_kernel(double* __restrict__ A, int n, int j0, int jb) {
  __shared__ double D[NB][NB + 1];
  __shared__ double x[NB];
  __shared__ double rinv[NB];
  __shared__ double part[4][NB];
  const int tid = threadIdx.x;
  double* y = A + (size_t)n * n;
  for (int e = tid; e < NB * NB; e += 256) {
    const int r = e / NB, c = e - r * NB;
    D[r][c] = (c <= r && r < jb) ? A[(size_t)(j0 + r) * n + j0 + c] : 0.0;
  }
  if (tid < NB) x[tid] = tid < jb ? y[j0 + tid] : 0.0;
  __syncthreads();
  if (tid < NB) rinv[tid] = tid < jb ? 1.0 / D[tid][tid] : 0.0;
  __syncthreads();
  if (tid < 32) {
    // lane holds x[lane] and x[lane+32]; right-looking back substitution with shuffles
    double x0 = x[tid], x1 = x[tid + 32];
#pragma unroll
    for (int c = NB - 1; c >= 0; c--) {
      const double src = (c >= 32) ? x1 : x0;
      const double xc = __shfl_sync(0xffffffffu, src, c & 31) * rinv[c];
      if (c >= 32) {
        if (tid == (c & 31)) x1 = xc;
        else if (tid < (c & 31)) x1 = fma(-D[c][tid + 32], xc, x1);
        x0 = fma(-D[c][tid], xc, x0);
      } else {
        if (tid == c) x0 = xc;
        else if (tid < c) x0 = fma(-D[c][tid], xc, x0);
      }
    }
    x[tid] = x0; x[tid + 32] = x1;
  }
  __syncthreads();
  const int i = blockIdx.x * NB + (tid & 63), g = tid >> 6;
  double s = 0.0;
  if (i < j0) {
#pragma unroll
    for (int r = 0; r < 16; r++) {
      const int rr = g * 16 + r;
      if (rr < jb) s = fma(A[(size_t)(j0 + rr) * n + i], x[rr], s);
    }
  }
  part[g][tid & 63] = s;
  __syncthreads();
  if (tid < NB) {
    const int ii = blockIdx.x * NB + tid;
    if (ii < j0) y[ii] -= part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
  }
  if (blockIdx.x == 0 && tid < jb) y[j0 + tid] = x[tid];
}

// ------------------------------------------------------------------------------------------------
// backward sweep, two-level version (replaces 94 dependent launches by 2 per 512 rows):
//   chol_bwd_diag_kernel : ONE CTA solves L_SS^T x_S = y_S for a super-block S of up to 512 rows, walking its
//                          64-blocks from the last to the first with y_S in shared memory;
//   chol_bwd_gemv_kernel : y[0:J0] -= L[S, 0:J0]^T x_S on all SMs.
// ------------------------------------------------------------------------------------------------
static constexpr int SB = 192;                                   // super-block rows (3 blocks of 64)
static constexpr int kBwdBlocks = (SB / NB) * (SB / NB + 1) / 2;  // lower-triangular 64x64 blocks: 6
static constexpr int kBwdSmem = (kBwdBlocks * NB * NB + SB + NB + NB + 4 * NB) * (int)sizeof(double);

__global__ void __launch_bounds__(256) chol_bwd_diag_kernel(double* __restrict__ A, int n, int J0, int JB, int vec) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* Lb = reinterpret_cast<double*>(smem_raw);     // [6][64][64]: block (bi,bj) at bi(bi+1)/2 + bj
  double* ys = Lb + kBwdBlocks * NB * NB;               // [SB]
  double* x = ys + SB;                                  // [64]
  double* rinv = x + NB;                                // [64]
  double* part = rinv + NB;                             // [4][64]
  const int tid = threadIdx.x;
  double* y = A + (size_t)n * n;
  // The whole lower triangle of the super-block (<= 192 KB) comes in with cp.async, every request in flight at
  // once: the serial walk below then never waits on DRAM (the per-64-block version paid a round trip per block).
  for (int e = tid; e < kBwdBlocks * NB * (NB / 2); e += 256) {
    const int blk = e / (NB * NB / 2), w = e % (NB * NB / 2);
    const int r = w / (NB / 2), c = 2 * (w % (NB / 2));
    const int bi = blk >= 3 ? 2 : (blk >= 1 ? 1 : 0), bj = blk - bi * (bi + 1) / 2;
    const int gr = bi * NB + r, gc = bj * NB + c;
    double* dst = Lb + (size_t)blk * NB * NB + r * NB + c;
    if (gr < JB && gc < JB) {
      const double* src = A + (size_t)(J0 + gr) * n + J0 + gc;
      if (vec) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src));
      } else {
        dst[0] = src[0];
        dst[1] = (gc + 1 < JB) ? src[1] : 0.0;
      }
    } else {
      dst[0] = 0.0; dst[1] = 0.0;
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int i = tid; i < SB; i += 256) ys[i] = i < JB ? y[J0 + i] : 0.0;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  const int nsb = (JB + NB - 1) / NB;
  for (int sb = nsb - 1; sb >= 0; sb--) {
    const int o = sb * NB;
    const int jb = min(NB, JB - o);
    const double* D = Lb + (size_t)(sb * (sb + 1) / 2 + sb) * NB * NB;
    __syncthreads();
    if (tid < NB) rinv[tid] = tid < jb ? 1.0 / D[tid * NB + tid] : 0.0;
    __syncthreads();
    if (tid < 32) {
      double x0 = ys[o + tid], x1 = ys[o + tid + 32];
#pragma unroll
      for (int c = NB - 1; c >= 0; c--) {
        const double src = (c >= 32) ? x1 : x0;
        const double xc = __shfl_sync(0xffffffffu, src, c & 31) * rinv[c];
        if (c >= 32) {
          if (tid == (c & 31)) x1 = xc;
          else if (tid < (c & 31)) x1 = fma(-D[c * NB + tid + 32], xc, x1);
          x0 = fma(-D[c * NB + tid], xc, x0);
        } else {
          if (tid == c) x0 = xc;
          else if (tid < c) x0 = fma(-D[c * NB + tid], xc, x0);
        }
      }
      x[tid] = x0; x[tid + 32] = x1;
      ys[o + tid] = x0; ys[o + tid + 32] = x1;
    }
    __syncthreads();
    const int g = tid >> 6, cl = tid & 63;
    for (int t = 0; t < sb; t++) {              // y_t -= L(sb,t)^T x_sb
      const double* Ls = Lb + (size_t)(sb * (sb + 1) / 2 + t) * NB * NB;
      double sum = 0.0;
#pragma unroll
      for (int r = 0; r < 16; r++) sum = fma(Ls[(g * 16 + r) * NB + cl], x[g * 16 + r], sum);
      part[g * NB + cl] = sum;
      __syncthreads();
      if (tid < NB) ys[t * NB + tid] -= part[tid] + part[NB + tid] + part[2 * NB + tid] + part[3 * NB + tid];
      __syncthreads();
    }
  }
  __syncthreads();
  for (int i = tid; i < JB; i += 256) y[J0 + i] = ys[i];
}

__global__ void __launch_bounds__(256) chol_bwd_gemv_kernel(double* __restrict__ A, int n, int J0, int JB,
                                                            const int* __restrict__ lo64) {
  __shared__ double xs[SB];
  __shared__ double part[4][NB];
  const int tid = threadIdx.x;
  if (lo64 != nullptr) {   // envelope: the super-block's rows start at column lo
    int lo = n;
    for (int b = J0 >> 6; b <= (J0 + JB - 1) >> 6; b++) lo = min(lo, lo64[b]);
    if ((int)blockIdx.x * NB + NB - 1 < lo) return;
  }
  double* y = A + (size_t)n * n;
  for (int i = tid; i < SB; i += 256) xs[i] = i < JB ? y[J0 + i] : 0.0;
  __syncthreads();
  const int i = blockIdx.x * NB + (tid & 63), g = tid >> 6;
  double s0 = 0.0, s1 = 0.0;
  if (i < J0) {
    // row group g handles rows g, g+4, g+8, ... of the super-block (two independent chains)
    int r = g;
#pragma unroll 4
    for (; r + 4 < JB; r += 8) {
      s0 = fma(A[(size_t)(J0 + r) * n + i], xs[r], s0);
      s1 = fma(A[(size_t)(J0 + r + 4) * n + i], xs[r + 4], s1);
    }
    if (r < JB) s0 = fma(A[(size_t)(J0 + r) * n + i], xs[r], s0);
  }
  part[g][tid & 63] = s0 + s1;
  __syncthreads();
  if (tid < NB) {
    const int ii = blockIdx.x * NB + tid;
    if (ii < J0) y[ii] -= part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
  }
}

// ------------------------------------------------------------------------------------------------
// backward sweep in ONE persistent CTA (n <= kBwdPersistMax): the sweep is a chain of n dependent steps with very
// little work per step, so 2 launches per 192 rows (and their launch gaps) cost more than the arithmetic.
// y lives in shared memory for the whole sweep.  Per 64-row block, from the last to the first:
//   * its diagonal block arrives by cp.async one block ahead (double buffer);
//   * columns are pre-scaled by 1 / L[c][c] so that the serial chain of the in-warp back-substitution is one shuffle
//     + one DFMA per column (no multiply by the reciprocal diagonal on the chain);
//   * all 512 threads subtract the block's contribution from the y entries to its left, reading L straight from
//     global memory (row-major L => consecutive threads read consecutive columns; 64 independent loads per thread).
//     In envelope mode only the columns from the block's first non-zero column on are touched.
// ------------------------------------------------------------------------------------------------
static constexpr int kBwdPersistMax = kCholChainsMaxRows;
static constexpr int kBwdPThreads = 512;
static constexpr int kBwdLd = 66;     // row stride of a staged 64x64 block (16-byte rows, spreads the banks)
static size_t bwd_persist_smem(int n) { return ((size_t)n + 64 + 2 * 64 * kBwdLd + 64 + 64) * sizeof(double); }

// One CTA sweeps one range (rows [b[c], b[c+1]), blocks of 64 rows counted from the range start).  The launches go
// down the elimination tree: the ranges of the top level first, the chains (level 0) last; the CTAs of one launch are
// the ranges of one level (c0 = first range of the level) and run concurrently.  A range keeps its OWN part of y in
// shared memory; the contributions of its rows to the y entries of the (lower-level) ranges they reach into go
// straight to global memory -- structurally every such entry is touched by one range of the level only, and a
// contribution that is exactly zero is not written, so that a block of rows shared by two ranges cannot lose an update.
// linv: rows of the inverse transposed diagonal blocks (written by the panel kernels' identity CTA): x_blk = L^-T y_blk
// is then a 64x64 matrix-vector product (8 threads per row + a 3-step shuffle reduction) instead of a 64-step chain of
// dependent shuffle + DFMA; without it the block is solved by in-warp back-substitution on the diagonal block itself.
__global__ void __launch_bounds__(kBwdPThreads, 1) chol_bwd_persist_kernel(double* __restrict__ A, int n,
                                                                           const int* __restrict__ lo64,
                                                                           const Ranges rg, int c0,
                                                                           const double* __restrict__ linv) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* Dsm = reinterpret_cast<double*>(smem_raw);      // [2][64][kBwdLd]
  double* xs = Dsm + 2 * 64 * kBwdLd;                      // [64]
  double* rinv = xs + 64;                                  // [64]
  double* ys = rinv + 64;                                  // [n (+pad)]
  const int tid = threadIdx.x;
  double* y = A + (size_t)n * n;
  const int nb64 = (n + 63) >> 6;
  const int rc = c0 + blockIdx.x;                          // the range this CTA sweeps
  const int rbeg = rg.b[rc], rend = rg.b[rc + 1];
  if (rend <= rbeg) return;
  const int sg0 = rc;                                      // own columns only; the lower ranges: chol_bwd_wide_kernel
  const int ylo = rbeg;                                    // ys[] holds the own range only
  const int nblk = (rend - rbeg + 63) >> 6;
  auto load_diag = [&](int blk, int buf) {
    const int o = rbeg + (blk << 6), jb = min(64, rend - o);
    double* D = Dsm + buf * 64 * kBwdLd;
    for (int e = tid; e < 64 * 32; e += kBwdPThreads) {
      const int r = e >> 5, c = 2 * (e & 31);
      double* dst = D + r * kBwdLd + c;
      if (linv != nullptr) {
        if (r < jb) {
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(linv + (size_t)(o + r) * 64 + c));
        } else {
          dst[0] = 0.0; dst[1] = 0.0;
        }
      } else if (r < jb && c <= r) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(A + (size_t)(o + r) * n + o + c));
      } else {
        dst[0] = 0.0; dst[1] = 0.0;
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  load_diag(nblk - 1, (nblk - 1) & 1);
  for (int i = ylo + tid; i < rend; i += kBwdPThreads) ys[i] = y[i];
  for (int blk = nblk - 1; blk >= 0; blk--) {
    const int o = rbeg + (blk << 6), jb = min(64, rend - o);
    double* D = Dsm + (blk & 1) * 64 * kBwdLd;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                       // diagonal block landed; previous update of ys complete
    if (blk > 0) load_diag(blk - 1, (blk - 1) & 1);
    if (linv != nullptr) {
      // x[i] = sum_{c >= i} (L^-T)[i][c] y[c]
      const int i = tid >> 3, part = tid & 7;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int c = part; c < 64; c += 16) {
        s0 = fma(D[i * kBwdLd + c], (c < jb) ? ys[o + c] : 0.0, s0);
        s1 = fma(D[i * kBwdLd + c + 8], (c + 8 < jb) ? ys[o + c + 8] : 0.0, s1);
      }
      double s = s0 + s1;
      s += shfl_xor_f64(s, 1);
      s += shfl_xor_f64(s, 2);
      s += shfl_xor_f64(s, 4);
      __syncthreads();                                     // every thread has read ys[o ..]
      if (part == 0) {
        xs[i] = (i < jb) ? s : 0.0;
        if (i < jb) ys[o + i] = s;
      }
      __syncthreads();
    } else {
      if (tid < 64) rinv[tid] = tid < jb ? 1.0 / D[tid * kBwdLd + tid] : 0.0;
      __syncthreads();
      // scale column c of the strictly lower part and y_c by 1 / L[c][c]
      for (int e = tid; e < 64 * 64; e += kBwdPThreads) {
        const int r = e >> 6, c = e & 63;
        if (c < r) D[r * kBwdLd + c] *= rinv[c];
      }
      if (tid < 64) xs[tid] = tid < jb ? ys[o + tid] * rinv[tid] : 0.0;
      __syncthreads();
      if (tid < 32) {
        double x0 = xs[tid], x1 = xs[tid + 32];
#pragma unroll
        for (int c = 63; c >= 0; c--) {
          const double xc = shfl_f64(c >= 32 ? x1 : x0, c & 31);
          if (c >= 32) {
            if (tid < (c & 31)) x1 = fma(-D[c * kBwdLd + tid + 32], xc, x1);
            x0 = fma(-D[c * kBwdLd + tid], xc, x0);
          } else if (tid < c) {
            x0 = fma(-D[c * kBwdLd + tid], xc, x0);
          }
        }
        xs[tid] = x0; xs[tid + 32] = x1;
        if (tid < jb) ys[o + tid] = x0;
        if (tid + 32 < jb) ys[o + tid + 32] = x1;
      }
      __syncthreads();
    }
    // y[c] -= sum_r L[o + r][c] x[r] for the columns left of the block, range by range
    const int bl0 = o >> 6, bl1 = (o + jb - 1) >> 6;      // the (absolute) 64-blocks of lo64 these rows fall into
    for (int sg = sg0; sg <= rc; sg++) {
      const int hi = min(rg.b[sg + 1], o);
      const int l = min(lo64[sg * nb64 + bl0], lo64[sg * nb64 + bl1]);
      const int lo = max(rg.b[sg], min(l, hi)) & ~1;
      for (int c = lo + tid; c < hi; c += kBwdPThreads) {
        const double* col = A + (size_t)o * n + c;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int r = 0;
#pragma unroll 4
        for (; r + 3 < jb; r += 4) {
          s0 = fma(col[(size_t)r * n], xs[r], s0);
          s1 = fma(col[(size_t)(r + 1) * n], xs[r + 1], s1);
          s2 = fma(col[(size_t)(r + 2) * n], xs[r + 2], s2);
          s3 = fma(col[(size_t)(r + 3) * n], xs[r + 3], s3);
        }
        for (; r < jb; r++) s0 = fma(col[(size_t)r * n], xs[r], s0);
        const double sum = (s0 + s1) + (s2 + s3);
        if (sg == rc) ys[c] -= sum;
        else if (sum != 0.0) y[c] -= sum;
      }
    }
  }
  __syncthreads();
  for (int i = ylo + tid; i < rend; i += kBwdPThreads) y[i] = ys[i];
}

// Second half of the backward sweep of a level above the chains: once the level's own x is known (kernel above), its
// rows' contributions to the y entries of all LOWER ranges, y[c] -= sum_r L[r][c] x[r], are one wide matrix-vector
// product with no dependency chain -- spread over the whole GPU instead of being walked by the one persistent CTA of the
// range (a top-level separator row reaches into two whole chains and four lower separators: ~2000 columns).
// grid = (64-column tiles of [0, first row of the level), ranges of the level); thread = one column, 4 row groups.
__global__ void __launch_bounds__(256) chol_bwd_wide_kernel(double* __restrict__ A, int n, const int* __restrict__ lo64,
                                                            const Ranges rg, int c0) {
  __shared__ double part[4][64];
  const int rc = c0 + blockIdx.y;
  const int rbeg = rg.b[rc], rend = rg.b[rc + 1];
  const int c = blockIdx.x * 64 + (threadIdx.x & 63), g = threadIdx.x >> 6;
  const int nb64 = (n + 63) >> 6;
  const double* __restrict__ y = A + (size_t)n * n;
  // the range the tile's columns lie in (tiles may straddle two ranges: resolved per column)
  int sg = 0;
  while (sg < rc && c >= rg.b[sg + 1]) sg++;
  double s0 = 0.0, s1 = 0.0;
  if (c < rg.b[c0] && sg < rc) {
    const int* lo = lo64 + (size_t)sg * nb64;
    for (int r0 = rbeg + 16 * g; r0 < rend; r0 += 64) {          // row group g takes 16 rows of every 64
      const int r1 = min(r0 + 16, rend);
      // rows whose 64-block has nothing at or left of this column in range sg are structurally zero here
      if (min(lo[r0 >> 6], lo[(r1 - 1) >> 6]) > c) continue;
      int r = r0;
#pragma unroll 4
      for (; r + 1 < r1; r += 2) {
        s0 = fma(A[(size_t)r * n + c], y[r], s0);
        s1 = fma(A[(size_t)(r + 1) * n + c], y[r + 1], s1);
      }
      if (r < r1) s0 = fma(A[(size_t)r * n + c], y[r], s0);
    }
  }
  part[g][threadIdx.x & 63] = s0 + s1;
  __syncthreads();
  if (threadIdx.x < 64 && c < rg.b[c0] && sg < rc) {
    const double sum = (part[0][threadIdx.x] + part[1][threadIdx.x]) + (part[2][threadIdx.x] + part[3][threadIdx.x]);
    if (sum != 0.0) A[(size_t)n * n + c] -= sum;      // structurally one range of the level per column: no race
  }
}

// Envelope summary per 64-block and range.  firstc[c * n + r] = first structurally non-zero column of row r inside
// range c (fill-in of a Cholesky factor stays inside these per-range row envelopes); a value
// outside [b[c], min(r, b[c+1] - 1)] means "row r has nothing in range c", except in the row's own range, where it is
// clamped.
//   hi64[c * nb + blk] = last row with an entry in a column of range c that is <= 64 blk + 63  (monotone inside c)
//   lo64[c * nb + blk] = first column of range c used by the rows of block blk  (INT_MAX: none)
__global__ void __launch_bounds__(1024) chol_env_kernel(const int* __restrict__ firstc, int n, const Ranges rg,
                                                        int* __restrict__ hi64, int* __restrict__ lo64) {
  // one CTA per range (the tree of a long trajectory has 15 of them)
  const int nb = (n + 63) >> 6;
  const int c = blockIdx.x;
  int* __restrict__ hi = hi64 + (size_t)c * nb;
  int* __restrict__ lo = lo64 + (size_t)c * nb;
  for (int blk = threadIdx.x; blk < nb; blk += 1024) {
    hi[blk] = min(64 * blk + 63, rg.b[c + 1] - 1);
    lo[blk] = 0x7fffffff;
  }
  __syncthreads();
  for (int r = threadIdx.x; r < n; r += 1024) {
    int f = firstc[(size_t)c * n + r];
    const int top = min(r, rg.b[c + 1] - 1);
    if (r >= rg.b[c] && r < rg.b[c + 1]) f = max(rg.b[c], min(f, r));   // own range: at least the diagonal
    else if (f < rg.b[c] || f > top) continue;
    atomicMax(&hi[f >> 6], r);
    atomicMin(&lo[r >> 6], f);
  }
  __syncthreads();
  if (threadIdx.x == 0 && rg.b[c + 1] > rg.b[c]) {
    int run = 0;
    for (int blk = rg.b[c] >> 6; blk <= (rg.b[c + 1] - 1) >> 6; blk++) {
      run = max(run, hi[blk]);
      hi[blk] = run;
    }
  }
}

// A[r][c] += sum of the chains' scratch copies, r in [sep, n], c in [sep, min(r, n - 1)]
__global__ void chol_merge_kernel(double* __restrict__ A, int n, int sep, const double* __restrict__ scr, int nscr,
                                  size_t scr_stride, int lds) {
  const int r = sep + blockIdx.x;
  const int cmax = min(r, n - 1);
  for (int c = sep + threadIdx.x; c <= cmax; c += blockDim.x) {
    double acc = 0.0;
    for (int q = 0; q < nscr; q++) acc += scr[q * scr_stride + (size_t)(r - sep) * lds + (c - sep)];
    A[(size_t)r * n + c] += acc;
  }
}

}  // namespace droid

using namespace droid;

static size_t chol_ws_ints(int n, int k) {
  const size_t nb = (size_t)((n > 0 ? n : 0) + 63) / 64;
  return (size_t)(k + 1) * (3 * nb + 1);   // panel check-in counters | hi64 | lo64, per range
}

static size_t chol_ws_head(int n, int nchains) {
  return (256 + chol_ws_ints(n, nchains > 0 ? nchains : 0) * sizeof(int) + 255) / 256 * 256;
}
// rows of the inverse transposed 64x64 diagonal blocks, one row of 64 doubles per matrix row
static size_t chol_ws_linv(int n) { return ((size_t)(n > 0 ? n : 0) * 64 * sizeof(double) + 255) / 256 * 256; }

// scratch of the levels whose ranges run concurrently: one (S + 1) x S block per level, S = rows behind the level;
// the blocks shrink from level to level, so twice the first one bounds their sum (checked at run time)
extern "C" size_t droid_chol_tree_workspace_bytes(int n, int nranges, int sep_rows) {
  const size_t scr = nranges > 2 ? 2 * (size_t)(sep_rows + 1) * (size_t)sep_rows * sizeof(double) : 0;
  return chol_ws_head(n, nranges > 0 ? nranges - 1 : 0) + chol_ws_linv(n) + scr;
}

// chains + one flat separator: any chain may touch any separator row, so every chain but the first gets a PRIVATE
// scratch copy of the separator block
extern "C" size_t droid_chol_chains_workspace_bytes(int n, int nchains, int sep_rows) {
  const size_t scr = nchains > 1 ? (size_t)(nchains - 1) * (size_t)(sep_rows + 1) * (size_t)sep_rows * sizeof(double) : 0;
  return chol_ws_head(n, nchains > 0 ? nchains : 0) + chol_ws_linv(n) + scr;
}

extern "C" size_t droid_chol_workspace_bytes(int n) { return droid_chol_chains_workspace_bytes(n, 0, 0); }

extern "C" int droid_chol_solve_f64(double* A, double* b, int n, int* info_dev, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  return droid_chol_solve_chains_f64(A, b, n, nullptr, 0, nullptr, info_dev, workspace, workspace_bytes, stream);
}

extern "C" int droid_chol_solve_env_f64(double* A, double* b, int n, const int* first_dev, int* info_dev,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  return droid_chol_solve_chains_f64(A, b, n, first_dev, 0, nullptr, info_dev, workspace, workspace_bytes, stream);
}

// frontend-sized systems of ba: damping + factorisation + both sweeps + info in ONE launch (no memset, no damping kernel)
int droid::launch_chol_small_damped(double* A, int n, float lm, float ep, int* info_dev, cudaStream_t st) {
  if (n <= 0 || n > kSmallMax) return DROID_ERR_INVALID_ARGUMENT;
  if (dev_ctx().first(kOnceCholSmall))
    cudaFuncSetAttribute(chol_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSmemMax);
  DROID_LAUNCH chol_small_kernel<<<1, ((n + 1 + 31) / 32) * 32, chol_small_smem(n), st>>>(A, n, info_dev, 1, lm, ep);
  return check_launch();
}

static constexpr int kSyrk128 = (2 * 2 * 16 * (128 + 2) + 128 * 128) * (int)sizeof(double);
static constexpr int kSyrk64 = (2 * 2 * 16 * (64 + 2) + 64 * 64) * (int)sizeof(double);
// once per device: shared-memory opt-ins, the pairs of streams and the events of the concurrent ranges
static void chol_setup(DevCtx& dc) {
  if (dc.first(kOnceChol)) {
    cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPanelSmem2);
    cudaFuncSetAttribute(chol_bwd_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem);
    cudaFuncSetAttribute(chol_syrk_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrk128);
    cudaFuncSetAttribute(chol_syrk_kernel<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrk128);
    cudaFuncSetAttribute(chol_syrk_kernel<64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrk64);
    cudaFuncSetAttribute(chol_syrk_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrk64);
    cudaFuncSetAttribute(chol_bwd_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)bwd_persist_smem(kBwdPersistMax));
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    for (int c = 0; c < kMaxChains; c++) {
      ChainStreams& x = dc.chain[c];
      cudaStreamCreateWithPriority(&x.side, cudaStreamNonBlocking, hi);
      cudaStreamCreateWithFlags(&x.main, cudaStreamNonBlocking);
      cudaEventCreateWithFlags(&x.fork, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&x.p, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&x.r, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&x.join, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&x.done, cudaEventDisableTiming);
    }
    cudaEventCreateWithFlags(&dc.chol_start, cudaEventDisableTiming);
  }
}

namespace {
using ChainCtx = droid::ChainStreams;
struct ChainState {
  int J, Jend;
  bool on_side, finished;
  bool r_recorded;          // cx->r was recorded by THIS range (an event of an earlier call must not be waited on in a capture)
  cudaStream_t main;
  ChainCtx* cx;
  ChainArgs ca;
  const int* hi64;
  int* arrive;
  bool prof;
};
}  // namespace

static int chol_solve_tree_impl(double* A, double* b, int n, const int* firstc_dev, int nranges,
                                const int* range_end, int nlevels, const int* level_first, int* info_dev,
                                void* workspace, size_t workspace_bytes, void* stream, bool private_scr);
static int chol_solve_tree_enqueue(double* A, double* b, int n, const int* firstc_dev, int nranges,
                                   const int* range_end, int nlevels, const int* level_first, int* info_dev,
                                   void* workspace, size_t workspace_bytes, void* stream, bool private_scr,
                                   bool capturing);

extern "C" int droid_chol_solve_chains_f64(double* A, double* b, int n, const int* firstc_dev, int nchains,
                                           const int* chain_end, int* info_dev, void* workspace,
                                           size_t workspace_bytes, void* stream) {
  if (nchains < 0 || nchains > kMaxChains || (nchains > 0 && (!chain_end || !firstc_dev))) return DROID_ERR_INVALID_ARGUMENT;
  // chains + ONE separator = a two-level tree (a single range when there are no chains)
  int range_end[kMaxChains + 1], level_first[3] = {0, nchains, nchains + 1};
  for (int c = 0; c < nchains; c++) range_end[c] = chain_end[c];
  range_end[nchains] = n;
  if (nchains == 0) { level_first[1] = 1; }
  return chol_solve_tree_impl(A, b, n, firstc_dev, nchains + 1, range_end, nchains == 0 ? 1 : 2, level_first, info_dev,
                              workspace, workspace_bytes, stream, true);
}

extern "C" int droid_chol_solve_tree_f64(double* A, double* b, int n, const int* firstc_dev, int nranges,
                                         const int* range_end, int nlevels, const int* level_first, int* info_dev,
                                         void* workspace, size_t workspace_bytes, void* stream) {
  return chol_solve_tree_impl(A, b, n, firstc_dev, nranges, range_end, nlevels, level_first, info_dev, workspace,
                              workspace_bytes, stream, false);
}

// private_scr: every range of a level but the first accumulates its updates of the rows behind the level in a scratch
// copy of its own (general coupling); otherwise ranges at odd positions share ONE copy per level (path structure).
// ------------------------------------------------------------------------------------------------
// The tree solve as a CUDA GRAPH.  A solve over 8 chains is ~400 small launches on 16 streams with ~250 event
// operations between them; one host thread needs ~13 us per chain step for that, which with 8 chains in flight is more
// than a step takes on the device (the chains were launch-bound: 117 us per 128 columns instead of 70).  The launch
// sequence depends only on (n, tree layout) and the pointers, so it is captured once -- on an internal stream, the
// caller's may be the legacy default stream, which cannot be captured -- and replayed with one cudaGraphLaunch into
// the caller's stream.  A few graphs are kept per device (the backend alternates between at most a couple of layouts).
// DROID_CHOL_GRAPH=0 or DROID_CHOL_TRACE=1: direct launches.
// ------------------------------------------------------------------------------------------------
namespace {
struct GraphKey {
  const void *A, *firstc, *info, *ws;
  size_t ws_bytes;
  int n, nranges, nlevels, priv;
  int range_end[kMaxRanges], level_first[kMaxRanges + 1];
};
struct GraphEntry {
  GraphKey key;
  cudaGraphExec_t exec = nullptr;
  unsigned long long launches = 0, stamp = 0;
  bool unusable = false;            // capture failed for this key once: do not try again
};
constexpr int kGraphSlots = 4;
GraphEntry g_graphs[64][kGraphSlots];
unsigned long long g_graph_clock = 0;
}  // namespace

static int chol_solve_tree_impl(double* A, double* b, int n, const int* firstc_dev, int nranges,
                                const int* range_end, int nlevels, const int* level_first, int* info_dev,
                                void* workspace, size_t workspace_bytes, void* stream, bool private_scr) {
  static const bool graph_on = []() { const char* e = getenv("DROID_CHOL_GRAPH"); return !(e && e[0] == '0'); }();
  static const bool trace_env = []() { const char* e = getenv("DROID_CHOL_TRACE"); return e && e[0] == '1'; }();
  if (!graph_on || trace_env || nranges < 3 || nranges > kMaxRanges || nlevels < 1 || nlevels > nranges || !range_end ||
      !level_first || n <= kSmallMax)
    return chol_solve_tree_enqueue(A, b, n, firstc_dev, nranges, range_end, nlevels, level_first, info_dev, workspace,
                                   workspace_bytes, stream, private_scr, false);
  DevCtx& dc = dev_ctx();
  std::lock_guard<std::recursive_mutex> call_lock(dc.call);
  GraphKey key;
  memset(&key, 0, sizeof(key));
  key.A = A; key.firstc = firstc_dev; key.info = info_dev; key.ws = workspace; key.ws_bytes = workspace_bytes;
  key.n = n; key.nranges = nranges; key.nlevels = nlevels; key.priv = private_scr ? 1 : 0;
  for (int c = 0; c < nranges; c++) key.range_end[c] = range_end[c];
  for (int l = 0; l <= nlevels; l++) key.level_first[l] = level_first[l];
  GraphEntry* slots = g_graphs[dc.device & 63];
  GraphEntry* hit = nullptr;
  GraphEntry* victim = &slots[0];
  for (int i = 0; i < kGraphSlots; i++) {
    if ((slots[i].exec != nullptr || slots[i].unusable) && memcmp(&slots[i].key, &key, sizeof(key)) == 0) hit = &slots[i];
    if (slots[i].stamp < victim->stamp) victim = &slots[i];
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (hit == nullptr) {
    chol_setup(dc);
    cudaStream_t origin = dc.chain[0].main;
    if (victim->exec != nullptr) { cudaGraphExecDestroy(victim->exec); victim->exec = nullptr; }
    victim->key = key; victim->unusable = true; victim->launches = 0;
    hit = victim;
    const unsigned long long count0 = g_launch_count.load();
    if (cudaStreamBeginCapture(origin, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
      const int rc = chol_solve_tree_enqueue(A, b, n, firstc_dev, nranges, range_end, nlevels, level_first, info_dev,
                                             workspace, workspace_bytes, origin, private_scr, true);
      cudaGraph_t g = nullptr;
      const cudaError_t e = cudaStreamEndCapture(origin, &g);
      if (rc == DROID_OK && e == cudaSuccess && g != nullptr) {
        cudaGraphExec_t exec = nullptr;
        if (cudaGraphInstantiate(&exec, g, 0) == cudaSuccess && exec != nullptr) {
          hit->exec = exec; hit->unusable = false;
          hit->launches = g_launch_count.load() - count0;
        }
      }
      if (g != nullptr) cudaGraphDestroy(g);
      if (rc != DROID_OK && (rc == DROID_ERR_INVALID_ARGUMENT || rc == DROID_ERR_WORKSPACE_TOO_SMALL)) {
        cudaGetLastError();
        g_launch_count.store(count0);
        return rc;                                  // the caller's mistake, not the capture's
      }
    }
    cudaGetLastError();
    g_launch_count.store(count0);                   // captured, not launched
  }
  hit->stamp = ++g_graph_clock;
  if (hit->exec == nullptr)                          // capture not possible here: direct launches
    return chol_solve_tree_enqueue(A, b, n, firstc_dev, nranges, range_end, nlevels, level_first, info_dev, workspace,
                                   workspace_bytes, stream, private_scr, false);
  if (cudaGraphLaunch(hit->exec, st) != cudaSuccess) { g_last_error = cudaGetLastError(); return DROID_ERR_CUDA; }
  g_launch_count += hit->launches;
  return DROID_OK;
}

static int chol_solve_tree_enqueue(double* A, double* b, int n, const int* firstc_dev, int nranges,
                                   const int* range_end, int nlevels, const int* level_first, int* info_dev,
                                   void* workspace, size_t workspace_bytes, void* stream, bool private_scr,
                                   bool capturing) {
  if (n < 0 || !info_dev) return DROID_ERR_INVALID_ARGUMENT;
  if (n == 0) return DROID_OK;
  if (!A || !b) return DROID_ERR_INVALID_ARGUMENT;
  if (b != A + (size_t)n * n) return DROID_ERR_INVALID_ARGUMENT;  // rhs must be row n of A
  if (nranges < 1 || nranges > kMaxRanges || !range_end || nlevels < 1 || nlevels > nranges || !level_first ||
      (nranges > 1 && !firstc_dev))
    return DROID_ERR_INVALID_ARGUMENT;
  if (level_first[0] != 0 || level_first[nlevels] != nranges || range_end[nranges - 1] != n) return DROID_ERR_INVALID_ARGUMENT;
  for (int l = 0; l < nlevels; l++)
    if (level_first[l + 1] <= level_first[l] || level_first[l + 1] - level_first[l] > kMaxChains) return DROID_ERR_INVALID_ARGUMENT;
  const int nchains = nranges - 1;      // ranges besides the last one (sizes the per-range tables: nchains + 1 entries)
  cudaStream_t st = (cudaStream_t)stream;
  static const bool trace_env = []() { const char* e = getenv("DROID_CHOL_TRACE"); return e && e[0] == '1'; }();
  const bool trace0 = trace_env && !capturing;
  cudaEvent_t tev0 = nullptr;
  if (trace0) { cudaEventCreate(&tev0); cudaEventRecord(tev0, st); }
  cudaMemsetAsync(info_dev, 0, sizeof(int), st);
  if (n <= kSmallMax) {
    if (nranges > 1) return DROID_ERR_INVALID_ARGUMENT;
    if (dev_ctx().first(kOnceCholSmall))
      cudaFuncSetAttribute(chol_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmallSmemMax);
    const size_t smem = chol_small_smem(n);
    DROID_LAUNCH chol_small_kernel<<<1, ((n + 1 + 31) / 32) * 32, smem, st>>>(A, n, info_dev, 0, 0.f, 0.f);
    return check_launch();
  }
  DevCtx& dc = dev_ctx();
  std::lock_guard<std::recursive_mutex> call_lock(dc.call);   // the side streams / events below are shared per device
  chol_setup(dc);
  const bool vec = (n % 2 == 0) && (((uintptr_t)A & 15) == 0);
  const int nb = (n + 63) / 64;
  // ranges in elimination-tree layout: range r = [range_end[r-1], range_end[r]), levels = runs of consecutive ranges
  Ranges rg;
  rg.k = nchains;
  rg.b[0] = 0;
  for (int c = 0; c < nranges; c++) {
    const int e = range_end[c];
    if (e <= rg.b[c] || e > n || (e & 1 && e != n)) return DROID_ERR_INVALID_ARGUMENT;
    rg.b[c + 1] = e;
  }
  const int sep = nlevels > 1 ? rg.b[level_first[1]] : n, seprows = n - sep;   // rows behind level 0
  if (nranges > 1 && (!vec || n > kBwdPersistMax)) return DROID_ERR_INVALID_ARGUMENT;
  const size_t ws_head = chol_ws_head(n, nchains) + chol_ws_linv(n);
  if (!workspace || workspace_bytes < ws_head) return DROID_ERR_WORKSPACE_TOO_SMALL;
  // scratch blocks of every level that has more than one range and something behind it: one (shared by the ranges at
  // odd positions) or one per range but the first
  size_t scr_off[kMaxRanges + 1] = {}, scr_stride[kMaxRanges + 1] = {}, scr_total = 0;
  for (int l = 0; l + 1 < nlevels; l++) {
    scr_off[l] = scr_total;
    const size_t S = (size_t)(n - rg.b[level_first[l + 1]]);
    const int cnt = level_first[l + 1] - level_first[l];
    scr_stride[l] = (S + 1) * S;
    if (cnt > 1) scr_total += (private_scr ? (size_t)(cnt - 1) : 1) * scr_stride[l];
  }
  if (workspace_bytes < ws_head + scr_total * sizeof(double)) return DROID_ERR_WORKSPACE_TOO_SMALL;
  // Envelope variant: the launches keep their dense shapes (the host never learns the band structure), CTAs whose
  // rows or columns lie outside the envelope return at once.  64x64 tiles throughout: the active window is narrow.
  int* arrive = reinterpret_cast<int*>(workspace);
  int* hi64 = nullptr;
  int* lo64 = nullptr;
  cudaMemsetAsync(arrive, 0, (size_t)(nchains + 1) * (nb + 1) * sizeof(int), st);
  if (firstc_dev != nullptr) {
    hi64 = arrive + (size_t)(nchains + 1) * (nb + 1);
    lo64 = hi64 + (size_t)(nchains + 1) * nb;
    DROID_LAUNCH chol_env_kernel<<<nchains + 1, 1024, 0, st>>>(firstc_dev, n, rg, hi64, lo64);
  }
  double* scr = nullptr;
  // the backward sweep of envelope / chain systems runs in one persistent CTA per range and takes the diagonal blocks'
  // inverses from the panel kernels (dense systems of this size keep the multi-CTA sweep: 0.9 vs 2.0 ms at n = 5994)
  const bool persist_bwd = vec && firstc_dev != nullptr && n <= kBwdPersistMax;
  double* linv = persist_bwd ? reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(workspace) + chol_ws_head(n, nchains)) : nullptr;
  if (scr_total > 0) {
    scr = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(workspace) + chol_ws_head(n, nchains) + chol_ws_linv(n));
    cudaMemsetAsync(scr, 0, scr_total * sizeof(double), st);
  }
  // Look-ahead: while the bulk of trailing update J runs on the chain's main stream, a second (high-priority) stream
  // updates the next panel's 128 columns first and factorises them, so the serial panel chain hides behind the
  // SYRK for as long as the SYRK is the longer of the two.
  //   side : P(0) F(0) P(1) [wait R(0)] F(1) P(2) [wait R(1)] F(2) ...      P = panel, F = first block column
  //   main :      [wait P(0)] R(0)   [wait P(1)] R(1) ...                    R = rest of the trailing update
  // Chain mode runs one such pair of streams per chain, all forked from the caller's stream.
  ChainStreams* ctx = dc.chain;
  cudaEvent_t ev_start = dc.chol_start;
  // one outer panel (128 columns) of one range: factorise it, update what lies below / to the right
  auto step = [&](ChainState& cs) {
    const int J = cs.J;
    const int kb = (cs.Jend - J) < PB ? (cs.Jend - J) : PB;
    auto lookahead_ok = [&](int JJ) {
      const int r0 = JJ + PB;
      if (!vec || r0 + PB > cs.Jend) return false;   // the next panel is a full one of the same range
      return n + 1 - r0 >= 4 * PB;   // at least a few tiles of trailing matrix left to overlap with
    };
    cudaStream_t m = cs.main;
    ChainCtx& cx = *cs.cx;
    if (cs.on_side && !lookahead_ok(J)) {   // leave look-ahead mode: everything back on the main stream
      cudaEventRecord(cx.join, cx.side);
      cudaStreamWaitEvent(m, cx.join, 0);
      cs.on_side = false;
    }
    cudaStream_t ps = cs.on_side ? cx.side : m;
    if (cs.prof) prof_mark(DROID_PROF_CHOL_PANEL, true, ps);
    for (int j0 = J; j0 < J + kb; j0 += NB) {
      const int jb = (J + kb - j0) < NB ? (J + kb - j0) : NB;
      const int rem_rows = n + 1 - (j0 + jb);  // >= 1 (the rhs row)
      // second half of the outer panel: the rank-64 update by the first half is applied by the panel kernel itself
      const bool fused = j0 > J && vec && (jb & 1) == 0;
      DROID_LAUNCH chol_panel_kernel<<<(rem_rows + PRT - 1) / PRT + (linv ? 1 : 0), kPanelThreads,
                                       fused ? kPanelSmem2 : kPanelSmem1, ps>>>(
          A, n, j0, jb, fused ? NB : 0, info_dev, hi64 ? cs.hi64 : nullptr, cs.arrive, vec ? 1 : 0, cs.ca.hole_lo,
          cs.ca.hole_hi, linv, cs.ca.lo64r);
      const int rest = J + kb - (j0 + jb);     // columns of this outer panel still to factorise
      if (rest > 0 && !(vec && rest <= NB && (rest & 1) == 0)) {
        // (only when the next sub-panel cannot fuse it) rank-64 update of the panel's remaining columns
        const int t64 = (rem_rows + 63) / 64;
        if (vec) DROID_LAUNCH chol_syrk_kernel<64, true><<<t64, 256, kSyrk64, ps>>>(A, n, j0, jb, rest, hi64 ? cs.hi64 : nullptr, cs.ca, 1 << 30, 0);
        else DROID_LAUNCH chol_syrk_kernel<64, false><<<t64, 256, kSyrk64, ps>>>(A, n, j0, jb, rest, hi64 ? cs.hi64 : nullptr, cs.ca, 1 << 30, 0);
      }
    }
    if (cs.prof) prof_mark(DROID_PROF_CHOL_PANEL, false, ps);
    const int r0 = J + kb;
    cs.J = r0;
    if (r0 >= cs.Jend) cs.finished = true;
    if (r0 >= n) return;
    const int rr = n + 1 - r0;
    const int* hi = hi64 ? cs.hi64 : nullptr;
    // tiles of 64 rows counted from r0; those entirely inside the hole are not enumerated
    int t64 = (rr + 63) / 64, na = 1 << 30, nskip = 0;
    if (cs.ca.hole_hi > cs.ca.hole_lo) {
      na = cs.ca.hole_lo > r0 ? (cs.ca.hole_lo - r0 + 63) / 64 : 0;
      nskip = (cs.ca.hole_hi - r0) / 64 - na;
      if (nskip < 0) nskip = 0;
      t64 -= nskip;
    }
    const int t128 = (rr + 127) / 128;
    if (cs.on_side) {
      const bool big = hi == nullptr && t128 * (t128 + 1) / 2 >= dc.sm_count;
      cudaEventRecord(cx.p, cx.side);
      if (cs.r_recorded) cudaStreamWaitEvent(cx.side, cx.r, 0);    // R(J-1)
      if (big) DROID_LAUNCH chol_syrk_kernel<128, true><<<t128, 256, kSyrk128, cx.side>>>(A, n, J, kb, PB, hi, cs.ca, 1 << 30, 0);
      else DROID_LAUNCH chol_syrk_kernel<64, true><<<2 * t64, 256, kSyrk64, cx.side>>>(A, n, J, kb, PB, hi, cs.ca, na, nskip);
      cudaStreamWaitEvent(m, cx.p, 0);
      {
        if (cs.prof) prof_mark(DROID_PROF_CHOL_SYRK, true, m);
        if (big) DROID_LAUNCH chol_syrk_kernel<128, true><<<(t128 - 1) * t128 / 2, 256, kSyrk128, m>>>(A, n, J, kb, -1, hi, cs.ca, 1 << 30, 0);
        else if (t64 > 2) DROID_LAUNCH chol_syrk_kernel<64, true><<<(t64 - 2) * (t64 - 1) / 2, 256, kSyrk64, m>>>(A, n, J, kb, -2, hi, cs.ca, na, nskip);
        if (cs.prof) prof_mark(DROID_PROF_CHOL_SYRK, false, m);
      }
      cudaEventRecord(cx.r, m);
      cs.r_recorded = true;
    } else {
      if (cs.prof) prof_mark(DROID_PROF_CHOL_SYRK, true, m);
      if (hi == nullptr && t128 * (t128 + 1) / 2 >= dc.sm_count) {
        if (vec) DROID_LAUNCH chol_syrk_kernel<128, true><<<t128 * (t128 + 1) / 2, 256, kSyrk128, m>>>(A, n, J, kb, 0, hi, cs.ca, 1 << 30, 0);
        else DROID_LAUNCH chol_syrk_kernel<128, false><<<t128 * (t128 + 1) / 2, 256, kSyrk128, m>>>(A, n, J, kb, 0, hi, cs.ca, 1 << 30, 0);
      } else {
        if (vec) DROID_LAUNCH chol_syrk_kernel<64, true><<<t64 * (t64 + 1) / 2, 256, kSyrk64, m>>>(A, n, J, kb, 0, hi, cs.ca, na, nskip);
        else DROID_LAUNCH chol_syrk_kernel<64, false><<<t64 * (t64 + 1) / 2, 256, kSyrk64, m>>>(A, n, J, kb, 0, hi, cs.ca, na, nskip);
      }
      if (cs.prof) prof_mark(DROID_PROF_CHOL_SYRK, false, m);
    }
  };
  // range c at position `idx` of level `lv`: it must leave alone the later ranges of its own level (the hole) and sends
  // its updates of everything behind the level either to A (even idx) or to the level's scratch block (odd idx): two
  // ranges of the same parity never touch the same higher-level block (neighbours in the path order alternate)
  auto open_range = [&](ChainState& cs, int c, int lv, int idx, cudaStream_t m, ChainCtx* cx, bool prof) {
    cs.J = rg.b[c]; cs.Jend = rg.b[c + 1];
    cs.main = m; cs.cx = cx; cs.prof = prof && !capturing; cs.r_recorded = false;
    cs.finished = cs.J >= cs.Jend;
    const int behind = lv + 1 < nlevels ? rg.b[level_first[lv + 1]] : n;     // first row behind this level
    cs.ca.hole_lo = cs.ca.hole_hi = 0; cs.ca.sep = behind; cs.ca.scr = nullptr; cs.ca.lds = n - behind;
    if (lv + 1 < nlevels) {
      cs.ca.hole_lo = rg.b[c + 1]; cs.ca.hole_hi = behind;
      if (scr != nullptr && level_first[lv + 1] - level_first[lv] > 1) {
        if (private_scr) { if (idx > 0) cs.ca.scr = scr + scr_off[lv] + (size_t)(idx - 1) * scr_stride[lv]; }
        else if (idx & 1) cs.ca.scr = scr + scr_off[lv];
      }
    }
    cs.hi64 = hi64 ? hi64 + (size_t)c * nb : nullptr;
    cs.ca.lo64r = lo64 ? lo64 + (size_t)c * nb : nullptr;
    cs.arrive = arrive + (size_t)c * (nb + 1);
    const int r0 = cs.J + PB;
    cs.on_side = vec && r0 + PB <= cs.Jend && n + 1 - r0 >= 4 * PB;
    if (cs.on_side) {
      cudaEventRecord(cx->fork, m);
      cudaStreamWaitEvent(cx->side, cx->fork, 0);
    }
  };
  auto close_range = [&](ChainState& cs) {
    if (cs.on_side) {
      cudaEventRecord(cs.cx->join, cs.cx->side);
      cudaStreamWaitEvent(cs.main, cs.cx->join, 0);
      cs.on_side = false;
    }
  };
  // DROID_CHOL_TRACE=1: device time of every stage of the solve (events on the caller's stream; synchronises; debug aid)
  const bool trace = trace0;
  cudaEvent_t tev[2 * kMaxRanges + 8];
  const char* tname[2 * kMaxRanges + 8];
  int ntev = 0;
  auto tmark = [&](const char* what) {
    if (!trace) return;
    cudaEventCreate(&tev[ntev]);
    cudaEventRecord(tev[ntev], st);
    tname[ntev++] = what;
  };
  if (trace) { tev[ntev] = tev0; tname[ntev++] = "set-up (memsets, envelope summary)"; }
  for (int lv = 0; lv < nlevels; lv++) {
    const int r_lo = level_first[lv], cnt = level_first[lv + 1] - r_lo;
    ChainState cs[kMaxChains];
    tmark("level: factorise");
    if (cnt == 1) {
      open_range(cs[0], r_lo, lv, 0, st, &ctx[0], true);
      while (!cs[0].finished) step(cs[0]);
      close_range(cs[0]);
      continue;
    }
    cudaEventRecord(ev_start, st);
    for (int c = 0; c < cnt; c++) {
      cudaStream_t m = c == 0 ? st : ctx[c].main;
      if (c > 0) cudaStreamWaitEvent(m, ev_start, 0);
      open_range(cs[c], r_lo + c, lv, c, m, &ctx[c], c == 0);
    }
    // the ranges' launches are issued round-robin so that all of them are in flight from the start
    for (bool any = true; any;) {
      any = false;
      for (int c = 0; c < cnt; c++) {
        if (cs[c].finished) continue;
        step(cs[c]);
        any = true;
      }
    }
    for (int c = 0; c < cnt; c++) {
      close_range(cs[c]);
      if (c > 0) {
        cudaEventRecord(ctx[c].done, cs[c].main);
        cudaStreamWaitEvent(st, ctx[c].done, 0);
      }
    }
    tmark("level: join + merge");
    if (lv + 1 < nlevels && scr != nullptr) {
      const int behind = rg.b[level_first[lv + 1]];
      DROID_LAUNCH chol_merge_kernel<<<n - behind + 1, 256, 0, st>>>(A, n, behind, scr + scr_off[lv], private_scr ? cnt - 1 : 1,
                                                                       scr_stride[lv], n - behind);
    }
  }
  if (!capturing) prof_mark(DROID_PROF_CHOL_BWD, true, st);
  struct BwdEnd { bool on; cudaStream_t s; ~BwdEnd() { if (on) prof_mark(DROID_PROF_CHOL_BWD, false, s); } } bwd_end{!capturing, st};
  if (persist_bwd) {
    for (int lv = nlevels - 1; lv >= 0; lv--) {
      const int cnt = level_first[lv + 1] - level_first[lv];
      tmark("backward: level");
      DROID_LAUNCH chol_bwd_persist_kernel<<<cnt, kBwdPThreads, bwd_persist_smem(n), st>>>(A, n, lo64, rg, level_first[lv], linv);
      if (lv > 0)
        DROID_LAUNCH chol_bwd_wide_kernel<<<dim3((rg.b[level_first[lv]] + 63) / 64, cnt), 256, 0, st>>>(A, n, lo64, rg,
                                                                                                       level_first[lv]);
    }
    if (trace) {
      tmark("end");
      cudaStreamSynchronize(st);
      for (int i = 0; i + 1 < ntev; i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, tev[i], tev[i + 1]);
        fprintf(stderr, "[chol trace] %-40s %8.1f us\n", tname[i], ms * 1e3f);
      }
      for (int i = 0; i < ntev; i++) cudaEventDestroy(tev[i]);
    }
    return check_launch();
  }
  for (int J0 = ((n - 1) / SB) * SB; J0 >= 0; J0 -= SB) {
    const int JB = (n - J0) < SB ? (n - J0) : SB;
    DROID_LAUNCH chol_bwd_diag_kernel<<<1, 256, kBwdSmem, st>>>(A, n, J0, JB, vec ? 1 : 0);
    if (J0 > 0) DROID_LAUNCH chol_bwd_gemv_kernel<<<(J0 + NB - 1) / NB, 256, 0, st>>>(A, n, J0, JB, lo64);
  }
  return check_launch();
}
