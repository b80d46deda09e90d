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

namespace droid {

static constexpr int NB = 64;
static constexpr int kSmallMax = 156;  // (n+1)*(n+1)*8 bytes <= ~197 KB of shared memory

// ------------------------------------------------------------------------------------------------
// small systems: whole solve in one CTA
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chol_small_kernel(double* __restrict__ A, int n,
                                                         int* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* S = reinterpret_cast<double*>(smem_raw);  // [(n+1)][ld], row n = rhs
  const int ld = n + 1;
  const int tid = threadIdx.x;
  __shared__ int fail;
  if (tid == 0) fail = 0;
  for (int e = tid; e < (n + 1) * n; e += 256) {
    const int r = e / n, c = e - r * n;
    if (c <= r) S[r * ld + c] = A[(size_t)r * n + c];
  }
  __syncthreads();
  for (int c = 0; c < n; c++) {
    const double d = S[c * ld + c];
    if (!(d > 0.0)) {
      if (tid == 0) fail = 1;
    }
    const double inv = 1.0 / sqrt(d);
    __syncthreads();
    // scale column c (rows c..n, row n is the rhs)
    for (int r = c + tid; r <= n; r += 256) S[r * ld + c] = (r == c) ? sqrt(d) : S[r * ld + c] * inv;
    __syncthreads();
    // trailing update of the lower triangle, rows c+1..n
    const int rem = n - c;  // rows c+1..n  -> rem rows, columns c+1..min(r, n-1)
    for (int e = tid; e < rem * rem; e += 256) {
      const int rr = e / rem, cc = e - rr * rem;
      const int r = c + 1 + rr, q = c + 1 + cc;
      if (q <= r && q < n) S[r * ld + q] -= S[r * ld + c] * S[q * ld + c];
    }
    __syncthreads();
  }
  // row n now holds y (L y = b).  Backward: L^T x = y, done by warp 0.
  if (tid < 32) {
    for (int c = n - 1; c >= 0; c--) {
      double s = 0.0;
      for (int t = c + 1 + tid; t < n; t += 32) s += S[t * ld + c] * S[n * ld + t];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (tid == 0) S[n * ld + c] = (S[n * ld + c] - s) / S[c * ld + c];
      __syncwarp();
    }
  }
  __syncthreads();
  for (int e = tid; e < (n + 1) * n; e += 256) {
    const int r = e / n, c = e - r * n;
    if (c <= r) A[(size_t)r * n + c] = S[r * ld + c];
  }
  if (tid == 0 && fail) *info = 1;
}

// ------------------------------------------------------------------------------------------------
// blocked path, step 1: factor the diagonal block (redundantly in every CTA, it is 64^3/3 flops) and
// solve this CTA's row block  X L_jj^T = A[rows, j0:j0+jb].
// grid.x = 1 + number of 64-row blocks in rows j0+jb .. n (row n = rhs).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chol_panel_kernel(double* __restrict__ A, int n, int j0, int jb,
                                                         int* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double (*D)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(smem_raw);
  double (*X)[NB + 1] = D + NB;
  __shared__ int fail;
  const int tid = threadIdx.x;
  if (tid == 0) fail = 0;
  for (int e = tid; e < jb * jb; e += 256) {
    const int r = e / jb, c = e - r * jb;
    D[r][c] = (c <= r) ? A[(size_t)(j0 + r) * n + j0 + c] : 0.0;
  }
  __syncthreads();
  for (int c = 0; c < jb; c++) {
    const double d = D[c][c];
    if (!(d > 0.0) && tid == 0) fail = 1;
    const double sq = sqrt(d), inv = 1.0 / sq;
    __syncthreads();
    for (int r = c + tid; r < jb; r += 256) D[r][c] = (r == c) ? sq : D[r][c] * inv;
    __syncthreads();
    const int rem = jb - c - 1;
    for (int e = tid; e < rem * rem; e += 256) {
      const int rr = e / rem, cc = e - rr * rem;
      if (cc <= rr) D[c + 1 + rr][c + 1 + cc] -= D[c + 1 + rr][c] * D[c + 1 + cc][c];
    }
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    for (int e = tid; e < jb * jb; e += 256) {
      const int r = e / jb, c = e - r * jb;
      if (c <= r) A[(size_t)(j0 + r) * n + j0 + c] = D[r][c];
    }
    if (tid == 0 && fail) *info = 1;
    return;
  }
  const int i0 = j0 + jb + (blockIdx.x - 1) * NB;
  const int rows = min(NB, n + 1 - i0);
  for (int e = tid; e < rows * jb; e += 256) {
    const int r = e / jb, c = e - r * jb;
    X[r][c] = A[(size_t)(i0 + r) * n + j0 + c];
  }
  __syncthreads();
  // right-looking triangular solve, one column per step
  for (int c = 0; c < jb; c++) {
    const double inv = 1.0 / D[c][c];
    for (int r = tid; r < rows; r += 256) X[r][c] *= inv;
    __syncthreads();
    const int rem = jb - c - 1;
    for (int e = tid; e < rows * rem; e += 256) {
      const int r = e / rem, t = c + 1 + (e - r * rem);
      X[r][t] -= X[r][c] * D[t][c];
    }
    __syncthreads();
  }
  for (int e = tid; e < rows * jb; e += 256) {
    const int r = e / jb, c = e - r * jb;
    A[(size_t)(i0 + r) * n + j0 + c] = X[r][c];
  }
}

// ------------------------------------------------------------------------------------------------
// blocked path, step 2: trailing update  C[bi][bj] -= X_bi X_bj^T  for all 64x64 tiles bi >= bj of
// rows/cols r0 = j0+jb .. (row n included as a row, never as a column).  One CTA per tile,
// 256 threads x (4x4) fp64 register tiles, operands staged k-major in shared memory.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chol_syrk_kernel(double* __restrict__ A, int n, int j0, int jb) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double (*Xi)[NB + 2] = reinterpret_cast<double (*)[NB + 2]>(smem_raw);  // [k][row]
  double (*Xj)[NB + 2] = Xi + NB;                                          // [k][col]
  // decode the lower-triangular tile index
  const int t = blockIdx.x;
  int bi = (int)((sqrtf(8.f * t + 1.f) - 1.f) * 0.5f);
  while ((bi + 1) * (bi + 2) / 2 <= t) bi++;
  while (bi * (bi + 1) / 2 > t) bi--;
  const int bj = t - bi * (bi + 1) / 2;
  const int r0 = j0 + jb;
  const int i0 = r0 + bi * NB, c0 = r0 + bj * NB;
  const int rows = min(NB, n + 1 - i0);
  const int cols = min(NB, n - c0);
  if (cols <= 0 || rows <= 0) return;
  const int tid = threadIdx.x;
  for (int e = tid; e < NB * NB; e += 256) {
    const int r = e / NB, k = e - r * NB;
    Xi[k][r] = (r < rows && k < jb) ? A[(size_t)(i0 + r) * n + j0 + k] : 0.0;
    Xj[k][r] = (r < cols && k < jb) ? A[(size_t)(c0 + r) * n + j0 + k] : 0.0;
  }
  __syncthreads();
  const int ty = tid >> 4, tx = tid & 15;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
#pragma unroll 8
  for (int k = 0; k < NB; k++) {
    const double2 a01 = *reinterpret_cast<const double2*>(&Xi[k][ty * 4]);
    const double2 a23 = *reinterpret_cast<const double2*>(&Xi[k][ty * 4 + 2]);
    const double2 b01 = *reinterpret_cast<const double2*>(&Xj[k][tx * 4]);
    const double2 b23 = *reinterpret_cast<const double2*>(&Xj[k][tx * 4 + 2]);
    const double av[4] = {a01.x, a01.y, a23.x, a23.y};
    const double bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
  }
#pragma unroll
  for (int a = 0; a < 4; a++) {
    const int r = ty * 4 + a;
    if (r >= rows) continue;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int c = tx * 4 + b;
      if (c >= cols) continue;
      if (c0 + c > i0 + r) continue;  // keep to the lower triangle
      A[(size_t)(i0 + r) * n + c0 + c] -= acc[a][b];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// backward sweep, one launch per 64-block from the last to the first:
//   x_j = L_jj^-T y_j  (every CTA redundantly, warp 0), then  y[i] -= sum_r L[j0+r][i] x_j[r], i < j0.
// y lives in row n of A and is overwritten by x.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) chol_bwd_kernel(double* __restrict__ A, int n, int j0, int jb) {
  __shared__ double D[NB][NB + 1];
  __shared__ double x[NB];
  const int tid = threadIdx.x;
  double* y = A + (size_t)n * n;
  for (int e = tid; e < jb * jb; e += 256) {
    const int r = e / jb, c = e - r * jb;
    D[r][c] = (c <= r) ? A[(size_t)(j0 + r) * n + j0 + c] : 0.0;
  }
  if (tid < jb) x[tid] = y[j0 + tid];
  __syncthreads();
  if (tid < 32) {
    for (int c = jb - 1; c >= 0; c--) {
      double s = 0.0;
      for (int t = c + 1 + tid; t < jb; t += 32) s += D[t][c] * x[t];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (tid == 0) x[c] = (x[c] - s) / D[c][c];
      __syncwarp();
    }
  }
  __syncthreads();
  const int i = blockIdx.x * 256 + tid;
  if (i < j0) {
    double s = 0.0;
#pragma unroll 8
    for (int r = 0; r < jb; r++) s = fma(A[(size_t)(j0 + r) * n + i], x[r], s);
    y[i] -= s;
  }
  if (blockIdx.x == 0 && tid < jb) y[j0 + tid] = x[tid];
}

}  // namespace droid

using namespace droid;

extern "C" size_t droid_chol_workspace_bytes(int n) {
  (void)n;
  return 256;  // the factorisation is in place; reserved for future use
}

extern "C" int droid_chol_solve_f64(double* A, double* b, int n, int* info_dev, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  (void)workspace; (void)workspace_bytes;
  if (n < 0 || !info_dev) return DROID_ERR_INVALID_ARGUMENT;
  if (n == 0) return DROID_OK;
  if (!A || !b) return DROID_ERR_INVALID_ARGUMENT;
  if (b != A + (size_t)n * n) return DROID_ERR_INVALID_ARGUMENT;  // rhs must be row n of A
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(info_dev, 0, sizeof(int), st);
  if (n <= kSmallMax) {
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(chol_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      attr_set = true;
    }
    const size_t smem = (size_t)(n + 1) * (n + 1) * sizeof(double);
    DROID_LAUNCH chol_small_kernel<<<1, 256, smem, st>>>(A, n, info_dev);
    return check_launch();
  }
  static bool attr2 = false;
  const size_t panel_smem = 2 * NB * (NB + 1) * sizeof(double);
  const size_t syrk_smem = 2 * NB * (NB + 2) * sizeof(double);
  if (!attr2) {
    cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)panel_smem);
    cudaFuncSetAttribute(chol_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)syrk_smem);
    attr2 = true;
  }
  for (int j0 = 0; j0 < n; j0 += NB) {
    const int jb = (n - j0) < NB ? (n - j0) : NB;
    const int rem_rows = n + 1 - (j0 + jb);  // >= 1 (the rhs row)
    const int nrb = (rem_rows + NB - 1) / NB;
    DROID_LAUNCH chol_panel_kernel<<<1 + nrb, 256, panel_smem, st>>>(A, n, j0, jb, info_dev);
    const int ncb = (n - (j0 + jb) + NB - 1) / NB;  // column blocks that remain
    if (ncb > 0) {
      // tiles (bi, bj), bj < ncb, bi >= bj, bi < nrb
      const int ntile = nrb * (nrb + 1) / 2;
      DROID_LAUNCH chol_syrk_kernel<<<ntile, 256, syrk_smem, st>>>(A, n, j0, jb);
    }
  }
  for (int j0 = ((n - 1) / NB) * NB; j0 >= 0; j0 -= NB) {
    const int jb = (n - j0) < NB ? (n - j0) : NB;
    const int g = j0 > 0 ? (j0 + 255) / 256 : 1;
    DROID_LAUNCH chol_bwd_kernel<<<g, 256, 0, st>>>(A, n, j0, jb);
  }
  return check_launch();
}
