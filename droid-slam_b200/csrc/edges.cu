// Edge selection on the device: the part of FactorGraph.add_proximity_factors
// (droid_slam/factor_graph.py:305-369) that follows the distance computation.  The reference runs it as Python
// double loops with an .item() per candidate; at 1000 keyframes that is seconds of host time around kernels that
// take microseconds.  Here the whole selection stays on the device and is stream-ordered:
//   1. prox_mask_kernel      d[i - rad < j] = inf, d[d > 100] = inf                                   (:318-319)
//   2. prox_exist_kernel     non-maximum suppression around the edges the graph already has           (:322-333)
//   3. prox_mandatory_kernel temporal neighbours (and stereo pairs) go in unconditionally             (:336-345)
//   4. radix sort of (distance, pair index)  -- cub::DeviceRadixSort, stable: ties in index order     (:347)
//   5. prox_greedy_kernel    one warp walks the sorted candidates, 32 at a time: accept the first one alive, suppress
//                            its neighbourhood (arithmetically for the 31 candidates in flight, through a byte map
//                            for later batches), until the distance exceeds thresh or the graph is full (:348-367)
// The result is the reference's edge list, in the reference's order (bit-exact against its own output, tests/).
// Index arithmetic follows Python: d[(i - t0) * W + (j - t1)] with a negative index wraps around.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace droid {
namespace {

__device__ __forceinline__ long long wrap_index(long long idx, long long n) { return idx < 0 ? idx + n : idx; }

__global__ void prox_mask_kernel(const float* __restrict__ d, float* __restrict__ dm, int* __restrict__ vals,
                                 unsigned char* __restrict__ sup, int n, int W, int t0, int t1, int rad) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const int i = t0 + idx / W, j = t1 + idx % W;
  const float v = d[idx];
  dm[idx] = (i - rad < j || v > 100.0f) ? __int_as_float(0x7f800000) : v;
  vals[idx] = idx;
  sup[idx] = 0;
}

// |di| + |dj| <= max(min(|i - j| - 2, nms), 0), inside [t0, t) x [t1, t)
__device__ __forceinline__ int nms_radius(int i, int j, int nms) { return max(min(abs(i - j) - 2, nms), 0); }

__global__ void prox_exist_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int n_exist,
                                  float* __restrict__ dm, int W, int t, int t0, int t1, int nms) {
  const int e = blockIdx.x;
  if (e >= n_exist) return;
  const int i = (int)ii[e], j = (int)jj[e];
  const int r = nms_radius(i, j, nms), side = 2 * nms + 1;
  for (int c = threadIdx.x; c < side * side; c += blockDim.x) {
    const int di = c / side - nms, dj = c % side - nms;
    if (abs(di) + abs(dj) > r) continue;
    const int i1 = i + di, j1 = j + dj;
    if (i1 >= t0 && i1 < t && j1 >= t1 && j1 < t) dm[(size_t)(i1 - t0) * W + (j1 - t1)] = __int_as_float(0x7f800000);
  }
}

// number of mandatory entries of frames 0 .. x-1: s per frame (stereo) + 2 min(i, rad + 1) temporal neighbours
__host__ __device__ inline long long mandatory_before(long long x, int rad, int s) {
  const long long m = rad + 1;
  const long long tri = x <= m + 1 ? x * (x - 1) / 2 : m * (m + 1) / 2 + (x - m - 1) * m;
  return s * x + 2 * tri;
}

__global__ void prox_mandatory_kernel(float* __restrict__ dm, int64_t* __restrict__ out_ii, int64_t* __restrict__ out_jj,
                                      int* __restrict__ status, int n, int W, int t, int t0, int t1, int rad, int stereo) {
  const int i = t0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= t) return;
  long long o = mandatory_before(i, rad, stereo ? 1 : 0) - mandatory_before(t0, rad, stereo ? 1 : 0);
  auto kill = [&](int j) {
    const long long idx = wrap_index((long long)(i - t0) * W + (j - t1), n);
    if (idx < 0 || idx >= n) { atomicOr(status, 1); return; }      // Python would raise IndexError
    dm[idx] = __int_as_float(0x7f800000);
  };
  if (stereo) {
    out_ii[o] = i; out_jj[o] = i; o++;
    kill(i);
  }
  for (int j = max(i - rad - 1, 0); j < i; j++) {
    out_ii[o] = i; out_jj[o] = j; o++;
    out_ii[o] = j; out_jj[o] = i; o++;
    kill(j);
  }
}

__global__ void __launch_bounds__(32) prox_greedy_kernel(const float* __restrict__ keys, const int* __restrict__ vals,
                                                         unsigned char* sup, int64_t* __restrict__ out_ii,
                                                         int64_t* __restrict__ out_jj, int* __restrict__ out_count,
                                                         int* __restrict__ status, int n, int W, int t, int t0, int t1,
                                                         int nms, float thresh, int max_factors, int mandatory,
                                                         int capacity) {
  const int lane = threadIdx.x;
  const int side = 2 * nms + 1;
  int count = mandatory;
  bool stop = false;
  const float kInf = __int_as_float(0x7f800000);
  // Software pipeline over batches of 32 sorted candidates: while batch b is walked, the (distance, pair) of batch b+2
  // and the suppression flag of batch b+1 are already on their way.  A flag fetched that early cannot know about the
  // candidates accepted while batch b is walked, so every acceptance is also tested arithmetically against the lane's
  // candidate of batch b+1 (like against the other candidates of batch b itself); acceptances of older batches reach
  // it through the byte map, which is fenced at the end of every batch.
  auto cand = [&](int k, float& key, int& idx) {
    key = k < n ? keys[k] : kInf;
    idx = k < n ? vals[k] : 0;
  };
  float key0, key1, key2;
  int idx0, idx1, idx2;
  cand(lane, key0, idx0);
  cand(32 + lane, key1, idx1);
  unsigned char sup0 = (lane < n && !(key0 > thresh)) ? __ldcg(sup + idx0) : (unsigned char)0;
  for (int base = 0; base < n && !stop; base += 32) {
    const int k = base + lane;
    cand(k + 64, key2, idx2);
    const bool close1 = (k + 32) < n && !(key1 > thresh);
    unsigned char sup1 = close1 ? __ldcg(sup + idx1) : (unsigned char)0;     // consumed in the next iteration
    const int ci = t0 + idx0 / W, cj = t1 + idx0 % W;
    const int ni = t0 + idx1 / W, nj = t1 + idx1 % W;
    const bool close = k < n && !(key0 > thresh);                // `if d[k] > thresh: continue`
    bool alive = close && sup0 == 0;
    if (__ballot_sync(0xffffffffu, !close) != 0u) stop = true;   // sorted: nothing behind this batch can pass
    unsigned live;
    while ((live = __ballot_sync(0xffffffffu, alive)) != 0u) {
      if (count > max_factors) { stop = true; break; }           // `if len(es) > max_factors: break`
      if (count + 2 > capacity) { if (lane == 0) atomicOr(status, 2); stop = true; break; }
      const int l = __ffs(live) - 1;
      const int i = __shfl_sync(0xffffffffu, ci, l), j = __shfl_sync(0xffffffffu, cj, l);
      if (lane == 0) {
        out_ii[count] = i; out_jj[count] = j;
        out_ii[count + 1] = j; out_jj[count + 1] = i;
      }
      count += 2;
      const int r = nms_radius(i, j, nms);
      // candidates of this batch and of the next one (their flags were read before this acceptance) ...
      if (alive && abs(ci - i) + abs(cj - j) <= r) alive = false;
      if (abs(ni - i) + abs(nj - j) <= r) sup1 = 1;
      if (lane == l) alive = false;
      // ... and everything behind them
      for (int c = lane; c < side * side; c += 32) {
        const int di = c / side - nms, dj = c % side - nms;
        if (abs(di) + abs(dj) > r) continue;
        const int i1 = i + di, j1 = j + dj;
        if (i1 >= t0 && i1 < t && j1 >= t1 && j1 < t) __stcg(sup + (size_t)(i1 - t0) * W + (j1 - t1), (unsigned char)1);
      }
    }
    __threadfence_block();
    __syncwarp();
    key0 = key1; idx0 = idx1; sup0 = sup1;
    key1 = key2; idx1 = idx2;
  }
  if (lane == 0) *out_count = count;
}

size_t sort_temp_bytes(int n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const float*)nullptr, (float*)nullptr, (const int*)nullptr,
                                  (int*)nullptr, n);
  return bytes;
}

struct ProxWs {
  float* dm; float* keys; int* vals; int* vals_sorted; unsigned char* sup; int* status; void* temp;
  size_t temp_bytes, total;
};

ProxWs carve_prox(void* base, int n) {
  ProxWs w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* q = base ? (void*)((char*)base + off) : nullptr;
    off += (bytes + 255) & ~(size_t)255;
    return q;
  };
  const size_t nn = n > 0 ? n : 1;
  w.status = (int*)take(256);
  w.dm = (float*)take(nn * sizeof(float));
  w.keys = (float*)take(nn * sizeof(float));
  w.vals = (int*)take(nn * sizeof(int));
  w.vals_sorted = (int*)take(nn * sizeof(int));
  w.sup = (unsigned char*)take(nn);
  w.temp_bytes = sort_temp_bytes((int)nn);
  w.temp = take(w.temp_bytes);
  w.total = off;
  return w;
}


// ------------------------------------------------------------------------------------------------
// FactorGraph.__filter_repeated_edges (droid_slam/factor_graph.py:43-54): keep[k] = (ii[k], jj[k]) is not one of the
// graph's existing edges (active + inactive).  The reference builds a Python set with one .item() per existing edge
// and tests every candidate with two more; here the existing pairs go into an open-addressing hash table in global
// memory (64-bit keys, linear probing, atomicCAS) and one CTA tests the candidates and compacts the survivors in
// their original order.  Duplicates INSIDE the candidate list are all kept, exactly like the reference (its set holds
// only the existing edges).
// ------------------------------------------------------------------------------------------------
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __forceinline__ unsigned long long pair_key(long long i, long long j) {
  return ((unsigned long long)(unsigned int)i << 32) | (unsigned long long)(unsigned int)j;
}
__device__ __forceinline__ unsigned int pair_hash(unsigned long long x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
  return (unsigned int)x;
}

__global__ void filter_insert_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int n,
                                     unsigned long long* __restrict__ table, unsigned int mask) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const unsigned long long key = pair_key(ii[e], jj[e]);
  unsigned int slot = pair_hash(key) & mask;
  while (true) {
    const unsigned long long prev = atomicCAS(&table[slot], kEmptyKey, key);
    if (prev == kEmptyKey || prev == key) return;
    slot = (slot + 1) & mask;
  }
}

__global__ void __launch_bounds__(1024) filter_query_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj,
                                                            int n, const unsigned long long* __restrict__ table,
                                                            unsigned int mask, int64_t* __restrict__ out_ii,
                                                            int64_t* __restrict__ out_jj, unsigned char* __restrict__ keep,
                                                            int* __restrict__ out_count) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int e = c0 + tid;
    int flag = 0;
    long long i = 0, j = 0;
    if (e < n) {
      i = ii[e]; j = jj[e];
      const unsigned long long key = pair_key(i, j);
      unsigned int slot = pair_hash(key) & mask;
      flag = 1;
      while (true) {
        const unsigned long long v = table[slot];
        if (v == key) { flag = 0; break; }
        if (v == kEmptyKey) break;
        slot = (slot + 1) & mask;
      }
      if (keep != nullptr) keep[e] = (unsigned char)flag;
    }
    // ordered compaction: exclusive scan of the flags over the CTA
    int x = flag;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const int pos = s_base + (warp > 0 ? s_warp[warp - 1] : 0) + x - flag;
    if (flag) { out_ii[pos] = i; out_jj[pos] = j; }
    __syncthreads();
    if (tid == 0) s_base += s_warp[31];
    __syncthreads();
  }
  if (tid == 0) *out_count = s_base;
}

static unsigned int filter_table_slots(int n_exist) {
  unsigned int cap = 64;
  while (cap < 2u * (unsigned int)(n_exist > 0 ? n_exist : 0) + 1u) cap <<= 1;
  return cap;
}

}  // namespace
}  // namespace droid

using namespace droid;

extern "C" size_t droid_proximity_workspace_bytes(int n_pairs) {
  if (n_pairs < 0) return 0;
  return carve_prox(nullptr, n_pairs).total + 256;
}

extern "C" int droid_proximity_mandatory_count(int t, int t0, int rad, int stereo) {
  if (t < t0 || t0 < 0 || rad < 0) return -1;
  return (int)(mandatory_before(t, rad, stereo ? 1 : 0) - mandatory_before(t0, rad, stereo ? 1 : 0));
}

extern "C" int droid_proximity_edges(const float* d, int t, int t0, int t1, int rad, int nms, float thresh,
                                     int max_factors, int stereo, const int64_t* ii_exist,
                                     const int64_t* jj_exist, int n_exist, int64_t* out_ii, int64_t* out_jj,
                                     int capacity, int* out_count, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  if (t0 < 0 || t1 < 0 || t < t0 || t < t1 || rad < 0 || nms < 0 || nms > 15 || n_exist < 0 || capacity < 0)
    return DROID_ERR_INVALID_ARGUMENT;
  if (!out_ii || !out_jj || !out_count || !workspace) return DROID_ERR_INVALID_ARGUMENT;
  if (n_exist > 0 && (!ii_exist || !jj_exist)) return DROID_ERR_INVALID_ARGUMENT;
  const long long nl = (long long)(t - t0) * (t - t1);
  if (nl > 0x7fffffffLL) return DROID_ERR_INVALID_ARGUMENT;
  const int n = (int)nl, W = t - t1;
  if (n > 0 && !d) return DROID_ERR_INVALID_ARGUMENT;
  const int mandatory = droid_proximity_mandatory_count(t, t0, rad, stereo);
  if (mandatory > capacity) return DROID_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)workspace & 255) != 0) {
    const size_t adj = 256 - ((uintptr_t)workspace & 255);
    if (workspace_bytes < adj) return DROID_ERR_WORKSPACE_TOO_SMALL;
    workspace = (char*)workspace + adj;
    workspace_bytes -= adj;
  }
  ProxWs w = carve_prox(workspace, n);
  if (workspace_bytes < w.total) return DROID_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(w.status, 0, sizeof(int), st);
  if (n > 0) {
    DROID_LAUNCH prox_mask_kernel<<<(n + 255) / 256, 256, 0, st>>>(d, w.dm, w.vals, w.sup, n, W, t0, t1, rad);
    if (n_exist > 0)
      DROID_LAUNCH prox_exist_kernel<<<n_exist, 64, 0, st>>>(ii_exist, jj_exist, n_exist, w.dm, W, t, t0, t1, nms);
  }
  if (t > t0)
    DROID_LAUNCH prox_mandatory_kernel<<<(t - t0 + 127) / 128, 128, 0, st>>>(w.dm, out_ii, out_jj, w.status, n, W, t, t0,
                                                                              t1, rad, stereo);
  if (n > 0) {
    size_t tb = w.temp_bytes;
    ++g_launch_count;
    if (cub::DeviceRadixSort::SortPairs(w.temp, tb, w.dm, w.keys, w.vals, w.vals_sorted, n, 0, 32, st) != cudaSuccess)
      return DROID_ERR_CUDA;
  }
  DROID_LAUNCH prox_greedy_kernel<<<1, 32, 0, st>>>(w.keys, w.vals_sorted, w.sup, out_ii, out_jj, out_count, w.status, n, W,
                                                    t, t0, t1, nms, thresh, max_factors, mandatory, capacity);
  return check_launch();
}

extern "C" size_t droid_filter_edges_workspace_bytes(int n_exist) {
  if (n_exist < 0) return 0;
  return (size_t)filter_table_slots(n_exist) * sizeof(unsigned long long) + 256;
}

extern "C" int droid_filter_repeated_edges(const int64_t* ii, const int64_t* jj, int n, const int64_t* ii_exist,
                                           const int64_t* jj_exist, int n_exist, int64_t* out_ii, int64_t* out_jj,
                                           unsigned char* keep, int* out_count, void* workspace, size_t workspace_bytes,
                                           void* stream) {
  if (n < 0 || n_exist < 0 || !out_count || !workspace) return DROID_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!ii || !jj || !out_ii || !out_jj)) return DROID_ERR_INVALID_ARGUMENT;
  if (n_exist > 0 && (!ii_exist || !jj_exist)) return DROID_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)workspace & 255) != 0) {
    const size_t adj = 256 - ((uintptr_t)workspace & 255);
    if (workspace_bytes < adj) return DROID_ERR_WORKSPACE_TOO_SMALL;
    workspace = (char*)workspace + adj;
    workspace_bytes -= adj;
  }
  const unsigned int slots = filter_table_slots(n_exist);
  if (workspace_bytes < (size_t)slots * sizeof(unsigned long long)) return DROID_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* table = reinterpret_cast<unsigned long long*>(workspace);
  cudaMemsetAsync(table, 0xff, (size_t)slots * sizeof(unsigned long long), st);
  if (n_exist > 0)
    DROID_LAUNCH filter_insert_kernel<<<(n_exist + 255) / 256, 256, 0, st>>>(ii_exist, jj_exist, n_exist, table, slots - 1);
  DROID_LAUNCH filter_query_kernel<<<1, 1024, 0, st>>>(ii, jj, n, table, slots - 1, out_ii, out_jj, keep, out_count);
  return check_launch();
}
