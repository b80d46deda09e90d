/* droid_b200.h — C ABI of libdroid_b200.so: the sm_100a implementation of DROID-SLAM's
 * `droid_backends` operator surface (reference: /root/reference/src/droid.cpp:237-250).
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in _host; all tensors are contiguous
 *    in the layout stated next to the argument (the reference enforces contiguity only,
 *    src/droid.cpp:84-85);
 *  - `stream` is a cudaStream_t passed as void*; every call is stream-ordered and never allocates
 *    caller-visible memory (outputs and workspaces are supplied by the caller; the Python host
 *    `droid_backends` allocates them with torch on the current stream).  No call synchronises the host,
 *    with ONE documented exception: droid_ba waits on `stream` once (a) when the reduced camera system has
 *    >= 1536 rows or a reduce_fn is set -- the host needs the chain layout / packed size (24 ints) to shape
 *    the launch sequence -- and (b) at the end of the call when DROID_BA_FLAG_CHECK is set and (a) did not
 *    already deliver the status word;
 *  - return value: DROID_OK or a negative DROID_ERR_* code (the Python host raises RuntimeError,
 *    mirroring TORCH_CHECK -> RuntimeError of the reference);
 *  - index tensors are int64 (torch.long), as in the reference;
 *  - dtype codes select the element type of volume / fmap tensors (the reference dispatches
 *    AT_DISPATCH_FLOATING_TYPES_AND_HALF, src/correlation_kernels.cu:145).
 */
#ifndef DROID_B200_H_
#define DROID_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DROID_OK 0
#define DROID_ERR_INVALID_ARGUMENT (-1) /* bad shape / null pointer / unsupported radius */
#define DROID_ERR_UNSUPPORTED_DTYPE (-2)
#define DROID_ERR_WORKSPACE_TOO_SMALL (-3)
#define DROID_ERR_CUDA (-4) /* a launch failed; see droid_last_cuda_error() */
#define DROID_ERR_NOT_IMPLEMENTED (-5)
#define DROID_ERR_BA_INPUT (-6) /* droid_ba with DROID_BA_FLAG_CHECK: the device-side validation failed (status word) */
#define DROID_ERR_STATIC_SCRATCH_TOO_SMALL (-7) /* a *_noalloc entry point: size beyond its built-in scratch */

#define DROID_DTYPE_F32 0
#define DROID_DTYPE_F16 1
#define DROID_DTYPE_F64 2

/* Library / build identification.  Returns e.g. "droid_b200 0.2 sm_100a". */
const char* droid_version(void);
/* cudaGetErrorString of the last failing launch seen by this library (thread-unsafe, debug aid). */
const char* droid_last_cuda_error(void);

/* Number of kernels this library has launched since it was loaded (bench.py's gpu_launches). */
unsigned long long droid_launch_count(void);

/* Per-phase device timing for bench.py's roofline: when enabled, the multi-kernel entry points record
 * CUDA events on the caller's stream around each phase.  droid_prof_read waits for the recorded events,
 * adds the elapsed milliseconds and interval counts per phase into ms[]/counts[] (nphase entries) and
 * clears the records.  Phases: */
#define DROID_PROF_BA_PREP 0    /* graph bookkeeping + edge poses + clearing the system       */
#define DROID_PROF_BA_ACCUM 1   /* ba_accum_kernel: H, v, C, w                                 */
#define DROID_PROF_BA_SCHUR 2   /* ba_schur_kernel: S = E Q E^T, v_S                           */
#define DROID_PROF_BA_REDUCE 3  /* multi-GPU reduction callback                                */
#define DROID_PROF_BA_SOLVE 4   /* damping + Cholesky factor + solves                          */
#define DROID_PROF_BA_UPDATE 5  /* dz back-substitution + retractions                          */
#define DROID_PROF_CORR_GEMM 6  /* corr_build                                                  */
#define DROID_PROF_CHOL_PANEL 8 /* inside BA_SOLVE: panel factorisations                         */
#define DROID_PROF_CHOL_SYRK 9  /* inside BA_SOLVE: trailing updates                               */
#define DROID_PROF_CHOL_BWD 10  /* inside BA_SOLVE: backward substitution                         */
#define DROID_PROF_NPHASE 12
int droid_prof_enable(int on);
int droid_prof_read(double* ms, int* counts, int nphase);

/* ---- correlation lookup ---------------------------------------------------------------------
 * Replaces corr_index_forward  (src/droid.cpp:170-178, src/correlation_kernels.cu:19-70,126-155).
 *   volume [B,H1,W1,H2,W2] dtype; coords [B,2,H1,W1] f32 (x,y), already scaled to this level;
 *   out    [B,2r+1,2r+1,H1,W1] dtype, FULLY WRITTEN (no pre-zeroing required).
 *   out[b, i, j, y, x] = bilinear(volume[b,y,x], x0 + i - r, y0 + j - r), zero outside. */
int droid_corr_index_forward(const void* volume, const float* coords, void* out, int dtype, int B,
                             int H1, int W1, int H2, int W2, int radius, void* stream);

/* Replaces corr_index_backward (src/droid.cpp:180-191, src/correlation_kernels.cu:73-124,157-185).
 *   corr_grad [B,2r+1,2r+1,H1,W1]; volume_grad [B,H1,W1,H2,W2] FULLY WRITTEN (zeros + window). */
int droid_corr_index_backward(const float* coords, const void* corr_grad, void* volume_grad,
                              int dtype, int B, int H1, int W1, int H2, int W2, int radius,
                              void* stream);

/* Fused 4-level pyramid lookup (one launch instead of 4; what CorrBlock.__call__ does,
 * droid_slam/modules/corr.py:40-50).  volumes[l] is level l = [B,H1,W1,H2>>l,W2>>l];
 * coords [B,2,H1,W1] are LEVEL-0 coordinates (divided by 2^l inside, as corr.py:47);
 * out [B, L*(2r+1)^2, H1, W1] = torch.cat of the per-level results on dim 1. */
int droid_corr_index_pyramid(const void* const* volumes_host, int num_levels, const float* coords,
                             void* out, int dtype, int B, int H1, int W1, int H2, int W2,
                             int radius, void* stream);

/* ---- on-the-fly ("alt") correlation ----------------------------------------------------------
 * Replaces altcorr_forward (src/droid.cpp:193-203, src/altcorr_kernel.cu:27-149,290-319).
 *   fmap1 [B,H1,W1,C], fmap2 [B,H2,W2,C] channel-last dtype; coords [B,N,H1,W1,2] f32 (x,y);
 *   out [B,N,(2r+1)^2,H1,W1] dtype, fully written; channel = iy + (2r+1)*ix. */
int droid_altcorr_forward(const void* fmap1, const void* fmap2, const float* coords, void* out,
                          int dtype, int B, int N, int H1, int W1, int H2, int W2, int C,
                          int radius, void* stream);

/* The same op with a caller-supplied workspace (droid_altcorr_forward_workspace_bytes; 0 = not needed for this shape),
 * which unlocks the tensor-core kernel (csrc/altcorr_tc.cu: radius 3, C = 128, W2 <= 128): the target map is copied to
 * fp16 in the workspace and the source map is converted tile by tile inside the kernel, both with a round-trip check --
 * AltCorrBlock hands `.float()` copies of an fp16 pyramid (droid_slam/modules/corr.py:96-120), so the conversion is
 * exact -- and correlated with tcgen05.mma; when a value is not an fp16 number a device-side flag routes the call to the
 * CUDA-core kernel instead, without a host round trip.  Same result contract as droid_altcorr_forward (<= 1e-5 of the
 * largest value).  This is the entry point the Python host's altcorr_forward binds. */
int droid_altcorr_forward_ws(const void* fmap1, const void* fmap2, const float* coords, void* out,
                             int dtype, int B, int N, int H1, int W1, int H2, int W2, int C,
                             int radius, void* workspace, size_t workspace_bytes, void* stream);
size_t droid_altcorr_forward_workspace_bytes(int B, int N, int H1, int W1, int H2, int W2, int C, int radius);

/* Replaces altcorr_backward (src/droid.cpp:205-217, src/altcorr_kernel.cu:152-286,321-356); f32 only
 * like the reference.  fmap1_grad/fmap2_grad fully written; coords_grad [B,N,H1,W1,2] is all zero
 * (the reference never writes it). */
int droid_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords,
                           const float* corr_grad, float* fmap1_grad, float* fmap2_grad,
                           float* coords_grad, int B, int N, int H1, int W1, int H2, int W2, int C,
                           int radius, void* stream);

/* Indexed alt-correlation: same result as altcorr_forward on
 * fmap1 = pyr0[ii], fmap2 = pyrL[jj] (modules/corr.py:106-125) without the gathered copies.
 *   pyr0 [F,H1,W1,C], pyrL [F,H2,W2,C] dtype; ii,jj [B] int64; output f32. */
/* fp16 pyramids with C = 128 and W2 <= 128 take the tensor-core kernel (TMA straight from the pyramid, no copies). */
int droid_altcorr_forward_indexed(const void* pyr0, const void* pyrL, const int64_t* ii,
                                  const int64_t* jj, const float* coords, float* out, int dtype,
                                  int F, int B, int N, int H1, int W1, int H2, int W2, int C,
                                  int radius, void* stream);

/* Every level of an AltCorrBlock in one launch (extension; droid_slam/modules/corr.py:106-125 is four calls + cat):
 *   pyr_host[l]: device pointer of level l, [F, H1>>l, W1>>l, C] channel-last; coords [B,N,H1,W1,2] at LEVEL 0 (divided
 *   by 2^l inside, exactly as corr.py:116); out [B, N, num_levels*49, H1, W1] f32, level l at channels 49 l .. 49 l + 48.
 * Tensor-core kernel only (f16, C = 128, radius 3, W1 <= 128, num_levels <= 4): DROID_ERR_NOT_IMPLEMENTED otherwise --
 * the caller then uses droid_altcorr_forward_indexed per level (the Python host does). */
int droid_altcorr_pyramid_indexed(const void* const* pyr_host, int num_levels, const int64_t* ii,
                                  const int64_t* jj, const float* coords, float* out, int dtype, int F, int B,
                                  int N, int H1, int W1, int C, int radius, void* stream);

/* ---- all-pairs correlation volume + pyramid (tcgen05/TMEM GEMM) -------------------------------
 * Replaces CorrBlock.corr + 3x avg_pool2d (droid_slam/modules/corr.py:24-38,63-71).
 *   fmap1,fmap2 [B,C,HW] f16 (C=128), UNSCALED (the /4 on each side is applied inside, in f16,
 *   exactly as corr.py:67-68);  lvl0..lvl3: [B,HW,H>>l,W>>l] f16.  H%8==0, W%8==0. */
int droid_corr_build(const void* fmap1, const void* fmap2, void* lvl0, void* lvl1, void* lvl2,
                     void* lvl3, int B, int C, int H, int W, void* workspace,
                     size_t workspace_bytes, void* stream);
size_t droid_corr_build_workspace_bytes(int B, int C, int H, int W);

/* ---- geometry --------------------------------------------------------------------------------
 * frame_distance (src/droid.cpp:120-136, src/droid_kernels.cu:518-657,1438-1460).
 *   poses [*,7] (tx,ty,tz,qx,qy,qz,qw), disps [nbuf,ht,wd], intrinsics [4], ii,jj [n], dist [n]. */
int droid_frame_distance(const float* poses, const float* disps, const float* intrinsics,
                         const int64_t* ii, const int64_t* jj, float* dist, int n, int ht, int wd,
                         float beta, void* stream);

/* Same result, bit for bit, with the pair-independent per-pixel work (normalised image coordinates, the `total`
 * reduction) done once per call in a small workspace instead of once per pixel-pair: what the Python host calls.
 *   workspace: droid_frame_distance_workspace_bytes(ht, wd) bytes of device memory, 16-byte aligned. */
size_t droid_frame_distance_workspace_bytes(int ht, int wd);
int droid_frame_distance_ws(const float* poses, const float* disps, const float* intrinsics,
                            const int64_t* ii, const int64_t* jj, float* dist, int n, int ht, int wd,
                            float beta, void* workspace, size_t workspace_bytes, void* stream);

/* frame_distance WITHOUT any allocation and without a caller-supplied workspace.  This entry point has NO
 * counterpart in the reference: it was added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result, bit for
 * bit, as droid_frame_distance / droid_frame_distance_ws.  The scratch of the workspace path lives in a file-scope
 * static __device__ array of DROID_FRAME_DISTANCE_NOALLOC_MAX_PIXELS * 16 + 256 bytes (257 KB, held by every GPU
 * that loads the library until exit); ht * wd beyond that maximum returns DROID_ERR_STATIC_SCRATCH_TOO_SMALL.  The
 * call performs no cudaMalloc / cudaFree and calls no allocating library routine; the caller passes inputs and
 * outputs as for droid_frame_distance.  All calls on a device SHARE the array (it keeps the last call's contents;
 * every call rewrites all of what it reads, so nothing needs clearing): calls that overlap on the GPU -- different
 * streams, or graphs replayed concurrently -- race on it and must be serialised by the caller. */
#define DROID_FRAME_DISTANCE_NOALLOC_MAX_PIXELS 16384
int droid_frame_distance_noalloc_max_pixels(void);
int droid_frame_distance_noalloc(const float* poses, const float* disps, const float* intrinsics,
                                 const int64_t* ii, const int64_t* jj, float* dist, int n, int ht, int wd,
                                 float beta, void* stream);

/* projmap (src/droid.cpp:139-154, src/droid_kernels.cu:427-516,1463-1488).
 *   coords [n,ht,wd,3] (channel 2 = 0), valid [n,ht,wd,1]; both fully written. */
int droid_projmap(const float* poses, const float* disps, const float* intrinsics,
                  const int64_t* ii, const int64_t* jj, float* coords, float* valid, int n, int ht,
                  int wd, void* stream);

/* reproject -- extension, SURVEY 8f N1: DepthVideo.reproject (droid_slam/depth_video.py:139-147) =
 * geom/projective_ops.py:96-125 with jacobian=False, i.e. the PYTHON semantics (not projmap's): intrinsics
 * [nbuf,4] per frame, Z < 0.1 -> 1, valid = Z > 0.2, stereo edges ii == jj use the baseline (-0.1,0,0).
 *   coords [n,ht,wd,2], valid [n,ht,wd,1]; both fully written.
 *   target [n,ht,wd,2] and motn [n,4,ht,wd] are both NULL or both given: motn = the motion features of
 *   FactorGraph.update (droid_slam/factor_graph.py:203-204), clamp(cat(coords - grid, target - coords), -64, 64). */
int droid_reproject(const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                    const int64_t* jj, const float* target, float* coords, float* valid, float* motn,
                    int n, int ht, int wd, void* stream);

/* iproj (src/droid.cpp:157-166, src/droid_kernels.cu:779-850,1518-1541). points [nm,ht,wd,3]. */
int droid_iproj(const float* poses, const float* disps, const float* intrinsics, float* points,
                int nm, int ht, int wd, void* stream);

/* depth_filter (src/droid.cpp:220-234, src/droid_kernels.cu:661-775,1491-1515).
 *   disps [nbuf,ht,wd]; ix [num] int64; thresh [num]; counter [num,ht,wd] fully written (0..6). */
int droid_depth_filter(const float* poses, const float* disps, const float* intrinsics,
                       const int64_t* ix, const float* thresh, float* counter, int num, int nbuf,
                       int ht, int wd, void* stream);

/* ---- edge selection (extension) ------------------------------------------------------
 * The part of FactorGraph.add_proximity_factors (droid_slam/factor_graph.py:305-369) behind the distance
 * computation, on the device and stream-ordered.
 *   d [(t-t0)*(t-t1)]: video.distance over meshgrid(arange(t0,t), arange(t1,t)), row-major (not modified);
 *   ii_exist/jj_exist [n_exist]: cat(active, bad, inactive) edges of the graph (:322-323);
 *   out_ii/out_jj [capacity]: the edge list handed to add_factors, in the reference's order (mandatory temporal /
 *   stereo edges, then the accepted pairs in both directions by ascending distance; ties in index order);
 *   out_count (device int): number of entries written.  capacity >= droid_proximity_mandatory_count(...) and
 *   >= max_factors + 3 never overflows.  Status word = first int of the workspace: bit0 = an index the Python code
 *   would have raised IndexError on, bit1 = capacity exhausted.
 *   workspace: droid_proximity_workspace_bytes((t-t0)*(t-t1)) bytes of device memory. */
size_t droid_proximity_workspace_bytes(int n_pairs);
int droid_proximity_mandatory_count(int t, int t0, int rad, int stereo);
int droid_proximity_edges(const float* d, int t, int t0, int t1, int rad, int nms, float thresh, int max_factors,
                          int stereo, const int64_t* ii_exist, const int64_t* jj_exist, int n_exist,
                          int64_t* out_ii, int64_t* out_jj, int capacity, int* out_count, void* workspace,
                          size_t workspace_bytes, void* stream);

/* FactorGraph.__filter_repeated_edges (droid_slam/factor_graph.py:43-54), the second half of SURVEY 8f N3: of the
 * candidate edges (ii[k], jj[k]), k < n, keep those that are not among the graph's existing edges
 * (ii_exist/jj_exist = cat(active, inactive)); duplicates inside the candidate list are all kept, like the reference.
 *   out_ii/out_jj [n]: the kept edges in their original order; keep [n] (may be NULL): 1 = kept; out_count (device
 *   int): number kept.  Indices must be non-negative and < 2^32.
 *   workspace: droid_filter_edges_workspace_bytes(n_exist) bytes (an open-addressing hash table of the existing pairs). */
size_t droid_filter_edges_workspace_bytes(int n_exist);
int droid_filter_repeated_edges(const int64_t* ii, const int64_t* jj, int n, const int64_t* ii_exist,
                                const int64_t* jj_exist, int n_exist, int64_t* out_ii, int64_t* out_jj,
                                unsigned char* keep, int* out_count, void* workspace, size_t workspace_bytes,
                                void* stream);

/* ---- convex upsampling (extension) -----------------------------------------------------
 * cvx_upsample (droid_slam/droid_net.py:21-35), as DepthVideo.upsample applies it (depth_video.py:123-127).
 *   data [batch,ht,wd,dim] float32; mask [batch,9*8*8,ht,wd] logits, float32 (mask_dtype 0) or float16 (1);
 *   out [batch,8*ht,8*wd,dim] float32, fully written. */
int droid_cvx_upsample(const float* data, const void* mask, float* out, int mask_dtype, int batch, int ht, int wd,
                       int dim, void* stream);

/* ---- segmented mean over edges (extension) -----------------------------------------------
 * scatter_mean(net, ix, dim=1) of GraphAgg.forward (droid_slam/droid_net.py:59-75; torch_scatter): out[m] = mean of
 * src[e] over the edges with ix[e] == m.  src [n,D], out [M,D] (every row written; 0 for an empty group), float32
 * (dtype 0, D % 4 == 0) or float16 (dtype 1, D % 8 == 0), 16-byte aligned; accumulation in float32, ascending edge
 * order.  Status word = first int of the workspace: bit0 = an index outside [0, M). */
size_t droid_scatter_mean_workspace_bytes(int n, int M);
int droid_scatter_mean(const void* src, const int64_t* ix, void* out, int dtype, int n, int M, long long D,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ---- dense bundle adjustment -------------------------------------------------------------------
 * Replaces ba (src/droid.cpp:88-117, src/droid_kernels.cu:1314-1434).
 *   poses [nbuf_p,7] and disps [nbuf,ht,wd] are UPDATED IN PLACE; disps_sens [nbuf,ht,wd];
 *   targets, weights [n_edges,2,ht,wd]; eta [n_eta,ht,wd] with one row per entry of
 *   kx = unique(ii U [t0,t1)) (so n_eta must equal |kx|; checked on the device, see status);
 *   dx_out [t1-t0,6], dz_out [n_eta,ht*wd] = updates of the LAST iteration (dz untouched when
 *   motion_only; eta / n_eta are ignored then, as in the reference).  Graph bookkeeping (unique/CSR) runs on
 *   the device; host synchronisation only as stated at the top of this file.
 *   flags & DROID_BA_FLAG_CHECK: the call returns DROID_ERR_BA_INPUT when the device-side validation found
 *   n_eta != |kx| (the reference raises on the broadcast, src/droid_kernels.cu:1398) or an edge index outside
 *   [0, nbuf) (undefined behaviour in the reference; such edges are otherwise dropped).
 *   Multi-GPU: when `reduce_fn` != NULL it is called once per iteration, stream-ordered,
 *   on the packed reduced system (n_doubles fp64 values) after local accumulation; the Python host
 *   plugs torch.distributed.all_reduce (NCCL) in there.  With frame_lo/frame_hi a rank only
 *   accumulates source frames kx in [frame_lo, frame_hi).  "All frames" is frame_hi == INT32_MAX or < 0 (or
 *   the zero-initialised pair (0, 0) when reduce_fn == NULL); any other pair is taken literally, so
 *   frame_lo == frame_hi is an EMPTY shard: the rank contributes nothing and only joins the reduction.
 *   The buffer handed to reduce_fn is the reduced system PACKED to its structural non-zeros (the row segments
 *   inside the envelope of the pose order + the right-hand side; identical layout on every rank because the order
 *   is a function of the edge list alone) -- 10-20x fewer bytes than the dense lower triangle at 1000 keyframes.
 *   edge_row (device int[n_edges], or NULL = identity): row of edge e inside `targets` / `weights`, for a rank
 *   that holds only the rows of the edges whose source frame lies in its shard (entries of other edges are
 *   never read). */
typedef int (*droid_reduce_fn)(void* user, double* buf, size_t n_doubles, void* stream);

typedef struct {
  int n_edges, nbuf, ht, wd, n_eta, t0, t1, iterations, motion_only;
  float lm, ep;
  int frame_lo, frame_hi;      /* shard of source frames handled by this rank              */
  droid_reduce_fn reduce_fn;   /* NULL on a single GPU                                      */
  void* reduce_user;
  const int* edge_row;         /* NULL, or row of every edge inside targets / weights       */
  int flags;                   /* DROID_BA_FLAG_*                                            */
} droid_ba_params;
#define DROID_BA_FLAG_CHECK 1  /* read the status word back (one small D2H + wait) and fail on bad input */

size_t droid_ba_workspace_bytes(const droid_ba_params* p);
int droid_ba(float* poses, float* disps, const float* intrinsics, const float* disps_sens,
             const float* targets, const float* weights, const float* eta, const int64_t* ii,
             const int64_t* jj, const droid_ba_params* p, float* dx_out, float* dz_out,
             void* workspace, size_t workspace_bytes, void* stream);

/* Device-side status word of the last droid_ba on this workspace (first 4 bytes of the workspace):
 * bit0 = n_eta != |kx|, bit1 = index out of range, bit2 = Cholesky failed in some iteration
 * (=> dx = 0 for that iteration, like src/droid_kernels.cu:1207-1210).  Word 3 = largest out-degree of a source
 * frame.  droid_ba reads it only under DROID_BA_FLAG_CHECK (or where it has to wait for the chain layout anyway). */
#define DROID_BA_STATUS_ETA_MISMATCH 1
#define DROID_BA_STATUS_BAD_INDEX 2
#define DROID_BA_STATUS_CHOL_FAILED 4

/* Dense SPD solve used by ba, exported for tests/bench: A [n,n] fp64 row-major (lower triangle
 * read; contents unspecified afterwards: untouched for n <= 156, L for larger n), b [n] fp64
 * (must be A + n*n; overwritten by x); info (device int) != 0 on failure.
 * workspace: droid_chol_workspace_bytes(n) bytes of device memory (may be NULL for n <= 156). */
int droid_chol_solve_f64(double* A, double* b, int n, int* info_dev, void* workspace,
                         size_t workspace_bytes, void* stream);

/* Same solve for a matrix with a known row envelope: first_dev[r] (device, int[n]) = first structurally
 * non-zero column of row r (entries left of it must be exactly 0).  Fill-in stays inside the envelope, so
 * all tiles outside it are skipped; first_dev == NULL is the dense solve.  Replaces the sparsity that the
 * reference gets from Eigen::SimplicialLLT on a SparseMatrix (src/droid_kernels.cu:1192-1201).
 * workspace: droid_chol_workspace_bytes(n) bytes. */
int droid_chol_solve_env_f64(double* A, double* b, int n, const int* first_dev, int* info_dev,
                             void* workspace, size_t workspace_bytes, void* stream);
size_t droid_chol_workspace_bytes(int n);

/* Same solve for a matrix in "chains + separator" form (the nested-dissection shape ba gives the reduced camera
 * system of long trajectories): the columns split into nchains ranges [chain_end[c-1], chain_end[c]) (chain_end is a
 * HOST array of even, increasing values, chain_end[-1] = 0) that are structurally decoupled from each other, followed
 * by the separator [chain_end[nchains-1], n).  The chains are factorised concurrently (one pair of streams each),
 * then the separator.  firstc_dev (device, int[(nchains+1) * n]): firstc[c*n + r] = first non-zero column of row r
 * inside range c (range nchains = the separator); a value outside the range, or > r, means "none".  nchains == 0 is
 * droid_chol_solve_env_f64 (firstc_dev may then be NULL = dense).  n must be even, A 16-byte aligned, n <= 12288.
 * workspace: droid_chol_chains_workspace_bytes(n, nchains, n - chain_end[nchains-1]) bytes. */
int droid_chol_solve_chains_f64(double* A, double* b, int n, const int* firstc_dev, int nchains,
                                const int* chain_end, int* info_dev, void* workspace, size_t workspace_bytes,
                                void* stream);
size_t droid_chol_chains_workspace_bytes(int n, int nchains, int sep_rows);

/* The general form: an ELIMINATION TREE of column ranges (nested dissection).  The columns split into nranges ranges
 * [range_end[r-1], range_end[r]) (host array, even increasing values, range_end[nranges-1] = n), laid out level by
 * level: level l holds the ranges level_first[l] .. level_first[l+1]-1 (host array of nlevels + 1 entries,
 * level_first[0] = 0, level_first[nlevels] = nranges, at most 8 ranges per level).  Ranges of one level are
 * structurally decoupled from each other and are factorised CONCURRENTLY; a range is coupled only to ranges of higher
 * levels, and two ranges of one level whose positions in the level have the same parity must not be coupled to the same
 * higher-level range (true for the separators of a path graph in path order: neighbours alternate).  The levels are
 * factorised one after the other; the backward sweep walks them in reverse.  firstc_dev as above, one row of n entries
 * per range.  Chains + one separator is the two-level case; the reduced camera system of a long trajectory uses
 * log2(chains) + 1 levels, which shortens the serial pivot chain from (chain + all separators) to
 * (chain + one separator per level).
 * workspace: droid_chol_tree_workspace_bytes(n, nranges, n - range_end[level_first[1] - 1]) bytes. */
int droid_chol_solve_tree_f64(double* A, double* b, int n, const int* firstc_dev, int nranges, const int* range_end,
                              int nlevels, const int* level_first, int* info_dev, void* workspace,
                              size_t workspace_bytes, void* stream);
size_t droid_chol_tree_workspace_bytes(int n, int nranges, int sep_rows);

#ifdef __cplusplus
}
#endif
#endif /* DROID_B200_H_ */
