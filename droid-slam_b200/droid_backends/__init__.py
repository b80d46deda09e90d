"""Drop-in replacement of DROID-SLAM's ``droid_backends`` extension module for NVIDIA B200 (sm_100a).

Same nine functions, positional signatures and return structures as the reference's pybind module
(/root/reference/src/droid.cpp:237-250), so ``droid_slam/modules/corr.py``, ``depth_video.py``,
``factor_graph.py`` and ``visualization.py`` can ``import droid_backends`` unchanged.

This file is only the host side: it validates arguments, allocates outputs / workspaces with torch on
the current CUDA stream and calls the C ABI of ``libdroid_b200.so`` (include/droid_b200.h) through
ctypes.  There is NO CPU or PyTorch fallback: if the shared library is missing or a tensor is not on
a CUDA device the call raises.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "..", "lib", "libdroid_b200.so")

_F32, _F16, _F64 = 0, 1, 2
_DT = {torch.float32: _F32, torch.float16: _F16, torch.float64: _F64}

_ERRORS = {
    -1: "invalid argument", -2: "unsupported dtype", -3: "workspace too small",
    -4: "CUDA launch failed", -5: "not implemented", -6: "device-side input validation failed",
    -7: "size exceeds the built-in scratch of a no-alloc entry point",
}

REDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                             ctypes.c_void_p)


class BaParams(ctypes.Structure):
    """Mirror of droid_ba_params (include/droid_b200.h)."""
    _fields_ = [
        ("n_edges", ctypes.c_int), ("nbuf", ctypes.c_int), ("ht", ctypes.c_int), ("wd", ctypes.c_int),
        ("n_eta", ctypes.c_int), ("t0", ctypes.c_int), ("t1", ctypes.c_int),
        ("iterations", ctypes.c_int), ("motion_only", ctypes.c_int),
        ("lm", ctypes.c_float), ("ep", ctypes.c_float),
        ("frame_lo", ctypes.c_int), ("frame_hi", ctypes.c_int),
        ("reduce_fn", REDUCE_FN), ("reduce_user", ctypes.c_void_p), ("edge_row", ctypes.c_void_p),
        ("flags", ctypes.c_int),
    ]


BA_FLAG_CHECK = 1
# Input validation of ba() (eta rows == |unique(ii U [t0,t1))|, edge indices inside the buffer) happens on the device;
# by default the host reads the verdict back and raises like the reference does on the eta broadcast
# (src/droid_kernels.cu:1398).  DROID_BA_CHECK=0 (or ba(..., _validate=False)) keeps frontend-sized calls free of the
# one host wait this costs; large systems wait for their chain layout anyway, so the check is free there.
_BA_CHECK_DEFAULT = os.environ.get("DROID_BA_CHECK", "1") != "0"


def _load():
    path = os.environ.get("DROID_B200_LIB", _LIB_PATH)
    if not os.path.exists(path):
        raise ImportError(
            "libdroid_b200.so not found at %s -- build it with `python droid-slam_b200/build.py` "
            "(there is no CPU / PyTorch fallback)" % os.path.abspath(path))
    lib = ctypes.CDLL(os.path.abspath(path))
    lib.droid_version.restype = ctypes.c_char_p
    lib.droid_last_cuda_error.restype = ctypes.c_char_p
    lib.droid_ba_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_ba_workspace_bytes.argtypes = [ctypes.POINTER(BaParams)]
    lib.droid_chol_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_chol_chains_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_chol_tree_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_corr_build_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_launch_count.restype = ctypes.c_ulonglong
    lib.droid_proximity_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_scatter_mean_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_filter_edges_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_frame_distance_workspace_bytes.restype = ctypes.c_size_t
    lib.droid_altcorr_forward_workspace_bytes.restype = ctypes.c_size_t
    return lib


_lib = _load()
__version__ = _lib.droid_version().decode()

# names every build of the library must export (checked by tests/test_abi.py against the header)
C_ABI_SYMBOLS = [
    "droid_version", "droid_last_cuda_error", "droid_launch_count", "droid_prof_enable", "droid_prof_read", "droid_corr_index_forward", "droid_corr_index_backward",
    "droid_corr_index_pyramid", "droid_altcorr_forward", "droid_altcorr_forward_ws",
    "droid_altcorr_forward_workspace_bytes", "droid_altcorr_backward", "droid_altcorr_pyramid_indexed",
    "droid_altcorr_forward_indexed", "droid_corr_build", "droid_corr_build_workspace_bytes",
    "droid_frame_distance", "droid_frame_distance_workspace_bytes", "droid_frame_distance_ws",
    "droid_frame_distance_noalloc_max_pixels", "droid_frame_distance_noalloc", "droid_projmap", "droid_reproject", "droid_iproj", "droid_depth_filter",
    "droid_ba_workspace_bytes", "droid_ba", "droid_chol_solve_f64", "droid_chol_solve_env_f64",
    "droid_chol_workspace_bytes", "droid_chol_solve_chains_f64", "droid_chol_chains_workspace_bytes",
    "droid_chol_solve_tree_f64", "droid_chol_tree_workspace_bytes",
    "droid_proximity_workspace_bytes", "droid_proximity_mandatory_count", "droid_proximity_edges",
    "droid_filter_edges_workspace_bytes", "droid_filter_repeated_edges",
    "droid_cvx_upsample", "droid_scatter_mean_workspace_bytes", "droid_scatter_mean",
]


PROF_PHASES = ["ba_prep", "ba_accum", "ba_schur", "ba_reduce", "ba_solve", "ba_update", "corr_gemm", "spare",
               "chol_panel", "chol_syrk", "chol_bwd", "spare2"]


def launch_count():
    """Kernels launched by libdroid_b200 since load (bench.py's gpu_launches)."""
    return int(_lib.droid_launch_count())


def prof_enable(on=True):
    _lib.droid_prof_enable(1 if on else 0)


def prof_read():
    """{phase: (milliseconds, intervals)} of the phases recorded since the last read (synchronises)."""
    n = len(PROF_PHASES)
    ms = (ctypes.c_double * n)()
    cnt = (ctypes.c_int * n)()
    _check(_lib.droid_prof_read(ms, cnt, n), "prof_read")
    return {PROF_PHASES[i]: (ms[i], cnt[i]) for i in range(n)}


def _check(rc, what):
    if rc != 0:
        msg = _ERRORS.get(rc, "error %d" % rc)
        if rc == -4:
            msg += ": " + _lib.droid_last_cuda_error().decode()
        raise RuntimeError("droid_backends.%s: %s" % (what, msg))


def _contig(what, **tensors):
    # reference: CHECK_CONTIGUOUS -> TORCH_CHECK -> RuntimeError (src/droid.cpp:84-85)
    for name, t in tensors.items():
        if not isinstance(t, torch.Tensor):
            raise RuntimeError("droid_backends.%s: %s must be a tensor" % (what, name))
        if not t.is_contiguous():
            raise RuntimeError("%s must be contiguous" % name)
        if not t.is_cuda:
            raise RuntimeError("droid_backends.%s: %s must be a CUDA tensor (no CPU fallback)" % (what, name))


def _dtype(what, t, allowed):
    if t.dtype not in allowed:
        raise RuntimeError("droid_backends.%s: expected dtype in %s, got %s" % (what, allowed, t.dtype))


def _p(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None and t.numel() > 0 else 0)


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _long(what, **tensors):
    for name, t in tensors.items():
        if t.dtype != torch.int64:
            raise RuntimeError("droid_backends.%s: %s must be int64" % (what, name))


# ------------------------------------------------------------------------------------------------
# correlation lookup                                         reference: src/droid.cpp:170-191
# ------------------------------------------------------------------------------------------------
def corr_index_forward(volume, coords, radius):
    _contig("corr_index_forward", volume=volume, coords=coords)
    _dtype("corr_index_forward", volume, (torch.float32, torch.float16, torch.float64))
    _dtype("corr_index_forward", coords, (torch.float32,))
    if volume.dim() != 5 or coords.dim() != 4:
        raise RuntimeError("droid_backends.corr_index_forward: volume must be 5-d, coords 4-d")
    B, H1, W1, H2, W2 = volume.shape
    rd = 2 * radius + 1
    with torch.cuda.device(volume.device):
        out = torch.empty((B, rd, rd, H1, W1), dtype=volume.dtype, device=volume.device)
        _check(_lib.droid_corr_index_forward(_p(volume), _p(coords), _p(out), _DT[volume.dtype], B,
                                             H1, W1, H2, W2, int(radius), _stream(volume)),
               "corr_index_forward")
    return [out]


def corr_index_backward(volume, coords, corr_grad, radius):
    _contig("corr_index_backward", volume=volume, coords=coords, corr_grad=corr_grad)
    _dtype("corr_index_backward", volume, (torch.float32, torch.float16, torch.float64))
    if corr_grad.dtype != volume.dtype:
        raise RuntimeError("droid_backends.corr_index_backward: corr_grad dtype must match volume")
    B, H1, W1, H2, W2 = volume.shape
    with torch.cuda.device(volume.device):
        vg = torch.empty_like(volume)
        _check(_lib.droid_corr_index_backward(_p(coords), _p(corr_grad), _p(vg), _DT[volume.dtype], B,
                                              H1, W1, H2, W2, int(radius), _stream(volume)),
               "corr_index_backward")
    return [vg]


def corr_index_pyramid(volumes, coords, radius=3):
    """Extension: all pyramid levels in one launch == torch.cat of the per-level lookups, dim 1.
    volumes[l]: [B,H1,W1,H2>>l,W2>>l]; coords [B,2,H1,W1] at level 0 (modules/corr.py:40-50)."""
    L = len(volumes)
    _contig("corr_index_pyramid", coords=coords, **{"volume%d" % i: v for i, v in enumerate(volumes)})
    v0 = volumes[0]
    _dtype("corr_index_pyramid", v0, (torch.float32, torch.float16))
    B, H1, W1, H2, W2 = v0.shape
    for l, v in enumerate(volumes):
        if v.dtype != v0.dtype or tuple(v.shape) != (B, H1, W1, H2 >> l, W2 >> l):
            raise RuntimeError("droid_backends.corr_index_pyramid: bad level %d" % l)
    rd = 2 * radius + 1
    ptrs = (ctypes.c_void_p * L)(*[v.data_ptr() for v in volumes])
    with torch.cuda.device(v0.device):
        out = torch.empty((B, L * rd * rd, H1, W1), dtype=v0.dtype, device=v0.device)
        _check(_lib.droid_corr_index_pyramid(ptrs, L, _p(coords), _p(out), _DT[v0.dtype], B, H1, W1,
                                             H2, W2, int(radius), _stream(v0)), "corr_index_pyramid")
    return out


# ------------------------------------------------------------------------------------------------
# alt correlation                                            reference: src/droid.cpp:193-217
# ------------------------------------------------------------------------------------------------
def altcorr_forward(fmap1, fmap2, coords, radius, _plain=False):
    """``_plain`` (tests): the entry point without a workspace, i.e. always the CUDA-core kernel."""
    _contig("altcorr_forward", fmap1=fmap1, fmap2=fmap2, coords=coords)
    _dtype("altcorr_forward", fmap1, (torch.float32,))
    _dtype("altcorr_forward", fmap2, (torch.float32,))
    _dtype("altcorr_forward", coords, (torch.float32,))
    B, N, H, W, _ = coords.shape
    _, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    rd = 2 * radius + 1
    with torch.cuda.device(fmap1.device):
        out = torch.empty((B, N, rd * rd, H, W), dtype=fmap1.dtype, device=fmap1.device)
        # tensor-core path (altcorr_tc.cu): fp16 copies of the maps + an exactness flag live in a workspace; the
        # CUDA-core kernel behind it takes over on the device when the maps are not fp16 values
        nws = 0 if _plain else int(_lib.droid_altcorr_forward_workspace_bytes(B, N, H1, W1, H2, W2, C, int(radius)))
        ws = _workspace(fmap1.device, nws, "altcorr") if nws else None
        _check(_lib.droid_altcorr_forward_ws(_p(fmap1), _p(fmap2), _p(coords), _p(out), _F32, B, N, H1,
                                             W1, H2, W2, C, int(radius), _p(ws) if nws else None,
                                             ctypes.c_size_t(ws.numel() if nws else 0), _stream(fmap1)),
               "altcorr_forward")
    return [out]


def altcorr_backward(fmap1, fmap2, coords, corr_grad, radius):
    _contig("altcorr_backward", fmap1=fmap1, fmap2=fmap2, coords=coords, corr_grad=corr_grad)
    for t in (fmap1, fmap2, coords, corr_grad):
        _dtype("altcorr_backward", t, (torch.float32,))
    B, N = coords.shape[0], coords.shape[1]
    _, H1, W1, C = fmap1.shape
    _, H2, W2, _ = fmap2.shape
    with torch.cuda.device(fmap1.device):
        g1 = torch.empty_like(fmap1)
        g2 = torch.empty_like(fmap2)
        gc = torch.empty((B, N, H1, W1, 2), dtype=fmap1.dtype, device=fmap1.device)
        _check(_lib.droid_altcorr_backward(_p(fmap1), _p(fmap2), _p(coords), _p(corr_grad), _p(g1),
                                           _p(g2), _p(gc), B, N, H1, W1, H2, W2, C, int(radius),
                                           _stream(fmap1)), "altcorr_backward")
    return [g1, g2, gc]


def altcorr_forward_indexed(pyr0, pyrl, ii, jj, coords, radius=3):
    """Extension: altcorr_forward(pyr0[ii].float(), pyrl[jj].float(), coords, r) without
    materialising the gathered fp32 copies (modules/corr.py:113-120).  pyr*: [F,H,W,C] f16/f32."""
    _contig("altcorr_forward_indexed", pyr0=pyr0, pyrl=pyrl, ii=ii, jj=jj, coords=coords)
    _long("altcorr_forward_indexed", ii=ii, jj=jj)
    _dtype("altcorr_forward_indexed", pyr0, (torch.float32, torch.float16))
    if pyrl.dtype != pyr0.dtype:
        raise RuntimeError("droid_backends.altcorr_forward_indexed: dtype mismatch")
    B, N, H, W, _ = coords.shape
    F, H1, W1, C = pyr0.shape
    _, H2, W2, _ = pyrl.shape
    rd = 2 * radius + 1
    with torch.cuda.device(pyr0.device):
        out = torch.empty((B, N, rd * rd, H, W), dtype=torch.float32, device=pyr0.device)
        _check(_lib.droid_altcorr_forward_indexed(_p(pyr0), _p(pyrl), _p(ii), _p(jj), _p(coords),
                                                  _p(out), _DT[pyr0.dtype], F, B, N, H1, W1, H2, W2,
                                                  C, int(radius), _stream(pyr0)),
               "altcorr_forward_indexed")
    return out


def altcorr_pyramid_indexed(pyramid, ii, jj, coords, radius=3):
    """Extension: every level of an AltCorrBlock in ONE launch == torch.cat([altcorr_forward(pyramid[0][ii].float(),
    pyramid[l][jj].float(), coords / 2**l, r) for l], dim=2) (modules/corr.py:106-125).  pyramid[l]: [F, H>>l, W>>l, C]
    float16, channel-last; coords [B, S, H, W, 2] at level 0; returns [B, S, L*(2r+1)^2, H, W] float32.  Shapes the
    tensor-core kernel does not cover fall back to one altcorr_forward_indexed call per level."""
    L = len(pyramid)
    _contig("altcorr_pyramid_indexed", ii=ii, jj=jj, coords=coords, **{"level%d" % i: v for i, v in enumerate(pyramid)})
    _long("altcorr_pyramid_indexed", ii=ii, jj=jj)
    p0 = pyramid[0]
    _dtype("altcorr_pyramid_indexed", p0, (torch.float32, torch.float16))
    F, H1, W1, C = p0.shape
    B, N, H, W, _ = coords.shape
    for l, v in enumerate(pyramid):
        if v.dtype != p0.dtype or tuple(v.shape) != (F, H1 >> l, W1 >> l, C):
            raise RuntimeError("droid_backends.altcorr_pyramid_indexed: bad level %d" % l)
    rd = 2 * radius + 1
    with torch.cuda.device(p0.device):
        if p0.dtype == torch.float16 and 1 <= L <= 4:
            out = torch.empty((B, N, L * rd * rd, H, W), dtype=torch.float32, device=p0.device)
            ptrs = (ctypes.c_void_p * L)(*[v.data_ptr() for v in pyramid])
            rc = _lib.droid_altcorr_pyramid_indexed(ptrs, L, _p(ii), _p(jj), _p(coords), _p(out), _DT[p0.dtype], F, B,
                                                    N, H1, W1, C, int(radius), _stream(p0))
            if rc == 0:
                return out
            if rc != -5:                                   # DROID_ERR_NOT_IMPLEMENTED: shape not covered
                _check(rc, "altcorr_pyramid_indexed")
        return torch.cat([altcorr_forward_indexed(p0, pyramid[l], ii, jj, (coords / 2 ** l).contiguous(), radius)
                          for l in range(L)], dim=2)


def corr_build(fmap1, fmap2, num_levels=4):
    """Extension: CorrBlock.corr + avg-pool pyramid (modules/corr.py:24-38,63-71) as one tcgen05 GEMM
    with a pooling epilogue.  fmap1, fmap2: [B, C, H, W] (or [1,B,C,H,W]) f16, unscaled.
    Returns the 4 pyramid levels [B, H, W, H>>l, W>>l] f16."""
    if fmap1.dim() == 5:
        fmap1 = fmap1.reshape((-1,) + tuple(fmap1.shape[2:]))
        fmap2 = fmap2.reshape((-1,) + tuple(fmap2.shape[2:]))
    _contig("corr_build", fmap1=fmap1, fmap2=fmap2)
    _dtype("corr_build", fmap1, (torch.float16,))
    _dtype("corr_build", fmap2, (torch.float16,))
    if num_levels != 4:
        raise RuntimeError("droid_backends.corr_build: num_levels must be 4")
    B, C, H, W = fmap1.shape
    with torch.cuda.device(fmap1.device):
        lv = [torch.empty((B, H, W, H >> l, W >> l), dtype=torch.float16, device=fmap1.device)
              for l in range(4)]
        nws = _lib.droid_corr_build_workspace_bytes(B, C, H, W)
        ws = torch.empty((nws,), dtype=torch.uint8, device=fmap1.device)
        _check(_lib.droid_corr_build(_p(fmap1), _p(fmap2), _p(lv[0]), _p(lv[1]), _p(lv[2]), _p(lv[3]),
                                     B, C, H, W, _p(ws), ctypes.c_size_t(nws), _stream(fmap1)),
               "corr_build")
    return lv


# ------------------------------------------------------------------------------------------------
# geometry                                                   reference: src/droid.cpp:120-166,220-234
# ------------------------------------------------------------------------------------------------
def frame_distance(poses, disps, intrinsics, ii, jj, beta, _plain=False):
    """``_plain=True`` calls the workspace-free C entry point (droid_frame_distance); same bits, slower."""
    _contig("frame_distance", poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _long("frame_distance", ii=ii, jj=jj)
    for t in (poses, disps, intrinsics):
        _dtype("frame_distance", t, (torch.float32,))
    n = ii.shape[0]
    ht, wd = disps.shape[1], disps.shape[2]
    with torch.cuda.device(poses.device):
        dist = torch.empty((n,), dtype=torch.float32, device=poses.device)
        if _plain or n <= 256:      # few pairs (the frontend's keyframe test): one launch, no per-call table, same bits
            _check(_lib.droid_frame_distance(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj), _p(dist), n, ht, wd,
                                             ctypes.c_float(beta), _stream(poses)), "frame_distance")
            return dist
        nws = int(_lib.droid_frame_distance_workspace_bytes(ht, wd))
        ws = _workspace(poses.device, nws, "frame_distance")
        _check(_lib.droid_frame_distance_ws(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj), _p(dist), n, ht, wd,
                                            ctypes.c_float(beta), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(nws),
                                            _stream(poses)), "frame_distance")
    return dist


def projmap(poses, disps, intrinsics, ii, jj):
    _contig("projmap", poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _long("projmap", ii=ii, jj=jj)
    for t in (poses, disps, intrinsics):
        _dtype("projmap", t, (torch.float32,))
    n = ii.shape[0]
    ht, wd = disps.shape[1], disps.shape[2]
    with torch.cuda.device(poses.device):
        coords = torch.empty((n, ht, wd, 3), dtype=torch.float32, device=poses.device)
        valid = torch.empty((n, ht, wd, 1), dtype=torch.float32, device=poses.device)
        _check(_lib.droid_projmap(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj), _p(coords),
                                  _p(valid), n, ht, wd, _stream(poses)), "projmap")
    return [coords, valid]


def reproject(poses, disps, intrinsics, ii, jj, target=None):
    """Extension: what ``DepthVideo.reproject(ii, jj)`` computes (droid_slam/depth_video.py:139-147,
    geom/projective_ops.py:96-125) in one launch.  intrinsics [nbuf,4] per frame (like ``video.intrinsics``).
    Returns [coords [n,ht,wd,2], valid [n,ht,wd,1]]; with ``target`` [n,ht,wd,2] also the motion features of
    ``FactorGraph.update`` (droid_slam/factor_graph.py:203-204) as a third tensor [n,4,ht,wd]."""
    _contig("reproject", poses=poses, disps=disps, intrinsics=intrinsics, ii=ii, jj=jj)
    _long("reproject", ii=ii, jj=jj)
    for t in (poses, disps, intrinsics):
        _dtype("reproject", t, (torch.float32,))
    if intrinsics.dim() != 2 or intrinsics.shape[1] != 4 or intrinsics.shape[0] < disps.shape[0]:
        raise RuntimeError("droid_backends.reproject: intrinsics must be [nbuf,4] (one row per frame)")
    n = ii.shape[0]
    ht, wd = disps.shape[1], disps.shape[2]
    if target is not None:
        _contig("reproject", target=target)
        _dtype("reproject", target, (torch.float32,))
        if tuple(target.shape) != (n, ht, wd, 2):
            raise RuntimeError("droid_backends.reproject: target must be [n,ht,wd,2]")
    with torch.cuda.device(poses.device):
        coords = torch.empty((n, ht, wd, 2), dtype=torch.float32, device=poses.device)
        valid = torch.empty((n, ht, wd, 1), dtype=torch.float32, device=poses.device)
        motn = torch.empty((n, 4, ht, wd), dtype=torch.float32, device=poses.device) if target is not None else None
        _check(_lib.droid_reproject(_p(poses), _p(disps), _p(intrinsics), _p(ii), _p(jj),
                                    _p(target) if target is not None else None, _p(coords), _p(valid),
                                    _p(motn) if motn is not None else None, n, ht, wd, _stream(poses)), "reproject")
    return [coords, valid] if motn is None else [coords, valid, motn]


def proximity_edges(d, t, t0, t1, rad, nms, thresh, max_factors, stereo, ii_exist, jj_exist):
    """Extension: the edge selection of ``FactorGraph.add_proximity_factors``
    (droid_slam/factor_graph.py:305-369) on the device.  ``d`` = ``video.distance(ii, jj, beta)`` over
    ``meshgrid(arange(t0, t), arange(t1, t))`` (float32 CUDA, row-major, left untouched); ``ii_exist`` / ``jj_exist`` =
    ``cat(graph.ii, graph.ii_bad, graph.ii_inac)`` and likewise for jj.  Returns ``(ii, jj)`` int64 CUDA tensors: the
    edge list the reference hands to ``add_factors``, in the same order.  One D2H read (the edge count)."""
    if not d.is_cuda or d.dtype != torch.float32 or not d.is_contiguous():
        raise RuntimeError("droid_backends.proximity_edges: d must be a contiguous float32 CUDA tensor")
    t, t0, t1 = int(t), int(t0), int(t1)
    n = (t - t0) * (t - t1)
    if t0 < 0 or t1 < 0 or t < t0 or t < t1 or d.numel() != n:
        raise RuntimeError("droid_backends.proximity_edges: d must hold (t - t0) * (t - t1) distances")
    dev = d.device
    ii_exist = ii_exist.to(device=dev, dtype=torch.long).contiguous()
    jj_exist = jj_exist.to(device=dev, dtype=torch.long).contiguous()
    if ii_exist.numel() != jj_exist.numel():
        raise RuntimeError("droid_backends.proximity_edges: ii_exist / jj_exist differ in length")
    with torch.cuda.device(dev):
        mand = int(_lib.droid_proximity_mandatory_count(t, t0, int(rad), 1 if stereo else 0))
        cap = max(mand, int(max_factors) + 1) + 4
        out_ii = torch.empty((cap,), dtype=torch.long, device=dev)
        out_jj = torch.empty((cap,), dtype=torch.long, device=dev)
        count = torch.zeros((1,), dtype=torch.int32, device=dev)
        nws = int(_lib.droid_proximity_workspace_bytes(n))
        ws = torch.empty((nws,), dtype=torch.uint8, device=dev)
        _check(_lib.droid_proximity_edges(_p(d), t, t0, t1, int(rad), int(nms), ctypes.c_float(float(thresh)),
                                          int(max_factors), 1 if stereo else 0, _p(ii_exist), _p(jj_exist),
                                          int(ii_exist.numel()), _p(out_ii), _p(out_jj), cap, _p(count),
                                          ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(nws), _stream(d)),
               "proximity_edges")
        k = int(count.item())
        off = (-ws.data_ptr()) % 256
        status = int(ws[off:off + 4].view(torch.int32).item())
    if status & 1:
        raise IndexError("droid_backends.proximity_edges: index out of range (t0, t1, rad as the Python code would raise)")
    if status & 2:
        raise RuntimeError("droid_backends.proximity_edges: edge capacity exhausted")
    return out_ii[:k], out_jj[:k]


def filter_repeated_edges(ii, jj, ii_exist, jj_exist):
    """Extension: ``FactorGraph._FactorGraph__filter_repeated_edges`` (droid_slam/factor_graph.py:43-54)
    on the device.  ``ii_exist`` / ``jj_exist`` = ``cat(graph.ii, graph.ii_inac)`` and likewise for jj.  Returns the
    candidate edges that are not in the graph yet, in their original order (int64 CUDA).  One D2H read (the count)."""
    _contig("filter_repeated_edges", ii=ii, jj=jj)
    _long("filter_repeated_edges", ii=ii, jj=jj)
    dev = ii.device
    ii_exist = ii_exist.to(device=dev, dtype=torch.long).contiguous()
    jj_exist = jj_exist.to(device=dev, dtype=torch.long).contiguous()
    if ii.numel() != jj.numel() or ii_exist.numel() != jj_exist.numel():
        raise RuntimeError("droid_backends.filter_repeated_edges: ii / jj differ in length")
    n, ne = int(ii.numel()), int(ii_exist.numel())
    with torch.cuda.device(dev):
        out_ii, out_jj = torch.empty_like(ii), torch.empty_like(jj)
        count = torch.zeros((1,), dtype=torch.int32, device=dev)
        nws = int(_lib.droid_filter_edges_workspace_bytes(ne))
        ws = torch.empty((nws,), dtype=torch.uint8, device=dev)
        _check(_lib.droid_filter_repeated_edges(_p(ii), _p(jj), n, _p(ii_exist), _p(jj_exist), ne, _p(out_ii), _p(out_jj),
                                                None, _p(count), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(nws),
                                                _stream(ii)), "filter_repeated_edges")
        k = int(count.item())
    return out_ii[:k], out_jj[:k]


def cvx_upsample(data, mask):
    """Extension: ``cvx_upsample`` of droid_slam/droid_net.py:21-35 in one pass.  data [batch,ht,wd,dim]
    float32, mask [batch, 9*8*8, ht, wd] (any shape with that many elements; float32 or float16 logits) ->
    [batch, 8*ht, 8*wd, dim] float32."""
    _contig("cvx_upsample", data=data, mask=mask)
    _dtype("cvx_upsample", data, (torch.float32,))
    _dtype("cvx_upsample", mask, (torch.float32, torch.float16))
    if data.dim() != 4:
        raise RuntimeError("droid_backends.cvx_upsample: data must be [batch,ht,wd,dim]")
    batch, ht, wd, dim = data.shape
    if mask.numel() != batch * 576 * ht * wd or mask.device != data.device:
        raise RuntimeError("droid_backends.cvx_upsample: mask must hold batch * 9*8*8 * ht * wd logits on data's device")
    with torch.cuda.device(data.device):
        out = torch.empty((batch, 8 * ht, 8 * wd, dim), dtype=torch.float32, device=data.device)
        _check(_lib.droid_cvx_upsample(_p(data), _p(mask), _p(out), 0 if mask.dtype == torch.float32 else 1,
                                       batch, ht, wd, dim, _stream(data)), "cvx_upsample")
    return out


def scatter_mean(src, index, dim=1, dim_size=None):
    """Extension: ``torch_scatter.scatter_mean(src, index, dim=1)`` as ``GraphAgg.forward`` uses it
    (droid_slam/droid_net.py:67): src [1, n, ...] (batch 1, like everywhere in DROID-SLAM's inference) float32 or
    float16, index [n] int64 -> [1, dim_size or index.max() + 1, ...].  Accumulates in float32."""
    if dim != 1 or src.dim() < 2 or src.shape[0] != 1:
        raise RuntimeError("droid_backends.scatter_mean: only dim=1 of a [1, n, ...] tensor (GraphAgg's use)")
    _contig("scatter_mean", src=src, index=index)
    _long("scatter_mean", index=index)
    _dtype("scatter_mean", src, (torch.float32, torch.float16))
    n = src.shape[1]
    if index.numel() != n:
        raise RuntimeError("droid_backends.scatter_mean: index must have one entry per slice of dim 1")
    M = int(dim_size) if dim_size is not None else (int(index.max().item()) + 1 if n > 0 else 0)
    D = 1
    for s_ in src.shape[2:]:
        D *= int(s_)
    V = 4 if src.dtype == torch.float32 else 8
    if D % V != 0:
        raise RuntimeError("droid_backends.scatter_mean: trailing size must be a multiple of %d" % V)
    with torch.cuda.device(src.device):
        out = torch.empty((1, M) + tuple(src.shape[2:]), dtype=src.dtype, device=src.device)
        nws = int(_lib.droid_scatter_mean_workspace_bytes(n, M))
        ws = torch.empty((nws,), dtype=torch.uint8, device=src.device)
        _check(_lib.droid_scatter_mean(_p(src), _p(index), _p(out), 0 if src.dtype == torch.float32 else 1, n, M,
                                       ctypes.c_longlong(D), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(nws),
                                       _stream(src)), "scatter_mean")
    return out


def iproj(poses, disps, intrinsics):
    _contig("iproj", poses=poses, disps=disps, intrinsics=intrinsics)
    for t in (poses, disps, intrinsics):
        _dtype("iproj", t, (torch.float32,))
    nm, ht, wd = disps.shape
    if poses.shape[0] < nm:
        raise RuntimeError("droid_backends.iproj: fewer poses than depth maps")
    with torch.cuda.device(disps.device):
        points = torch.empty((nm, ht, wd, 3), dtype=torch.float32, device=disps.device)
        _check(_lib.droid_iproj(_p(poses), _p(disps), _p(intrinsics), _p(points), nm, ht, wd,
                                _stream(disps)), "iproj")
    return points


def depth_filter(poses, disps, intrinsics, ix, thresh):
    _contig("depth_filter", poses=poses, disps=disps, intrinsics=intrinsics, ix=ix, thresh=thresh)
    _long("depth_filter", ix=ix)
    for t in (poses, disps, intrinsics, thresh):
        _dtype("depth_filter", t, (torch.float32,))
    num = ix.shape[0]
    nbuf, ht, wd = disps.shape
    with torch.cuda.device(disps.device):
        counter = torch.empty((num, ht, wd), dtype=torch.float32, device=disps.device)
        _check(_lib.droid_depth_filter(_p(poses), _p(disps), _p(intrinsics), _p(ix), _p(thresh),
                                       _p(counter), num, nbuf, ht, wd, _stream(disps)),
               "depth_filter")
    return counter


# ------------------------------------------------------------------------------------------------
# dense bundle adjustment                                     reference: src/droid.cpp:88-117
# ------------------------------------------------------------------------------------------------
_ws_cache = {}


def _workspace(device, nbytes, tag="ba"):
    key = (device.type, device.index, tag)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty((int(nbytes * 1.25) + 1024,), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations,
       lm, ep, motion_only, _shard=None, _reduce=None, _edge_row=None, _validate=None):
    """Updates ``poses`` and ``disps`` in place; returns [dx, dz] of the last iteration
    (src/droid_kernels.cu:1314-1434).  ``_shard=(lo, hi)`` / ``_reduce`` / ``_edge_row`` are used by ba_sharded.
    Raises RuntimeError when ``eta`` does not have one row per optimised frame or an edge index lies outside the
    buffer (``_validate=False`` / DROID_BA_CHECK=0: no read-back, such edges are dropped, missing eta rows skipped)."""
    _contig("ba", poses=poses, disps=disps, intrinsics=intrinsics, disps_sens=disps_sens,
            targets=targets, weights=weights, ii=ii, jj=jj)
    if not isinstance(eta, torch.Tensor) or not eta.is_cuda:
        raise RuntimeError("droid_backends.ba: eta must be a CUDA tensor")
    if not eta.is_contiguous():  # the reference .view()s eta without checking (:1398); be explicit
        raise RuntimeError("eta must be contiguous")
    _long("ba", ii=ii, jj=jj)
    for t in (poses, disps, intrinsics, disps_sens, targets, weights, eta):
        _dtype("ba", t, (torch.float32,))
    nbuf, ht, wd = disps.shape
    n_edges = ii.shape[0]
    HW = ht * wd
    n_eta = eta.numel() // HW if HW > 0 else 0
    P = int(t1) - int(t0)
    prm = BaParams(n_edges=n_edges, nbuf=nbuf, ht=ht, wd=wd, n_eta=n_eta, t0=int(t0), t1=int(t1),
                   iterations=int(iterations), motion_only=1 if motion_only else 0, lm=float(lm),
                   ep=float(ep), frame_lo=0, frame_hi=0x7FFFFFFF, reduce_fn=REDUCE_FN(0),
                   reduce_user=None, edge_row=None,
                   flags=BA_FLAG_CHECK if (_BA_CHECK_DEFAULT if _validate is None else _validate) else 0)
    if _edge_row is not None:
        if (_edge_row.dtype != torch.int32 or not _edge_row.is_cuda or not _edge_row.is_contiguous()
                or _edge_row.numel() != n_edges):
            raise RuntimeError("droid_backends.ba: edge_row must be a contiguous CUDA int32 tensor with one entry per edge")
        prm.edge_row = _edge_row.data_ptr()
    elif targets.shape[0] != n_edges or weights.shape[0] != n_edges:
        raise RuntimeError("droid_backends.ba: targets / weights need one row per edge")
    if _shard is not None:
        prm.frame_lo, prm.frame_hi = int(_shard[0]), int(_shard[1])
    cb = None
    if _reduce is not None:
        cb = REDUCE_FN(_reduce)
        prm.reduce_fn = cb
    if poses.shape[0] < t1:
        raise RuntimeError("droid_backends.ba: t1 exceeds the pose buffer")
    with torch.cuda.device(poses.device):
        nws = _lib.droid_ba_workspace_bytes(ctypes.byref(prm))
        if nws == 0:
            raise RuntimeError("droid_backends.ba: invalid argument")
        ws = _workspace(poses.device, nws)
        dx = torch.zeros((max(P, 0), 6), dtype=torch.float32, device=poses.device)
        dz = torch.zeros((n_eta, HW), dtype=torch.float32, device=poses.device)
        rc = _lib.droid_ba(_p(poses), _p(disps), _p(intrinsics), _p(disps_sens), _p(targets),
                           _p(weights), _p(eta), _p(ii), _p(jj), ctypes.byref(prm), _p(dx), _p(dz),
                           ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(ws.numel()), _stream(poses))
        if rc == -6:
            st = ba_status(poses.device)
            why = []
            if st & 1:
                why.append("eta has %d rows but unique(ii U [t0,t1)) has another count (the reference fails on this "
                           "broadcast, src/droid_kernels.cu:1398)" % n_eta)
            if st & 2:
                why.append("an edge index lies outside [0, %d)" % nbuf)
            raise RuntimeError("droid_backends.ba: " + "; ".join(why or ["device-side input validation failed"]))
        _check(rc, "ba")
    return [dx, dz]


def ba_status(device=None):
    """Device-side status word of the last ba() on this device (forces a D2H copy; debug aid)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ws = _ws_cache.get((device.type, device.index, "ba"))
    if ws is None:
        return 0
    off = (-ws.data_ptr()) % 256
    return int(ws[off:off + 4].view(torch.int32).item())


def ba_tree(device=None):
    """Elimination tree the last large ba() on this device factorised its reduced camera system with:
    (range_end, level_first) as droid_chol_solve_tree_f64 takes them, ([], []) for a single chain.  Forces a D2H copy;
    debug / test aid."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ws = _ws_cache.get((device.type, device.index, "ba"))
    if ws is None:
        return [], []
    off = (-ws.data_ptr()) % 256
    st = ws[off:off + 256].view(torch.int32).cpu().tolist()
    if st[8] != 0x0b200c0d or st[18] <= 0:
        return [], []
    nr, nl = st[18], st[19]
    return st[20:20 + nr], st[36:36 + nl + 1]


def ba_chains(device=None):
    """End rows of the chains (the level-0 ranges of ba_tree; the last one = first separator row); empty = one chain."""
    range_end, level_first = ba_tree(device)
    return range_end[:level_first[1]] if level_first else []


def chol_solve_tree(A, b, first, range_end, level_first):
    """Extension (tests/bench): droid_chol_solve_tree_f64 -- SPD solve over an elimination tree of column ranges.
    first: CUDA int32 [len(range_end), n]; range_end / level_first: lists of ints (see include/droid_b200.h)."""
    n = A.shape[0]
    nr, nl = len(range_end), len(level_first) - 1
    if first.dtype != torch.int32 or first.numel() != nr * n or not first.is_cuda:
        raise RuntimeError("first must be a CUDA int32 tensor of len(range_end) * n entries")
    with torch.cuda.device(A.device):
        buf = torch.empty((n * n + n,), dtype=torch.float64, device=A.device)
        buf[: n * n] = A.reshape(-1)
        buf[n * n:] = b.reshape(-1)
        info = torch.zeros((1,), dtype=torch.int32, device=A.device)
        sep_rows = n - int(range_end[level_first[1] - 1]) if nl > 1 else 0
        wsb = int(_lib.droid_chol_tree_workspace_bytes(n, nr, sep_rows))
        ws = torch.empty((wsb,), dtype=torch.uint8, device=A.device)
        ends = (ctypes.c_int * nr)(*[int(e) for e in range_end])
        lf = (ctypes.c_int * (nl + 1))(*[int(e) for e in level_first])
        _check(_lib.droid_chol_solve_tree_f64(_p(buf), ctypes.c_void_p(buf.data_ptr() + 8 * n * n), n, _p(first.contiguous()),
                                              nr, ends, nl, lf, _p(info), _p(ws), ctypes.c_size_t(wsb), _stream(A)),
               "chol_solve_tree")
    return buf[n * n:].clone(), info


def chol_solve(A, b, first=None, chain_end=None):
    """Extension (tests/bench): solves A x = b for SPD fp64 A on the device; returns (x, info).
    first (int32 [n], optional): first non-zero column of every row -> envelope variant.
    chain_end (list of ints, optional) + first of shape [len(chain_end) + 1, n]: chains + separator variant
    (droid_chol_solve_chains_f64)."""
    n = A.shape[0]
    if chain_end is not None:
        k = len(chain_end)
        if first is None or first.dtype != torch.int32 or first.numel() != (k + 1) * n or not first.is_cuda:
            raise RuntimeError("first must be a CUDA int32 tensor of (len(chain_end) + 1) * n entries")
        with torch.cuda.device(A.device):
            buf = torch.empty((n * n + n,), dtype=torch.float64, device=A.device)
            buf[: n * n] = A.reshape(-1)
            buf[n * n:] = b.reshape(-1)
            info = torch.zeros((1,), dtype=torch.int32, device=A.device)
            wsb = int(_lib.droid_chol_chains_workspace_bytes(n, k, n - int(chain_end[-1])))
            ws = torch.empty((wsb,), dtype=torch.uint8, device=A.device)
            ends = (ctypes.c_int * k)(*[int(e) for e in chain_end])
            _check(_lib.droid_chol_solve_chains_f64(_p(buf), ctypes.c_void_p(buf.data_ptr() + 8 * n * n), n,
                                                    _p(first.contiguous()), k, ends, _p(info), _p(ws),
                                                    ctypes.c_size_t(wsb), _stream(A)), "chol_solve_chains")
        return buf[n * n:].clone(), info
    with torch.cuda.device(A.device):
        buf = torch.empty((n * n + n,), dtype=torch.float64, device=A.device)
        buf[: n * n] = A.reshape(-1)
        buf[n * n:] = b.reshape(-1)
        info = torch.zeros((1,), dtype=torch.int32, device=A.device)
        wsb = int(_lib.droid_chol_workspace_bytes(n))
        ws = torch.empty((wsb,), dtype=torch.uint8, device=A.device)
        bptr = ctypes.c_void_p(buf.data_ptr() + 8 * n * n)
        if first is None:
            _check(_lib.droid_chol_solve_f64(_p(buf), bptr, n, _p(info), _p(ws), ctypes.c_size_t(wsb),
                                             _stream(A)), "chol_solve")
        else:
            if first.dtype != torch.int32 or first.numel() != n or not first.is_cuda:
                raise RuntimeError("first must be a CUDA int32 tensor of n entries")
            _check(_lib.droid_chol_solve_env_f64(_p(buf), bptr, n, _p(first.contiguous()), _p(info), _p(ws),
                                                 ctypes.c_size_t(wsb), _stream(A)), "chol_solve_env")
    return buf[n * n:].clone(), info


def ba_sharded(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1,
               iterations, lm, ep, motion_only, group=None, edge_row=None, bounds=None):
    """Multi-GPU global BA.  Poses, depth maps and the edge list are replicated; rank r accumulates the
    source frames kx in its contiguous shard [bounds[r], bounds[r+1]) (default: shard_bounds, balanced by work), the
    reduced camera system -- packed to its structural non-zeros -- is summed with one NCCL all-reduce per iteration,
    every rank solves it (replicated, deterministic) and updates all poses plus the depths of its own frames; the
    depth maps are all-gathered at the end so every rank returns the same state as single-GPU ba().
    ``targets`` / ``weights`` hold either one row per edge (replicated), or only the rows of THIS rank's edges
    (``shard_edges``; in edge order) -- then ``edge_row`` (from ``shard_edge_rows``) maps an edge to its row; it is
    computed here when omitted."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    nbuf = disps.shape[0]
    if bounds is None:
        bounds = shard_bounds(ii, nbuf, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    if edge_row is None and targets.shape[0] != ii.shape[0]:
        edge_row, count = shard_edge_rows(ii, lo, hi)
        if count != targets.shape[0] or count != weights.shape[0]:
            raise RuntimeError("droid_backends.ba_sharded: targets / weights must hold one row per edge, or one row per "
                               "edge of this rank's shard (%d), got %d" % (count, targets.shape[0]))

    def _reduce(user, buf, n, stream):
        t = _as_tensor(buf, n, poses.device)
        dist.all_reduce(t, group=group)
        return 0

    out = ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations,
             lm, ep, motion_only, _shard=(lo, hi), _reduce=_reduce, _edge_row=edge_row)
    if not motion_only and world > 1:
        gather_depth_shards(disps, bounds, rank, group)
    return out


def shard_edges(ii, lo, hi):
    """Mask of the edges whose source frame lies in the shard [lo, hi): the rows of targets / weights a rank needs."""
    return (ii >= lo) & (ii < hi)


def shard_edge_rows(ii, lo, hi):
    """(edge_row, count): edge_row[e] (int32) = row of edge e among the shard's edges (meaningless for edges outside
    the shard, which the kernels never touch), count = number of edges in the shard (one D2H read)."""
    mask = shard_edges(ii, lo, hi)
    rows = torch.cumsum(mask.to(torch.int32), 0, dtype=torch.int32) - 1
    return rows.clamp_(min=0).contiguous(), int(mask.sum().item())


def shard_range(nbuf, world, rank):
    """Contiguous source-frame shard [lo, hi) of rank `rank` and the (padded) shard length: equal frame counts."""
    per = (nbuf + world - 1) // world
    return min(nbuf, rank * per), min(nbuf, (rank + 1) * per), per


def shard_bounds(ii, nbuf, world):
    """Contiguous source-frame shards balanced by WORK instead of by frame count ("greedy balance on
    out-degree"): frame k costs deg_k * (deg_k + 12) (linearisation ~ deg, Schur Gram ~ deg^2), the cut points are
    the world-quantiles of the cumulative cost.  Returns world + 1 frame indices; a pure function of the edge list,
    so every rank computes the same cuts.  ``ii`` may be a CPU or CUDA tensor (one small D2H read)."""
    deg = torch.bincount(ii.clamp(0, nbuf - 1), minlength=nbuf).to(torch.float64)
    cost = torch.cumsum(deg * (deg + 12.0), 0)
    total = float(cost[-1].item()) if nbuf > 0 else 0.0
    if total <= 0.0:
        return [shard_range(nbuf, world, r)[0] for r in range(world)] + [nbuf]
    q = torch.arange(1, world, dtype=torch.float64, device=cost.device) * (total / world)
    cuts = (torch.searchsorted(cost, q) + 1).clamp(max=nbuf).tolist() if world > 1 else []
    b = [0] + [int(c) for c in cuts] + [nbuf]
    for i in range(1, len(b)):                  # monotone
        b[i] = max(b[i], b[i - 1])
    return b


def gather_depth_shards(disps, bounds, rank, group=None):
    """Every rank updated only disps[bounds[rank]:bounds[rank+1]]; all-gather the shards (padded to the longest) so
    that all ranks hold the full buffer."""
    import torch.distributed as dist
    world = len(bounds) - 1
    per = max(1, max(int(bounds[r + 1]) - int(bounds[r]) for r in range(world)))
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    flat = disps.new_zeros((per * world,) + tuple(disps.shape[1:]))
    dist.all_gather_into_tensor(flat, _pad_rows(disps[lo:hi], per), group=group)
    for r in range(world):
        a, b = int(bounds[r]), int(bounds[r + 1])
        if r != rank and b > a:
            disps[a:b].copy_(flat[r * per:r * per + (b - a)])
    return disps


def _pad_rows(t, rows):
    if t.shape[0] == rows:
        return t.contiguous()
    out = t.new_zeros((rows,) + tuple(t.shape[1:]))
    out[: t.shape[0]] = t
    return out


class _CudaArray:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3}


def _as_tensor(ptr, n, device):
    return torch.as_tensor(_CudaArray(int(ptr), int(n)), device=device)
