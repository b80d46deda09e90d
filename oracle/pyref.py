"""TEST INFRASTRUCTURE ONLY -- loader for the UNMODIFIED Python reference staged under oracle/_ref/pyref/
(oracle/stage_pyref.sh) plus lietorch_backends (oracle/build_lietorch.sh).

Only tests/ and bench.py's cpu_baseline / --impl reference legs import this.  It gives the reference's own callers of
the hot path -- droid_slam/modules/corr.py (CorrBlock, AltCorrBlock), depth_video.py (DepthVideo.distance / ba /
reproject), factor_graph.py (FactorGraph), geom/ba.py (BA) -- bound to a `droid_backends` module of the caller's
choice (this repo's ctypes host or the reference extension oracle/_ref/droid_backends_ref), which is how the drop-in
claim is checked: same Python, two back ends, same numbers.

Imports the hot path cannot satisfy offline are stand-ins under oracle/standins/ (torch_scatter, matplotlib,
data_readers; each says what it replaces).  The ONE semantic accommodation is made in memory, only for cpu():
geom/projective_ops.py:105 builds the stereo baseline with a hard-coded device="cuda"; on CPU tensors that token is
replaced by the pose tensor's device while the module is loaded (the staged file stays unmodified).
"""
import importlib
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
PYREF = os.path.join(HERE, "_ref", "pyref")
STANDINS = os.path.join(HERE, "standins")

_REF_MODULES = ("droid_backends", "lietorch", "modules", "modules.corr", "modules.extractor", "modules.gru",
                "modules.clipping", "geom", "geom.ba", "geom.chol", "geom.projective_ops", "geom.graph_utils",
                "geom.losses", "depth_video", "factor_graph", "droid_net", "droid_frontend", "droid_backend",
                "motion_filter", "trajectory_filler", "droid", "torch_scatter", "matplotlib", "matplotlib.pyplot",
                "data_readers", "data_readers.rgbd_utils")


def available():
    """(ok, why): the staged reference and lietorch_backends are both present."""
    if not os.path.isfile(os.path.join(PYREF, "droid_slam", "modules", "corr.py")):
        return False, "oracle/_ref/pyref missing: run `bash oracle/stage_pyref.sh` in the build container"
    so = [f for f in os.listdir(os.path.join(HERE, "_ref")) if f.startswith("lietorch_backends") and f.endswith(".so")]
    if not so:
        return False, "oracle/_ref/lietorch_backends*.so missing: run `bash oracle/build_lietorch.sh`"
    return True, ""


def _paths():
    for p in (os.path.join(HERE, "_ref"), STANDINS, PYREF, os.path.join(PYREF, "droid_slam")):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    # the reference's own module names win over same-named modules of other trees already imported
    for missing in ("matplotlib", "torch_scatter"):
        try:
            if missing not in sys.modules:
                importlib.import_module(missing)
        except Exception:  # noqa: BLE001
            pass


def load(backend, cpu=False):
    """Imports the staged reference with `droid_backends` bound to `backend` (a module object) and returns a namespace
    of its modules: corr, depth_video, factor_graph, droid_net, pops, ba, lietorch.  A second call re-binds the back
    end of the already imported modules (use(backend))."""
    ok, why = available()
    if not ok:
        raise ImportError(why)
    _paths()
    prev = sys.modules.get("droid_backends")
    sys.modules["droid_backends"] = backend
    try:
        if cpu and "geom.projective_ops" not in sys.modules:
            import geom  # noqa: F401  (package)
            path = os.path.join(PYREF, "droid_slam", "geom", "projective_ops.py")
            src = open(path).read()
            token = 'device="cuda"'
            if src.count(token) != 1:
                raise ImportError("geom/projective_ops.py: expected exactly one %s" % token)
            mod = types.ModuleType("geom.projective_ops")
            mod.__file__ = path
            mod.__package__ = "geom"
            sys.modules["geom.projective_ops"] = mod
            exec(compile(src.replace(token, "device=Gij.data.device"), path, "exec"), mod.__dict__)
            sys.modules["geom"].projective_ops = mod
        ns = types.SimpleNamespace()
        ns.lietorch = importlib.import_module("lietorch")
        ns.pops = importlib.import_module("geom.projective_ops")
        ns.ba = importlib.import_module("geom.ba")
        ns.corr = importlib.import_module("modules.corr")
        ns.droid_net = importlib.import_module("droid_net")
        ns.depth_video = importlib.import_module("depth_video")
        ns.factor_graph = importlib.import_module("factor_graph")
    finally:
        if prev is not None:
            sys.modules["droid_backends"] = prev
        else:
            sys.modules.pop("droid_backends", None)
    use(ns, backend)
    return ns


def use(ns, backend):
    """Points every staged module that holds a `droid_backends` global at `backend`."""
    for m in (ns.corr, ns.depth_video):
        m.droid_backends = backend
    ns.backend = backend
    return ns
