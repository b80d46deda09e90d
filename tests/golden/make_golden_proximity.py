"""Golden vectors for the edge selection of FactorGraph.add_proximity_factors.

Runs the UNMODIFIED reference method (/root/reference/droid_slam/factor_graph.py:305-369) on CPU in this container:
the module's imports that are absent here (lietorch, matplotlib, modules.corr -> droid_backends, geom.projective_ops)
are satisfied by empty stand-in modules -- the method itself uses none of them -- and `self` is a plain namespace whose
`video.distance` returns a stored distance vector and whose `add_factors` records the edge list.

  python tests/golden/make_golden_proximity.py      ->  tests/golden/ref_proximity.npz
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import proximity_cases as PC  # noqa: E402


def load_reference_class():
    for name in ("lietorch", "matplotlib", "matplotlib.pyplot", "modules", "modules.corr", "geom", "geom.projective_ops"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["lietorch"].SE3 = object
    sys.modules["modules.corr"].CorrBlock = object
    sys.modules["modules.corr"].AltCorrBlock = object
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["geom"].projective_ops = sys.modules["geom.projective_ops"]
    sys.path.insert(0, "/root/reference/droid_slam")
    import factor_graph  # the reference module, untouched
    return factor_graph.FactorGraph


def run_reference(FG, case):
    out = {}
    video = types.SimpleNamespace(counter=types.SimpleNamespace(value=case["t"]), stereo=case["stereo"],
                                  distance=lambda ii, jj, beta: torch.from_numpy(case["d"].copy()))
    as_t = lambda a: torch.as_tensor(a, dtype=torch.long)  # noqa: E731
    half = len(case["ii_exist"]) // 2
    self = types.SimpleNamespace(
        video=video, device="cpu", max_factors=case["max_factors"],
        ii=as_t(case["ii_exist"][:half]), jj=as_t(case["jj_exist"][:half]),
        ii_bad=as_t([]), jj_bad=as_t([]),
        ii_inac=as_t(case["ii_exist"][half:]), jj_inac=as_t(case["jj_exist"][half:]),
        add_factors=lambda ii, jj, remove=False: out.update(ii=ii.numpy().copy(), jj=jj.numpy().copy()))
    FG.add_proximity_factors(self, t0=case["t0"], t1=case["t1"], rad=case["rad"], nms=case["nms"], beta=0.25,
                             thresh=case["thresh"], remove=False)
    return out["ii"].astype(np.int64), out["jj"].astype(np.int64)


def main():
    FG = load_reference_class()
    store = {}
    for key in sorted(PC.CASES):
        case = PC.make_case(key)
        ii, jj = run_reference(FG, case)
        store[key + "_ii"], store[key + "_jj"] = ii, jj
        store[key + "_digest"] = PC.digest(case)
        print(key, "edges", len(ii))
    np.savez_compressed(os.path.join(HERE, "ref_proximity.npz"), **store)


if __name__ == "__main__":
    main()
