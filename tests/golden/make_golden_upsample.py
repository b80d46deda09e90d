"""Golden vectors for cvx_upsample (droid_slam/droid_net.py:21-35), produced by the UNMODIFIED reference
function on CPU in this container: droid_net.py is imported with empty stand-in modules for the imports that are absent
here (the function itself is plain torch).

  python tests/golden/make_golden_upsample.py      ->  tests/golden/ref_cvx_upsample.npz
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import upsample_cases as UC  # noqa: E402


def load_reference():
    need = {"modules": [], "modules.extractor": ["BasicEncoder"], "modules.corr": ["CorrBlock"], "modules.gru": ["ConvGRU"],
            "modules.clipping": ["GradientClip"], "lietorch": ["SE3"], "geom": [], "geom.ba": ["BA"],
            "geom.projective_ops": [], "geom.graph_utils": ["graph_to_edge_list", "keyframe_indicies"],
            "torch_scatter": ["scatter_mean"]}
    for name, attrs in need.items():
        m = sys.modules.setdefault(name, types.ModuleType(name))
        for a in attrs:
            setattr(m, a, object)
    sys.modules["geom"].projective_ops = sys.modules["geom.projective_ops"]
    sys.path.insert(0, "/root/reference/droid_slam")
    import droid_net  # the reference module, untouched
    return droid_net.cvx_upsample


def main():
    f = load_reference()
    store = {}
    for key in sorted(UC.CASES):
        data, mask = UC.make_case(key)
        out = f(torch.from_numpy(data), torch.from_numpy(mask)).numpy()
        store[key] = out
        store[key + "_digest"] = UC.digest(data, mask)
        print(key, out.shape)
    np.savez_compressed(os.path.join(HERE, "ref_cvx_upsample.npz"), **store)


if __name__ == "__main__":
    main()
