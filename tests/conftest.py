import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "_ref"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib_built():
    """Builds libdroid_b200.so if it is missing (nvcc cross-compiles without a GPU)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("droid_build", os.path.join(ROOT, "droid-slam_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build()


@pytest.fixture(scope="session")
def backends(lib_built):
    import droid_backends
    return droid_backends


@pytest.fixture(scope="session")
def ref_ext():
    """The unmodified reference extension built by oracle/build_ref.sh (oracle/_ref travels to the GPU box with the
    snapshot).  Only `-m gpu` tests ask for it, and there a missing / unloadable extension is a FAILURE, not a skip:
    the live comparisons are the strongest parity evidence this repo has."""
    try:
        import torch  # noqa: F401
        import droid_backends_ref
        return droid_backends_ref
    except Exception as ex:  # noqa: BLE001
        pytest.fail("the reference extension oracle/_ref/droid_backends_ref*.so cannot be imported (%r): run "
                    "`bash oracle/build_ref.sh` in the build container before `gpurun`" % (ex,))
