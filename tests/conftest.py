import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def apply_ref(spec, renv):
    """Apply a scene spec (mr_vamp_b200.scenes) to the reference oracle environment."""
    for item in spec:
        kind = item[0]
        if kind == "sphere":
            renv.add_sphere(*[float(np.float32(v)) for v in item[1:5]])
        elif kind == "cuboid":
            renv.add_cuboid(item[1])
        elif kind == "capsule":
            renv.add_capsule(item[1])
        elif kind == "heightfield":
            renv.add_heightfield(item[1], item[2])
        else:
            raise ValueError(kind)
    return renv


@pytest.fixture(scope="session")
def ref():
    from oracle import ref as r

    if not r.available():
        r.build()
    if not r.available():
        pytest.skip("oracle/_ref/libvamp_ref.so not built (needs the reference checkout)")
    return r


@pytest.fixture(scope="session")
def vb():
    import mr_vamp_b200.api as api

    api.lib()
    return api


@pytest.fixture(scope="session")
def gpu_ctx(vb):
    return vb.default_context()
