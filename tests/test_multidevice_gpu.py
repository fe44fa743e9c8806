"""One process, several devices: a context over two GPUs (vb_ctx_create with a device list) shards the host-buffer
batches over its slots; results must equal the one-device context's.  Skipped on a one-GPU box (the driver's GPU test
tier); run with `gpurun --gpus 2`."""
import ctypes as C

import numpy as np
import pytest

from mr_vamp_b200 import scenes

pytestmark = pytest.mark.gpu


def _device_count():
    lib = C.CDLL("libcudart.so.12")
    n = C.c_int(0)
    return n.value if lib.cudaGetDeviceCount(C.byref(n)) == 0 else 0


def test_two_device_context_matches_one_device(vb):
    try:
        n_dev = _device_count()
    except OSError:
        pytest.skip("libcudart not loadable from python")
    if n_dev < 2:
        pytest.skip(f"{n_dev} CUDA device(s) visible: needs 2")
    ctx1, ctx2 = vb.Context([0]), vb.Context([0, 1])
    assert ctx2.n_devices == 2
    spec = scenes.random_scene(7)
    for name in ("panda", "ur5", "baxter"):
        r1, r2 = vb.Robot(name, ctx=ctx1), vb.Robot(name, ctx=ctx2)
        e1, e2 = scenes.apply(spec, vb.Environment(ctx1)), scenes.apply(spec, vb.Environment(ctx2))
        rng = np.random.default_rng(3)
        q = (r1.lower_bounds + rng.random((200_003, r1.dimension), dtype=np.float32) * r1.ranges).astype(np.float32)
        assert np.array_equal(r1.fkcc_batch(q, e1), r2.fkcc_batch(q, e2)), name
        assert np.array_equal(r1.validate_batch(q, e1), r2.validate_batch(q, e2)), name
        a, b = q[:20_001], q[20_001:40_002]
        assert np.array_equal(r1.validate_motion_batch(a, b, e1), r2.validate_motion_batch(a, b, e2)), name
        assert np.allclose(r1.fk_batch(q[:1001]), r2.fk_batch(q[:1001]), rtol=0, atol=0), name
    # and a context on the SECOND device alone (kernel attributes are per device)
    ctx_b = vb.Context([1])
    rb = vb.Robot("panda", ctx=ctx_b)
    eb = scenes.apply(spec, vb.Environment(ctx_b))
    r1 = vb.Robot("panda", ctx=ctx1)
    e1 = scenes.apply(spec, vb.Environment(ctx1))
    q = (r1.lower_bounds + np.random.default_rng(4).random((50_000, 7), dtype=np.float32) * r1.ranges).astype(np.float32)
    assert np.array_equal(r1.fkcc_batch(q, e1), rb.fkcc_batch(q, eb))
