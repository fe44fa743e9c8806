"""Allocation-free motion validation (vb_validate_motion_noalloc_d): scratch in a file-scope static device array.

Collected last on purpose.  The new entry point is called through the C ABI twice in
a row, on different batches, and must give exactly the verdict words of vb_validate_motion_d (integer/bit work:
bit-exact); the second call also proves that what the first left in the shared array does not leak into the next.
"""
import ctypes as C

import numpy as np
import pytest

from mr_vamp_b200 import scenes

pytestmark = pytest.mark.gpu


def _edges(robot, seed, n):
    rng = np.random.default_rng(seed)
    a = (robot.lower_bounds + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)
    step = rng.normal(size=(n, robot.dimension)).astype(np.float32)
    step *= (rng.uniform(0.02, 1.5, (n, 1)) / np.linalg.norm(step, axis=1, keepdims=True)).astype(np.float32)
    b = np.clip(a + step, robot.lower_bounds, robot.lower_bounds + robot.ranges).astype(np.float32)
    return a, b


class _Dev:
    """Caller-owned device buffers through the C ABI (vb_dev_alloc / upload / download / free)."""

    def __init__(self, vb, ctx):
        self.vb, self.ctx, self.ptrs = vb, ctx, []

    def alloc(self, nbytes):
        p = C.c_void_p()
        self.vb._check(self.vb.lib().vb_dev_alloc(self.ctx.handle, 0, nbytes, C.byref(p)))
        self.ptrs.append(p)
        return p

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.alloc(arr.nbytes)
        self.vb._check(self.vb.lib().vb_dev_upload(self.ctx.handle, 0, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes))
        return p

    def download(self, p, dtype, count):
        out = np.zeros(count, dtype=dtype)
        self.vb._check(self.vb.lib().vb_dev_download(self.ctx.handle, 0, out.ctypes.data_as(C.c_void_p), p, out.nbytes))
        return out

    def free(self):
        for p in self.ptrs:
            self.vb.lib().vb_dev_free(self.ctx.handle, 0, p)
        self.ptrs = []


@pytest.mark.parametrize("name", ["panda", "ur5"])
def test_noalloc_motion_matches_allocating_entry_point(vb, gpu_ctx, name):
    L = vb.lib()
    robot = vb.Robot(name)
    env = scenes.apply(scenes.random_scene(11), vb.Environment())
    assert L.vb_noalloc_max_edges() == 262144
    dev = _Dev(vb, gpu_ctx)
    try:
        # two calls in a row; the second batch is smaller, so stale counts / prefixes of the first stay behind it
        for seed, n in ((21, 5000), (22, 1777)):
            a, b = _edges(robot, seed, n)
            a_d, b_d = dev.upload(a), dev.upload(b)
            words = (n + 31) // 32
            want_d, got_d = dev.alloc(4 * words), dev.alloc(4 * words)
            vb._check(L.vb_validate_motion_d(gpu_ctx.handle, 0, robot.robot_id, None, env.handle, a_d, b_d, n, 0, want_d, None))
            vb._check(L.vb_validate_motion_noalloc_d(gpu_ctx.handle, 0, robot.robot_id, None, env.handle, a_d, b_d, n, 0, got_d))
            want = dev.download(want_d, np.uint32, words)
            got = dev.download(got_d, np.uint32, words)
            ok = vb.unpack_bits(want, n)
            assert ok.any() and not ok.all(), "degenerate batch: the comparison would prove nothing"
            assert np.array_equal(got, want), f"{name}: {int((vb.unpack_bits(got, n) != ok).sum())} of {n} verdicts differ"
    finally:
        dev.free()


def test_noalloc_rejects_oversized_batches_and_attachments(vb, gpu_ctx):
    L = vb.lib()
    robot = vb.Robot("panda")
    env = scenes.apply(scenes.cage(), vb.Environment())
    dev = _Dev(vb, gpu_ctx)
    try:
        p = dev.alloc(64)
        rc = L.vb_validate_motion_noalloc_d(gpu_ctx.handle, 0, robot.robot_id, None, env.handle, p, p, 262145, 0, p)
        assert rc == 1 and b"262144" in L.vb_last_error()
        att = vb.Attachment([0, 0, 0.1], [0, 0, 0, 1])
        att.add_spheres([vb.Sphere([0, 0, 0], 0.03)])
        env.attach(att)
        rc = L.vb_validate_motion_noalloc_d(gpu_ctx.handle, 0, robot.robot_id, None, env.handle, p, p, 1, 0, p)
        assert rc == 4 and b"attachment" in L.vb_last_error()
    finally:
        dev.free()
