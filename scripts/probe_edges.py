"""Quick device-side probe (not the bench): validate_motion over the library's synthetic RRTC-shaped edges.
usage: python scripts/probe_edges.py [robot] [mbm|cage|clutter]"""
import ctypes as C
import os
import sys

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

ctx = api.default_context()
L = api.lib()
name = sys.argv[1] if len(sys.argv) > 1 else "panda"
which = sys.argv[2] if len(sys.argv) > 2 else "mbm"
robot = api.Robot(name)
if which == "mbm":
    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{name}.json.gz"))
    probs = mbm["problems"]
    spec = scenes.problem_to_spec(probs["table_pick" if "table_pick" in probs else sorted(probs)[0]][0])
else:
    spec = {"cage": scenes.cage(), "clutter": scenes.random_scene(7)}[which]
env = scenes.apply(spec, api.Environment())
n = 1 << 22
dim = robot.dimension
a, b, bits = C.c_void_p(), C.c_void_p(), C.c_void_p()
api._check(L.vb_dev_alloc(ctx.handle, 0, n * dim * 4, C.byref(a)))
api._check(L.vb_dev_alloc(ctx.handle, 0, n * dim * 4, C.byref(b)))
api._check(L.vb_dev_alloc(ctx.handle, 0, n // 8, C.byref(bits)))
api._check(L.vb_random_edges_d(ctx.handle, 0, robot.robot_id, 7, 0, n, C.c_float(0.05), C.c_float(1.0), a, b))
for rep in range(3):
    api._check(L.vb_timer_start(ctx.handle, 0))
    api._check(L.vb_validate_motion_d(ctx.handle, 0, robot.robot_id, None, env.handle, a, b, n, robot.resolution, bits, None))
    ms = C.c_double()
    api._check(L.vb_timer_stop(ctx.handle, 0, C.byref(ms)))
cnt = C.c_uint64()
api._check(L.vb_popcount_d(ctx.handle, 0, bits, n, C.byref(cnt)))
print(f"{name:7s} {which:8s} edges={n} {ms.value:8.2f} ms  {n / ms.value / 1e3:8.2f} Medges/s  valid={cnt.value / n:.3f}", flush=True)
