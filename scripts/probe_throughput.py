"""Quick device-side throughput probe (not the bench): fkcc over synthetic configurations."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

ctx = api.default_context()
L = api.lib()
cases = [("panda", "mbm"), ("panda", "empty"), ("panda", "cage"), ("panda", "clutter"), ("ur5", "clutter"),
         ("fetch", "clutter"), ("baxter", "clutter"), ("sphere", "clutter")]
if len(sys.argv) > 2:
    cases = [(sys.argv[1], sys.argv[2])]
for name, spec_name in cases:
    robot = api.Robot(name)
    if spec_name == "mbm":
        import os
        mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{name}.json.gz"))
        spec = scenes.problem_to_spec(next(iter(mbm["problems"].values()))[0] if name != "panda" else mbm["problems"]["table_pick"][0])
    else:
        spec = {"empty": [], "cage": scenes.cage(), "clutter": scenes.random_scene(7)}[spec_name]
    env = scenes.apply(spec, api.Environment())
    n = 1 << 24
    bits = C.c_void_p()
    api._check(L.vb_dev_alloc(ctx.handle, 0, n // 8, C.byref(bits)))
    for rep in range(3):
        api._check(L.vb_timer_start(ctx.handle, 0))
        api._check(L.vb_fkcc_random_d(ctx.handle, 0, robot.robot_id, None, env.handle, 42, 0, n, bits))
        ms = C.c_double()
        api._check(L.vb_timer_stop(ctx.handle, 0, C.byref(ms)))
    cnt = C.c_uint64()
    api._check(L.vb_popcount_d(ctx.handle, 0, bits, n, C.byref(cnt)))
    print(f"{name:7s} {spec_name:8s} n={n} {ms.value:8.2f} ms  {n / ms.value / 1e3:8.2f} Mcfg/s  valid={cnt.value / n:.3f}", flush=True)
    api._check(L.vb_dev_free(ctx.handle, 0, bits))
