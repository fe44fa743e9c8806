"""Dev tool: work counters of the Panda warp walk (library built with -DVB_STATS, see vb_walk.cuh).
usage: VAMP_B200_LIB=build/lib_stats.so python scripts/walk_stats.py [mbm|empty|clutter|cage]"""
import ctypes as C
import os
import sys

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

ctx = api.default_context()
L = api.lib()
rname = os.environ.get("VB_ROBOT", "panda")
robot = api.Robot(rname)
for which in (sys.argv[1:] or ["mbm"]):
    if which == "mbm":
        mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{rname}.json.gz"))
        probs = mbm["problems"]
        spec = scenes.problem_to_spec(probs["table_pick" if "table_pick" in probs else sorted(probs)[0]][0])
    else:
        spec = {"empty": [], "clutter": scenes.random_scene(7), "cage": scenes.cage()}[which]
    env = scenes.apply(spec, api.Environment())
    n = 1 << 22
    bits = C.c_void_p()
    api._check(L.vb_dev_alloc(ctx.handle, 0, n // 8, C.byref(bits)))
    out = (C.c_ulonglong * 32)()
    L.vb_debug_walk_stats(out, 1)
    api._check(L.vb_fkcc_random_d(ctx.handle, 0, robot.robot_id, None, env.handle, 42, 0, n, bits))
    api._check(L.vb_sync(ctx.handle))
    L.vb_debug_walk_stats(out, 0)
    v = [float(x) for x in out]
    w = v[0]
    print(f"== {which}: {int(w)} warp steps; live lanes/step {v[19] / w:.1f}; collided after self {v[18] / w:.2f}")
    print(f" self: fired (lane,group) per step {v[14] / w:.1f}; distinct groups fired per step {v[15] / w:.1f}")
    print(f" pair: units/step {v[10] / w:.1f}; passes/step {v[11] / w:.2f}; b tests/step {v[12] / w:.1f}; "
          f"lane efficiency {v[12] / max(32 * v[13], 1):.2f} (sum b / 32 max b); mean b/unit {v[12] / max(v[10], 1):.1f}")
    print(f" env: items/step {v[1] / w:.1f} (cert {v[16] / w:.1f}); passes/step {v[2] / w:.2f}; candidates/step {v[3] / w:.1f}; "
          f"lane efficiency {v[3] / max(32 * v[4], 1):.2f}; cand/item {v[3] / max(v[1], 1):.2f}; guard hits/step {v[5] / w:.1f}")
    print(f" fine: units/step {v[6] / w:.1f}; passes/step {v[7] / w:.2f}; tests/step {v[8] / w:.1f}; "
          f"lane efficiency {v[8] / max(32 * v[9], 1):.2f}; tests/unit {v[8] / max(v[6], 1):.2f}")
    sys.stdout.flush()
    api._check(L.vb_dev_free(ctx.handle, 0, bits))
