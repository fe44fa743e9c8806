"""Dev tool: per-phase cycle shares of the Panda warp walk (library built with -DVB_PHASE_CLOCKS, see vb_walk.cuh).
usage: VAMP_B200_LIB=build/lib_clk.so python scripts/phase_clocks.py [mbm|empty|clutter] [robot]"""
import ctypes as C
import os
import sys

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

names = {0: "config + pair lookup", 1: "chain", 2: "const groups + env_fetch issue", 3: "self guards", 4: "self pushes",
         5: "filter stage", 6: "pair stage", 7: "env classify", 8: "env pushes", 9: "env + fine stages", 10: "barrier",
         11: "config synthesis", 12: "check_warp total"}
ctx = api.default_context()
L = api.lib()
rname = sys.argv[2] if len(sys.argv) > 2 else "panda"
robot = api.Robot(rname)
which = sys.argv[1] if len(sys.argv) > 1 else "mbm"
if which == "mbm":
    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{rname}.json.gz"))
    probs = mbm["problems"]
    spec = scenes.problem_to_spec(probs["table_pick" if "table_pick" in probs else sorted(probs)[0]][0])
else:
    spec = {"empty": [], "clutter": scenes.random_scene(7)}[which]
env = scenes.apply(spec, api.Environment())
n = 1 << 22
bits = C.c_void_p()
api._check(L.vb_dev_alloc(ctx.handle, 0, n // 8, C.byref(bits)))
out = (C.c_ulonglong * 16)()
for rep in range(2):
    L.vb_debug_phase_clocks(out, 1)
    api._check(L.vb_fkcc_random_d(ctx.handle, 0, robot.robot_id, None, env.handle, 42, 0, n, bits))
    api._check(L.vb_sync(ctx.handle))
L.vb_debug_phase_clocks(out, 0)
v = list(out)
warps = n // 32
total = v[10] + v[11] + v[12]
print(f"{which}: cycles per warp-iteration {total / warps:.0f}")
for k in sorted(names):
    print(f"  {names[k]:32s} {v[k] / warps:8.0f} cycles  {100.0 * v[k] / total:5.1f}%")
