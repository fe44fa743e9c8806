"""Dev probe (not the bench): vb_fkcc_d over explicit device-resident configurations, MBM table_pick scene 1.
usage: [VAMP_B200_LIB=...] python scripts/probe_fkcc.py [robot] [n]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "panda"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 25
ctx = api.default_context()
L = api.lib()
robot = api.Robot(name)
mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{name}.json.gz"))
probs = mbm["problems"]
spec = scenes.problem_to_spec(probs["table_pick"][0] if "table_pick" in probs else next(iter(probs.values()))[0])
env = scenes.apply(spec, api.Environment())
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(42)
lo, rg = torch.tensor(robot.lower_bounds, device=dev), torch.tensor(robot.ranges, device=dev)
cfg = torch.rand((n, robot.dimension), device=dev, generator=gen).mul_(rg).add_(lo)
bits = torch.zeros((n + 31) // 32, device=dev, dtype=torch.int32)
torch.cuda.synchronize()
best = 1e9
for rep in range(6):
    ms = C.c_double()
    api._check(L.vb_timer_start(ctx.handle, 0))
    api._check(L.vb_fkcc_d(ctx.handle, 0, robot.robot_id, None, env.handle, cfg.data_ptr(), n, bits.data_ptr(), None))
    api._check(L.vb_timer_stop(ctx.handle, 0, C.byref(ms)))
    if rep:
        best = min(best, ms.value)
cnt = C.c_uint64()
api._check(L.vb_popcount_d(ctx.handle, 0, bits.data_ptr(), n, C.byref(cnt)))
print(f"{name:7s} mbm explicit n={n} best {best:8.3f} ms  {n / best / 1e3:8.2f} Mcfg/s  valid={cnt.value}", flush=True)
