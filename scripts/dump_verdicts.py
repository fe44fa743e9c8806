"""Dev tool: verdict words (+ optionally margins) of vb_fkcc_d over the probe's explicit configurations, to a .npy file.
usage: [VAMP_B200_LIB=...] python scripts/dump_verdicts.py out.npy [n] [margin]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from mr_vamp_b200 import api, scenes

out = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 25
want_margin = len(sys.argv) > 3
ctx = api.default_context()
L = api.lib()
robot = api.Robot("panda")
mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, "mbm_panda.json.gz"))
env = scenes.apply(scenes.problem_to_spec(mbm["problems"]["table_pick"][0]), api.Environment())
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(42)
lo, rg = torch.tensor(robot.lower_bounds, device=dev), torch.tensor(robot.ranges, device=dev)
cfg = torch.rand((n, 7), device=dev, generator=gen).mul_(rg).add_(lo)
bits = torch.zeros((n + 31) // 32, device=dev, dtype=torch.int32)
mg = torch.zeros(n, device=dev, dtype=torch.float32) if want_margin else None
for rep in range(3):  # repeated: a race would show as run-to-run differences too
    api._check(L.vb_fkcc_d(ctx.handle, 0, robot.robot_id, None, env.handle, cfg.data_ptr(), n, bits.data_ptr(),
                           mg.data_ptr() if want_margin else None))
    api._check(L.vb_sync(ctx.handle))
    w = bits.cpu().numpy().view(np.uint32).copy()
    print(out, "rep", rep, "valid", int(np.unpackbits(w.view(np.uint8)).sum()), flush=True)
np.save(out, w)
if want_margin:
    np.save(out.replace(".npy", "_margin.npy"), mg.cpu().numpy())
    np.save(out.replace(".npy", "_cfg.npy"), cfg.cpu().numpy())
