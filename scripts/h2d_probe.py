"""Dev probe for the end-to-end scaling question: host->device copy rate per rank when N ranks copy at
once from pinned memory, with and without binding the rank's thread + staging memory to the GPU's NUMA node.
usage (under torchrun / one process per GPU): python scripts/h2d_probe.py [bind]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from mr_vamp_b200 import api, sharding, hostbind

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
bind = len(sys.argv) > 1 and sys.argv[1] == "bind"
info = hostbind.bind_to_gpu(local) if bind else hostbind.describe(local)
info.pop('_cpus', None)
ctx = api.Context([local])
L, h = api.lib(), ctx.handle
comm = sharding.NcclComm(ctx, rank, world, f"/tmp/vb_h2d_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}")
n = 1 << 30
buf = api.pinned_empty((n // 4,), np.float32)
buf[:] = 1.0
d = C.c_void_p()
api._check(L.vb_dev_alloc(h, 0, n, C.byref(d)))
for rep in range(3):
    comm.barrier()
    t0 = time.perf_counter()
    api._check(L.vb_dev_upload(h, 0, d, buf.ctypes.data_as(C.c_void_p), n))
    api._check(L.vb_sync(h))
    dt = time.perf_counter() - t0
rates = comm.allgather_small(np.array([n / dt / 1e9], dtype=np.float32))[:, 0].view(np.float32)
if rank == 0:
    print(f"world {world} bind={bind}: H2D GB/s per rank {[round(float(r), 1) for r in rates]} sum {float(rates.sum()):.1f}; rank0 {info}", flush=True)
comm.barrier()
comm.close()
