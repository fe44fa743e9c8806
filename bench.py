#!/usr/bin/env python
"""Benchmark of the fkcc hot path (BASELINE.json: "Panda collision-checked configs/sec & edges
validated/sec at 1/2/4/8 B200").

A step = one pass of batched Panda fkcc (Robot::fkcc, robots/panda/fk.hh:1335-6276) over
`--configs` synthetic configurations per GPU against the MotionBenchMaker table_pick scene 1
(BASELINE.json configs[1]).  Prints ONE JSON line (rank 0):

  value      configs/s, whole job, inputs already resident in HBM (CUDA events on the launch stream)
  e2e        the same metric through the C-ABI call a host application makes (vb_fkcc): pinned HOST
             configurations in, verdict words out, copies inside the timed region
  roofline   FP32 (CUDA-core FMA pipe) roofline of the fkcc kernel, algorithmic flops W measured
             on the device (vb_fkcc_work_d); plus the HBM view of the same launch
  cpu_baseline  the reference's own AVX2 code (oracle/_ref) timed on this box's host cores

`--impl reference` times the reference CPU implementation instead (bounded sample per step).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "panda_fkcc_configs_per_s"
UNIT = "configs/s"
WORKLOAD = "panda_fkcc_mbm_table_pick_scene1"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--configs", type=int, default=100_000_000, help="configurations per GPU per step")
    ap.add_argument("--edges", type=int, default=1 << 22, help="edges per GPU per step for the edges/s half of the metric")
    ap.add_argument("--no-robots", action="store_true", help="skip the UR5 / Fetch / Baxter block (configs[3]) of the N = 1 line")
    ap.add_argument("--ref-sample", type=int, default=0, help="reference arm: configurations per step (0 = auto)")
    return ap.parse_args()


def bench_config(n: int) -> dict:
    """`config` of the JSON line: the workload, identical for both arms (run-specific figures live elsewhere)."""
    return {"workload": WORKLOAD, "configs_per_gpu_per_step": n, "robot": "PandaBase<0,0,0>",
            "environment": "MotionBenchMaker panda table_pick scene 1 (10 z-aligned cuboids + 2 z-aligned capsules)",
            "inputs": "library stream vb_random_configs(seed 42): uniform in the joint ranges",
            "l2": f"inputs {28 * n / 1e9:.2f} GB per step, larger than the 126 MB L2",
            "sharding": "configurations sharded by rank (contiguous ranges), environment + grid replicated per GPU; "
                        "one collective: ncclAllGather of the verdict words inside the timed step (N > 1)"}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def bench_scene():
    from mr_vamp_b200 import scenes

    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, "mbm_panda.json.gz"))
    problem = mbm["problems"]["table_pick"][0]
    return scenes.problem_to_spec(problem), problem


def host_configs(n: int, seed: int) -> np.ndarray:
    """Uniform configurations in the Panda joint ranges (scale_configuration, robots/panda/fk.hh:14-37): the library's
    synthetic stream (host-side generator of the C ABI; no device needed)."""
    from mr_vamp_b200 import api

    out = np.zeros((n, 7), dtype=np.float32)
    api._check(api.lib().vb_random_configs(api.ROBOT_IDS["panda"], seed, 0, n, api._ptr(out)))
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows: list[list[str]] = []
        self._stop = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                    capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.03)

    def __enter__(self):
        self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self) -> dict:
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_info() -> dict:
    model, smt = None, None
    try:
        with open("/proc/cpuinfo") as fh:
            for ln in fh:
                if ln.startswith("model name"):
                    model = ln.split(":", 1)[1].strip()
                    break
        with open("/sys/devices/system/cpu/smt/active") as fh:
            smt = fh.read().strip() == "1"
    except Exception:
        pass
    return {"model": model, "smt_active": smt, "logical_cpus": os.cpu_count()}


# ---- reference CPU implementation (oracle/_ref = the reference's own headers compiled for the host) ----
def ref_env(spec):
    from oracle import ref

    renv = ref.Env()
    for item in spec:
        if item[0] == "sphere":
            renv.add_sphere(*[float(np.float32(v)) for v in item[1:5]])
        elif item[0] == "cuboid":
            renv.add_cuboid(item[1])
        elif item[0] == "capsule":
            renv.add_capsule(item[1])
    return ref, renv


def time_reference(spec, cfg: np.ndarray, threads: int, packed: bool, passes: int = 1) -> float:
    """configs/s of the reference's fkcc on `threads` host threads over `passes` passes of `cfg`."""
    ref, renv = ref_env(spec)
    fn = ref.fkcc_packed if packed else ref.fkcc
    fn("panda_0_0", renv, cfg[: 8 * 1024], threads)  # touch code + per-thread environment copies
    t0 = time.perf_counter()
    for _ in range(passes):
        fn("panda_0_0", renv, cfg, threads)
    return passes * cfg.shape[0] / (time.perf_counter() - t0)


def auto_sample(spec, threads: int, seconds: float) -> int:
    probe = host_configs(8 * 16384 * max(1, threads // 8), 7)
    rate = time_reference(spec, probe, threads, True)
    n = int(rate * seconds)
    return max(8 * 1024, n // (8 * threads) * 8 * threads)


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    spec, _ = bench_scene()
    threads = os.cpu_count() or 1
    n = args.ref_sample or min(auto_sample(spec, threads, 3.0), 200_000_000)
    cfg = host_configs(n, 42)
    ref, renv = ref_env(spec)
    for _ in range(args.warmup):
        ref.fkcc_packed("panda_0_0", renv, cfg[: max(8 * threads, n // 8 // 8 * 8)], threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.fkcc_packed("panda_0_0", renv, cfg, threads)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(args.configs),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference", "cpu": cpu_info(),
                         "sample": f"bounded sample of the workload: {n} configurations per step x {args.steps} steps, oracle/_ref (reference "
                                   f"headers, g++ -O3 AVX2+FMA), reference fkcc<8> with 8 distinct configurations per block (its native "
                                   f"throughput mode; one verdict per block), {threads} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ---- the B200 arm ----------------------------------------------------------------------------------
class Dev:
    """Caller-owned device buffers through the C ABI (no torch: the library is the only CUDA user of this process)."""

    def __init__(self, api, ctx):
        self.api, self.L, self.h, self.ptrs = api, api.lib(), ctx.handle, []

    def alloc(self, nbytes: int):
        p = C.c_void_p()
        self.api._check(self.L.vb_dev_alloc(self.h, 0, max(16, nbytes), C.byref(p)))
        self.ptrs.append(p)
        return p

    def upload(self, dst, arr: np.ndarray):
        self.api._check(self.L.vb_dev_upload(self.h, 0, dst, arr.ctypes.data_as(C.c_void_p), arr.nbytes))
        self.api._check(self.L.vb_sync(self.h))

    def download(self, src, dtype, count: int) -> np.ndarray:
        out = np.zeros(count, dtype=dtype)
        self.api._check(self.L.vb_dev_download(self.h, 0, out.ctypes.data_as(C.c_void_p), src, out.nbytes))
        return out

    def free(self, p):
        self.L.vb_dev_free(self.h, 0, p)
        self.ptrs = [q for q in self.ptrs if q is not p]


def fill_parallel(fn, n: int, threads: int):
    """fn(first, count) over [0, n) on `threads` host threads (ctypes releases the GIL)."""
    chunk = (n + threads - 1) // threads
    ts = [threading.Thread(target=fn, args=(lo, min(chunk, n - lo))) for lo in range(0, n, chunk)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()


def edge_work(api, L, h, dev, rid, eh, a: np.ndarray, b: np.ndarray, resolution: int, dim: int):
    """Mean algorithmic flops per edge: sum of W(config) over the interpolants the reference visits, in
    rake order until the first invalid pass (inclusive), + 2 dim per interpolant.  W(config) from vb_fkcc_work_d."""
    v = (b - a).astype(np.float32)
    dist = np.sqrt((v.astype(np.float32) ** 2).sum(axis=1, dtype=np.float32)).astype(np.float32)
    n_pass = np.maximum(np.ceil(dist / np.float32(8) * np.float32(resolution)), 1).astype(np.int64)
    rows, owner, pidx = [], [], []
    pct = (np.arange(1, 9, dtype=np.float32) / np.float32(8))[None, :, None]
    for e in range(a.shape[0]):
        back = v[e] / np.float32(8 * n_pass[e])
        x0 = a[e][None, None, :] + v[e][None, None, :] * pct  # [1][8][dim]
        p = np.arange(n_pass[e], dtype=np.float32)[:, None, None]
        rows.append((x0 - p * back[None, None, :]).reshape(-1, dim).astype(np.float32))
        owner.append(np.full(8 * n_pass[e], e))
        pidx.append(np.repeat(np.arange(n_pass[e]), 8))
    q = np.ascontiguousarray(np.concatenate(rows))
    owner, pidx = np.concatenate(owner), np.concatenate(pidx)
    m = q.shape[0]
    q_d, w_d, b_d = dev.alloc(q.nbytes), dev.alloc(4 * m), dev.alloc(4 * ((m + 31) // 32))
    dev.upload(q_d, q)
    api._check(L.vb_fkcc_work_d(h, 0, rid, None, eh, q_d, m, b_d, w_d))
    api._check(L.vb_sync(h))
    w = dev.download(w_d, np.float32, m).astype(np.float64)
    ok = api.unpack_bits(dev.download(b_d, np.uint32, (m + 31) // 32), m)
    for p in (q_d, w_d, b_d):
        dev.free(p)
    first_bad = np.full(a.shape[0], np.iinfo(np.int64).max)
    np.minimum.at(first_bad, owner[~ok], pidx[~ok])
    reached = pidx <= first_bad[owner]
    per_edge = np.bincount(owner[reached], weights=w[reached] + 2 * dim, minlength=a.shape[0])
    return float(per_edge.mean()), float(reached.sum() / a.shape[0])


def run_b200(args):
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
    from mr_vamp_b200 import api, scenes, sharding

    rank, local_rank, world = dist_env()
    L = api.lib()
    ctx = api.Context([local_rank])  # fails loudly when no CUDA device is usable: there is no CPU path
    h = ctx.handle
    id_path = f"/tmp/vb_nccl_id_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"
    comm = sharding.NcclComm(ctx, rank, world, id_path)
    dev = Dev(api, ctx)
    spec, _ = bench_scene()
    env = scenes.apply(spec, api.Environment(ctx))
    robot = api.Robot("panda", ctx=ctx)  # PandaBase<0,0,0>: MBM scenes are built around the origin
    rid, eh, dim = robot.robot_id, env.handle, robot.dimension
    n = args.configs
    host_threads = max(1, min(16, os.cpu_count() or 1))

    # inputs: configurations rank*n .. rank*n + n - 1 of the library's synthetic stream (uniform in the joint ranges,
    # seed 42), written into pinned host memory and uploaded: 2.8 GB resident in HBM at 1e8 (larger than L2)
    cfg_h = api.pinned_empty((n, dim), np.float32)
    fill_parallel(lambda lo, c: api._check(L.vb_random_configs(rid, 42, rank * n + lo, c, api._ptr(cfg_h[lo:lo + c]))), n, host_threads)
    cfg_d = dev.alloc(cfg_h.nbytes)
    dev.upload(cfg_d, cfg_h)
    n_words = (n + 31) // 32
    bits_d = dev.alloc(4 * n_words)
    gathered_d = dev.alloc(4 * n_words * world) if world > 1 else None

    def step_device():
        api._check(L.vb_fkcc_d(h, 0, rid, None, eh, cfg_d, n, bits_d, None))
        if world > 1:
            # the one exchange step of the path: all-gather of the verdict words (NCCL over NVLink, issued by the
            # library on its own stream right behind the kernel)
            api._check(L.vb_comm_allgather_bits_d(h, bits_d, gathered_d, n_words))

    def timed(fn, steps):
        ms = C.c_double()
        comm.barrier()
        api._check(L.vb_timer_start(h, 0))
        for _ in range(steps):
            fn()
        api._check(L.vb_timer_stop(h, 0, C.byref(ms)))
        comm.barrier()
        return comm.max_float(ms.value)  # CUDA events on the launch stream, max over ranks

    for _ in range(max(args.warmup, 3)):
        step_device()
    launches0 = ctx.launch_count()
    with ClockSampler(local_rank) as clocks:
        ms_total = timed(step_device, args.steps)
    launches = ctx.launch_count() - launches0
    value = world * n * args.steps / (ms_total * 1e-3)
    bits = dev.download(bits_d, np.uint32, n_words)
    gather_ok = None
    if world > 1:
        # every rank holds every rank's words; rank r's block must equal what rank r computed
        mine = dev.download(C.c_void_p(gathered_d.value + 4 * n_words * rank), np.uint32, n_words)
        gather_ok = bool(np.array_equal(mine, bits))
    ok_all = api.unpack_bits(bits, n)
    valid_fraction = float(ok_all.mean())

    # ---- end to end through the host-buffer C-ABI call (vb_fkcc): pinned host in, verdict words out
    bits_h = api.pinned_empty((n_words,), np.uint32)

    def step_host():
        api._check(L.vb_fkcc(h, rid, None, eh, api._ptr(cfg_h), n, api._ptr(bits_h, api._u32p), None))

    e2e_steps = max(2, min(args.steps, 5))
    step_host()
    ms_e2e = timed(step_host, e2e_steps)
    e2e_value = world * n * e2e_steps / (ms_e2e * 1e-3)
    same = bool(np.array_equal(bits_h, bits))

    # the same through the device-synthesised path (0 input bytes: what a sampler-on-device planner sends)
    def step_synth():
        api._check(L.vb_fkcc_random_d(h, 0, rid, None, eh, 42, rank * n, n, bits_d))
        if world > 1:
            api._check(L.vb_comm_allgather_bits_d(h, bits_d, gathered_d, n_words))

    step_synth()
    ms_synth = timed(step_synth, 3)
    synth_value = world * n * 3 / (ms_synth * 1e-3)
    synth_same = bool(np.array_equal(dev.download(bits_d, np.uint32, n_words), bits))

    # ---- algorithmic flops per configuration, measured on the device over a sample of the same inputs
    n_w = min(n, 1 << 20)
    work_d, wbits_d = dev.alloc(4 * n_w), dev.alloc(4 * ((n_w + 31) // 32))
    api._check(L.vb_fkcc_work_d(h, 0, rid, None, eh, cfg_d, n_w, wbits_d, work_d))
    api._check(L.vb_sync(h))
    w_mean = float(dev.download(work_d, np.float32, n_w).astype(np.float64).mean())
    fp32_measured = C.c_double()
    api._check(L.vb_fp32_peak_d(h, 0, C.byref(fp32_measured)))

    # ---- second half of the metric: edges validated / s (planning::validate_motion, RRTC-shaped edges 0.05..1 rad)
    n_edges = args.edges
    ea_h, eb_h = api.pinned_empty((n_edges, dim), np.float32), api.pinned_empty((n_edges, dim), np.float32)
    fill_parallel(lambda lo, c: api._check(L.vb_random_edges(rid, 42, rank * n_edges + lo, c, 0.05, 1.0, api._ptr(ea_h[lo:lo + c]),
                                                             api._ptr(eb_h[lo:lo + c]))), n_edges, host_threads)
    ea_d, eb_d = dev.alloc(ea_h.nbytes), dev.alloc(eb_h.nbytes)
    dev.upload(ea_d, ea_h)
    dev.upload(eb_d, eb_h)
    e_words = (n_edges + 31) // 32
    ebits_d = dev.alloc(4 * e_words)
    ebits_h = api.pinned_empty((e_words,), np.uint32)

    def step_edges():
        api._check(L.vb_validate_motion_d(h, 0, rid, None, eh, ea_d, eb_d, n_edges, 0, ebits_d, None))

    def step_edges_host():
        api._check(L.vb_validate_motion(h, rid, None, eh, api._ptr(ea_h), api._ptr(eb_h), n_edges, 0, api._ptr(ebits_h, api._u32p), None))

    step_edges()
    e_launch0 = ctx.launch_count()
    ms_edges = timed(step_edges, 5)
    e_launches = ctx.launch_count() - e_launch0
    edges_value = world * n_edges * 5 / (ms_edges * 1e-3)
    ebits = dev.download(ebits_d, np.uint32, e_words)
    step_edges_host()
    ms_edges_host = timed(step_edges_host, 3)
    edges_e2e = world * n_edges * 3 / (ms_edges_host * 1e-3)
    edges_same = bool(np.array_equal(ebits_h, ebits))
    edge_ok = api.unpack_bits(ebits, n_edges)
    n_we = min(n_edges, 8192)
    w_edge, interp_per_edge = edge_work(api, L, h, dev, rid, eh, np.array(ea_h[:n_we]), np.array(eb_h[:n_we]), robot.resolution, dim)

    if rank != 0:
        comm.barrier()
        comm.close()
        return

    ms_step = ms_total / args.steps
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peaks = json.load(fh)
    except Exception:
        pass
    sm_max_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = 148 * 128 * 2 * sm_max_mhz * 1e6 / 1e12  # TFLOP/s: SMs x FP32 lanes x FMA x clock
    peak_source = f"148 SMs x 128 FP32 lanes x 2 x {sm_max_mhz:.0f} MHz (nominal; MEASURED_PEAKS.json has no FP32 entry)"
    achieved = w_mean * n / (ms_step * 1e-3) / 1e12
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_achieved = (4.0 * dim + 0.125) * n / (ms_step * 1e-3) / 1e9
    edges_per_gpu_s = edges_value / world
    edges_achieved = w_edge * edges_per_gpu_s / 1e12

    # DRAM traffic and pipe utilisation of the fkcc kernel from the committed ncu --set full capture of this kernel
    # (NOT measured in this run: bytes per configuration there x the configurations of one launch here)
    traffic, ncu_note, fma_pipe = None, None, None
    for name in ("r2_panda_fkcc_mbm_traffic.json", "r1_panda_fkcc_mbm_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                prof = json.load(fh)
            traffic = float(prof["bytes_per_config"]) * n
            fma_pipe = prof.get("fma_pipe_pct")
            ncu_note = f"scaled from the committed capture profiles/{name} (bytes per configuration x configurations per launch)"
            break
        except Exception:
            continue

    # ---- parity of the TIMED stream against the oracle (outside the timed region): every 100th configuration,
    # every 64th edge; the band is judged by the oracle's clearance (oracle/clear)
    parity = edges_parity = None
    cpu = edges_cpu = None
    robots_extra = None
    threads = os.cpu_count() or 1
    try:
        ref_mod, renv = ref_env(spec)
        idx = np.arange(0, n, 100)
        want, clr = ref_mod.fkcc_clearance("panda_0_0", renv, np.ascontiguousarray(cfg_h[idx]), threads)
        diff = ok_all[idx] != want
        parity = {"checked": int(idx.size), "mismatch_hard": int((diff & (clr > 1e-5)).sum()), "within_band": int((diff & (clr <= 1e-5)).sum()),
                  "band_m": 1e-5, "judge": "oracle clearance (oracle/clear: reference walk with instrumented primitives)",
                  "sample": "every 100th configuration of the timed stream (rank 0's shard)"}
        eidx = np.arange(0, n_edges, 64)
        ewant, eclr = ref_mod.motion_clearance("panda_0_0", renv, np.ascontiguousarray(ea_h[eidx]), np.ascontiguousarray(eb_h[eidx]), None, threads)
        ediff = edge_ok[eidx] != ewant
        edges_parity = {"checked": int(eidx.size), "mismatch_hard": int((ediff & (eclr > 1e-5)).sum()),
                        "within_band": int((ediff & (eclr <= 1e-5)).sum()), "sample": "every 64th edge of the timed stream"}
    except Exception as exc:
        parity = {"checked": 0, "unavailable": str(exc)}

    # ---- CPU baseline on this box's host cores (bounded samples of the same workloads), N = 1 only
    if world == 1:
        try:
            n_cpu = auto_sample(spec, threads, 8.0)
            # bounded sample: at most the step's own configurations (already on the host), walked several times
            passes = max(1, -(-n_cpu // n))
            sample = np.asarray(cfg_h[: min(n_cpu, n) // (8 * threads) * (8 * threads)])
            packed = time_reference(spec, sample, threads, True, passes)
            n_b = max(8 * threads, min(n_cpu // 8, sample.shape[0]))
            bcast = time_reference(spec, sample[:n_b], threads, False)
            one = time_reference(spec, sample[: max(8, min(sample.shape[0], n_cpu // threads // 4) // 8 * 8)], 1, True)
            cpu = {"value": packed, "unit": UNIT, "cores": threads, "kind": "reference",
                   "sample": f"first {sample.shape[0]} configurations of the same stream x {passes} passes, reference fkcc<8> with 8 "
                             f"distinct configurations per block (one verdict per block), {threads} threads",
                   "per_config_verdicts_value": bcast, "one_thread": one, "cpu": cpu_info()}
            ref_mod, renv = ref_env(spec)
            n_ce = min(n_edges, 1 << 18)
            ca, cb = np.ascontiguousarray(ea_h[:n_ce]), np.ascontiguousarray(eb_h[:n_ce])
            ref_mod.validate_motion("panda_0_0", renv, ca[:4096], cb[:4096], threads)
            t0 = time.perf_counter()
            ev = ref_mod.validate_motion("panda_0_0", renv, ca, cb, threads)
            dt = time.perf_counter() - t0
            t0 = time.perf_counter()
            ref_mod.validate_motion("panda_0_0", renv, ca[: n_ce // 16], cb[: n_ce // 16], 1)
            dt1 = time.perf_counter() - t0
            edges_cpu = {"value": n_ce / dt, "unit": "edges/s", "cores": threads, "kind": "reference",
                         "sample": f"first {n_ce} edges of the same stream, reference validate_motion<8>, {threads} threads",
                         "one_thread": (n_ce // 16) / dt1, "verdicts_match_gpu": bool(np.array_equal(ev, edge_ok[:n_ce]))}
        except Exception as exc:  # the checker is optional for the number, never for the product
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {exc}"}
        if not args.no_robots:
            robots_extra = other_robots(api, L, ctx, dev, fp32_peak, threads)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": bench_config(n),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 4 * dim * n, "d2h_bytes_per_step": 4 * n_words,
                "steps": e2e_steps, "matches_device_path": same,
                "device_synthesised_inputs": {"value": synth_value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                                              "note": "vb_fkcc_random_d: configurations generated in registers from (seed, index)",
                                              "matches_device_path": synth_same}},
        "gpu_launches": int(launches),
        "clocks": clocks.summary(),
        "parity": parity,
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                     "frac_is": "ALGORITHMIC flops of the reference walk (W, SURVEY.md 8(d)) per second over the FP32 peak; the kernel "
                                "skips part of that work (distance grid, joint-pair tables), so this is not pipe utilisation",
                     "fma_pipe_util_pct_ncu": fma_pipe,
                     "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)", "traffic_source": ncu_note,
                     "flops_per_config": w_mean, "peak_source": peak_source,
                     "peak_measured": fp32_measured.value, "frac_measured": achieved / fp32_measured.value if fp32_measured.value else None,
                     "peak_measured_source": "vb_fp32_peak_d: 8 independent FMA chains per thread, 8 x 256-thread CTAs per SM, best of 3, same run",
                     "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                             "bytes_per_config": 4.0 * dim + 0.125}},
        "cpu_baseline": cpu,
        "edges": {"metric": "panda_edges_validated_per_s", "value": edges_value, "unit": "edges/s", "edges_per_gpu_per_step": n_edges,
                  "ms_per_step": ms_edges / 5, "gpu_launches": int(e_launches), "edge_lengths_rad": [0.05, 1.0],
                  "valid_fraction": float(edge_ok.mean()),
                  "l2": f"inputs {8 * dim * n_edges / 1e6:.0f} MB per step" + (", larger than the 126 MB L2" if 8 * dim * n_edges > 126e6 else ""),
                  "e2e": {"value": edges_e2e, "unit": "edges/s", "h2d_bytes_per_step": 8 * dim * n_edges, "d2h_bytes_per_step": 4 * e_words,
                          "matches_device_path": edges_same},
                  "roofline": {"bound": "fp32", "achieved": edges_achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                               "frac": edges_achieved / fp32_peak, "flops_per_edge": w_edge, "interpolants_reached_per_edge": interp_per_edge,
                               "definition": "W(edge) = sum of W(config) over the interpolants the reference visits in rake order until the "
                                             "first invalid pass + 2 dim per interpolant (SURVEY.md 8(d)), on a sample of 8192 edges",
                               "peak_source": peak_source},
                  "parity": edges_parity, "cpu_baseline": edges_cpu},
        "extra": {"valid_fraction": valid_fraction, "gather_verified": gather_ok, "robots": robots_extra},
    }
    print(json.dumps(line), flush=True)
    comm.barrier()
    comm.close()
    try:
        os.remove(id_path)
    except OSError:
        pass


def other_robots(api, L, ctx, dev, fp32_peak, threads):
    """BASELINE.json configs[3]: UR5 / Fetch / Baxter fkcc on scene 1 of their MotionBenchMaker sets, device-resident
    library stream; algorithmic flops, roofline fraction and the reference CPU beside each."""
    from mr_vamp_b200 import scenes

    out = {}
    h = ctx.handle
    for name in ("ur5", "fetch", "baxter"):
        try:
            robot = api.Robot(name, ctx=ctx)
            mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{name}.json.gz"))
            probs = mbm["problems"]
            key = "table_pick" if "table_pick" in probs else next(iter(probs))
            spec = scenes.problem_to_spec(probs[key][0])
            env = scenes.apply(spec, api.Environment(ctx))
            n = 1 << 24
            q_d, b_d = dev.alloc(4 * robot.dimension * n), dev.alloc(n // 8)
            api._check(L.vb_random_configs_d(h, 0, robot.robot_id, 7, 0, n, q_d))
            ms = C.c_double()
            for rep in range(4):
                if rep == 1:
                    api._check(L.vb_sync(h))
                    api._check(L.vb_timer_start(h, 0))
                api._check(L.vb_fkcc_d(h, 0, robot.robot_id, None, env.handle, q_d, n, b_d, None))
            api._check(L.vb_timer_stop(h, 0, C.byref(ms)))
            rate = 3 * n / (ms.value * 1e-3)
            ok = api.unpack_bits(dev.download(b_d, np.uint32, n // 32), n)
            n_w = 1 << 18
            w_d = dev.alloc(4 * n_w)
            api._check(L.vb_fkcc_work_d(h, 0, robot.robot_id, None, env.handle, q_d, n_w, b_d, w_d))
            api._check(L.vb_sync(h))
            w_mean = float(dev.download(w_d, np.float32, n_w).astype(np.float64).mean())
            rec = {"scene": f"MotionBenchMaker {name} {key} scene 1", "configs_per_s": rate, "configs": n, "valid_fraction": float(ok.mean()),
                   "flops_per_config": w_mean, "roofline_frac": w_mean * rate / 1e12 / fp32_peak}
            try:
                ref_mod, renv = ref_env(spec)
                q = robot.random_configs(7, 0, 1 << 17)
                want, clr = ref_mod.fkcc_clearance(name, renv, q, threads)
                diff = ok[: q.shape[0]] != want
                rec["parity"] = {"checked": int(q.shape[0]), "mismatch_hard": int((diff & (clr > 1e-5)).sum()),
                                 "within_band": int((diff & (clr <= 1e-5)).sum())}
                big = np.ascontiguousarray(np.tile(q, (8, 1)))
                ref_mod.fkcc_packed(name, renv, big[: 8 * 1024 * threads], threads)
                t0 = time.perf_counter()
                ref_mod.fkcc_packed(name, renv, big, threads)
                rec["cpu_baseline"] = {"value": big.shape[0] / (time.perf_counter() - t0), "unit": UNIT, "cores": threads, "kind": "reference",
                                       "sample": f"{big.shape[0]} configurations, reference fkcc<8> packed blocks"}
            except Exception as exc:
                rec["cpu_baseline"] = {"value": None, "sample": f"unavailable: {exc}"}
            out[name] = rec
            for p in (q_d, b_d, w_d):
                dev.free(p)
        except Exception as exc:
            out[name] = {"error": str(exc)}
    return out


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
