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
    ap.add_argument("--edges", type=int, default=1 << 20, help="edges per GPU for the edges/s side metric")
    ap.add_argument("--ref-sample", type=int, default=0, help="reference arm: configurations per step (0 = auto)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def bench_scene():
    from mr_vamp_b200 import scenes

    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, "mbm_panda.json.gz"))
    problem = mbm["problems"]["table_pick"][0]
    return scenes.problem_to_spec(problem), problem


def host_configs(n: int, seed: int) -> np.ndarray:
    """Uniform configurations in the Panda joint ranges (scale_configuration, robots/panda/fk.hh:14-37)."""
    lo = np.array([-2.9671, -1.8326, -2.9671, -3.1416, -2.9671, -0.0873, -2.9671], dtype=np.float32)
    rg = np.array([5.9342, 3.6652, 5.9342, 3.2289, 5.9342, 3.9096, 5.9342], dtype=np.float32)
    rng = np.random.default_rng(seed)
    return (lo + rng.random((n, 7), dtype=np.float32) * rg).astype(np.float32)


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
        "config": {"workload": WORKLOAD, "configs_per_step": n, "robot": "PandaBase<0,0,0>",
                   "mode": "reference fkcc<8>, 8 distinct configurations per block (its native throughput mode; one verdict per block)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
                         "sample": f"{n} configurations per step x {args.steps} steps, oracle/_ref (reference headers, g++ -O3 AVX2+FMA)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ---- the B200 arm ----------------------------------------------------------------------------------
def run_b200(args):
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
    import torch
    import torch.distributed as dist

    from mr_vamp_b200 import api, scenes

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the fkcc path has no CPU implementation")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = api.lib()
    ctx = api.Context([local_rank])
    h = ctx.handle
    spec, _ = bench_scene()
    env = scenes.apply(spec, api.Environment(ctx))
    robot = api.Robot("panda", ctx=ctx)  # PandaBase<0,0,0>: MBM scenes are built around the origin
    rid, eh = robot.robot_id, env.handle
    n = args.configs
    dev = torch.device("cuda", local_rank)

    # inputs: uniform configurations in the joint ranges, resident in HBM (2.8 GB at 1e8: larger than L2)
    gen = torch.Generator(device=dev).manual_seed(42 + rank)
    lo = torch.tensor(robot.lower_bounds, device=dev)
    rg = torch.tensor(robot.ranges, device=dev)
    cfg_d = torch.rand((n, 7), device=dev, generator=gen, dtype=torch.float32).mul_(rg).add_(lo)
    bits_d = torch.zeros((n + 31) // 32, device=dev, dtype=torch.int32)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        api._check(L.vb_sync(h))

    # N > 1: the one exchange step of the path -- all-gather of the verdict words (NCCL over NVLink, issued
    # by the library on its own stream right behind the kernel)
    n_words = (n + 31) // 32
    gathered_d = None
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = (C.c_ubyte * 128)()
            api._check(L.vb_comm_unique_id(buf))
            uid = torch.tensor(list(buf), dtype=torch.uint8)
        uid = uid.to(dev)
        dist.broadcast(uid, src=0)
        raw = (C.c_ubyte * 128)(*uid.cpu().tolist())
        api._check(L.vb_comm_init(h, world, rank, raw))
        gathered_d = torch.zeros(n_words * world, device=dev, dtype=torch.int32)

    def step_device():
        api._check(L.vb_fkcc_d(h, 0, rid, None, eh, cfg_d.data_ptr(), n, bits_d.data_ptr(), None))
        if world > 1:
            api._check(L.vb_comm_allgather_bits_d(h, bits_d.data_ptr(), gathered_d.data_ptr(), n_words))

    def timed(fn, steps):
        ms = C.c_double()
        barrier()
        api._check(L.vb_timer_start(h, 0))
        for _ in range(steps):
            fn()
        api._check(L.vb_timer_stop(h, 0, C.byref(ms)))
        barrier()
        t = torch.tensor([ms.value], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        step_device()
    launches0 = ctx.launch_count()
    with ClockSampler(local_rank) as clocks:
        ms_total = timed(step_device, args.steps)
    launches = ctx.launch_count() - launches0
    value = world * n * args.steps / (ms_total * 1e-3)
    gather_ok = None
    if world > 1:
        # every rank holds every rank's words; rank r's block must equal what rank r computed
        mine = gathered_d[rank * n_words:(rank + 1) * n_words]
        gather_ok = bool(torch.equal(mine, bits_d))
    valid = C.c_uint64()
    api._check(L.vb_popcount_d(h, 0, bits_d.data_ptr(), n, C.byref(valid)))

    # ---- end to end through the host-buffer C-ABI call (vb_fkcc): pinned host in, verdict words out
    n_e2e = n
    cfg_h = torch.empty((n_e2e, 7), dtype=torch.float32, pin_memory=True)
    cfg_h.copy_(cfg_d[:n_e2e])
    bits_h = torch.zeros((n_e2e + 31) // 32, dtype=torch.int32, pin_memory=True)
    torch.cuda.synchronize()

    def step_host():
        api._check(L.vb_fkcc(h, rid, None, eh, C.cast(cfg_h.data_ptr(), api._f32p), n_e2e,
                             C.cast(bits_h.data_ptr(), api._u32p), None))

    e2e_steps = max(2, min(args.steps, 5))
    step_host()
    ms_e2e = timed(step_host, e2e_steps)
    e2e_value = world * n_e2e * e2e_steps / (ms_e2e * 1e-3)
    same = bool(torch.equal(bits_h, bits_d[: bits_h.numel()].cpu()))

    # ---- algorithmic flops per configuration, measured on the device over a sample of the same inputs
    n_w = min(n, 1 << 20)
    work_d = torch.zeros(n_w, device=dev, dtype=torch.float32)
    wbits_d = torch.zeros((n_w + 31) // 32, device=dev, dtype=torch.int32)
    api._check(L.vb_fkcc_work_d(h, 0, rid, None, eh, cfg_d.data_ptr(), n_w, wbits_d.data_ptr(), work_d.data_ptr()))
    api._check(L.vb_sync(h))
    w_mean = float(work_d.double().mean().item())
    fp32_measured = C.c_double()
    api._check(L.vb_fp32_peak_d(h, 0, C.byref(fp32_measured)))

    # ---- side metric: edges validated / s (planning::validate_motion, RRTC-shaped edges 0.05..1 rad)
    n_edges = args.edges
    ea = cfg_d[:n_edges].contiguous()
    d = torch.randn((n_edges, 7), device=dev, generator=gen)
    d = d / d.norm(dim=1, keepdim=True) * (0.05 + 0.95 * torch.rand((n_edges, 1), device=dev, generator=gen))
    eb = torch.minimum(torch.maximum(ea + d, lo), lo + rg).contiguous()
    ebits = torch.zeros((n_edges + 31) // 32, device=dev, dtype=torch.int32)
    torch.cuda.synchronize()

    def step_edges():
        api._check(L.vb_validate_motion_d(h, 0, rid, None, eh, ea.data_ptr(), eb.data_ptr(), n_edges, 0,
                                          ebits.data_ptr(), None))

    step_edges()
    ms_edges = timed(step_edges, 3)
    edges_value = world * n_edges * 3 / (ms_edges * 1e-3)
    ecount = C.c_uint64()
    api._check(L.vb_popcount_d(h, 0, ebits.data_ptr(), n_edges, C.byref(ecount)))
    edge_valid = ecount.value / n_edges

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
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
    achieved = w_mean * n / (ms_step * 1e-3) / 1e12
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_achieved = (28.0 + 0.125) * n / (ms_step * 1e-3) / 1e9

    # DRAM traffic of the fkcc kernel per launch, from the committed ncu --set full capture (bytes per configuration
    # there x the configurations of one launch here); None when the profile is absent
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r1_panda_fkcc_mbm_traffic.json")) as fh:
            traffic = float(json.load(fh)["bytes_per_config"]) * n
    except Exception:
        pass

    # ---- CPU baseline on this box's host cores (bounded sample of the same workload)
    cpu = None
    if world == 1:
        try:
            threads = os.cpu_count() or 1
            n_cpu = auto_sample(spec, threads, 8.0)
            # bounded sample: at most the step's own configurations (already on the host), walked several times
            passes = max(1, -(-n_cpu // n_e2e))
            sample = cfg_h[: min(n_cpu, n_e2e) // (8 * threads) * (8 * threads)].numpy()
            packed = time_reference(spec, sample, threads, True, passes)
            n_b = max(8 * threads, min(n_cpu // 8, sample.shape[0]))
            bcast = time_reference(spec, sample[:n_b], threads, False)
            try:
                ref_mod, renv = ref_env(spec)
                n_ce = min(n_edges, 1 << 17)
                ea_h, eb_h = ea[:n_ce].cpu().numpy(), eb[:n_ce].cpu().numpy()
                ref_mod.validate_motion("panda_0_0", renv, ea_h[:4096], eb_h[:4096], threads)
                t0 = time.perf_counter()
                ev = ref_mod.validate_motion("panda_0_0", renv, ea_h, eb_h, threads)
                cpu_edges = n_ce / (time.perf_counter() - t0)
                edge_match = bool(np.array_equal(ev, api.unpack_bits(ebits[: (n_ce + 31) // 32].cpu().numpy().view(np.uint32), n_ce)))
            except Exception:
                cpu_edges, edge_match = None, None
            cpu = {"value": packed, "unit": UNIT, "cores": threads, "kind": "reference",
                   "sample": f"first {sample.shape[0]} configurations of the same stream x {passes} passes, reference fkcc<8> with 8 "
                             f"distinct configurations per block (one verdict per block), {threads} threads",
                   "per_config_verdicts_value": bcast, "edges_per_s": cpu_edges,
                   "edge_verdicts_match_gpu": edge_match}
        except Exception as exc:  # the checker is optional for the number, never for the product
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {exc}"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "configs_per_gpu_per_step": n, "robot": "PandaBase<0,0,0>",
                   "environment": "MotionBenchMaker panda table_pick scene 1 (10 z-aligned cuboids + 2 z-aligned capsules)",
                   "l2": f"inputs {28 * n / 1e9:.2f} GB per step, larger than the 126 MB L2", "valid_fraction": valid.value / n,
                   "sharding": "configurations sharded by rank (contiguous ranges), environment + grid replicated per GPU; "
                               "one collective: ncclAllGather of the verdict words inside the timed step (N > 1)",
                   "gather_verified": gather_ok},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 28 * n_e2e, "d2h_bytes_per_step": 4 * ((n_e2e + 31) // 32),
                "steps": e2e_steps, "matches_device_path": same},
        "gpu_launches": int(launches),
        "clocks": clocks.summary(),
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                     "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write; HBM-side, the path is FP32-bound)",
                     "flops_per_config": w_mean,
                     "peak_source": f"148 SMs x 128 FP32 lanes x 2 x {sm_max_mhz:.0f} MHz (nominal; MEASURED_PEAKS.json has no FP32 entry)",
                     "peak_measured": fp32_measured.value, "frac_measured": achieved / fp32_measured.value if fp32_measured.value else None,
                     "peak_measured_source": "vb_fp32_peak_d: 8 independent FMA chains per thread, 8 x 256-thread CTAs per SM, best of 3, same run",
                     "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                             "bytes_per_config": 28.125}},
        "cpu_baseline": cpu,
        "extra": {"edges_per_s": edges_value, "edges_per_gpu": n_edges, "edge_lengths_rad": [0.05, 1.0],
                  "edge_valid_fraction": edge_valid},
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
