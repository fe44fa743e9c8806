"""BASELINE.json configs beyond the bench line, as parity + measurement cases.
Each case checks verdicts against the reference on a sample and prints device throughput next to the
reference's CPU throughput on this box; the numbers are collected in gpurun_out/r2_configs.json (copied to
profiles/r2_configs.json after the box visit).  The band is judged by the oracle's clearance (conftest.judge).
These are NOT the bench line (bench.py measures configs[1])."""
import json
import os
import time

import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref, judge

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RESULTS = {}


def _cfgs(robot, seed, n):
    rng = np.random.default_rng(seed)
    return (robot.lower_bounds + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)


def _edges(robot, seed, n, lo=0.05, hi=1.0):
    a = _cfgs(robot, seed, n)
    rng = np.random.default_rng(seed + 1)
    d = rng.normal(size=a.shape).astype(np.float32)
    d *= (rng.uniform(lo, hi, (n, 1)) / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return a, np.clip(a + d, robot.lower_bounds, robot.lower_bounds + robot.ranges).astype(np.float32)


def _timed(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def _record(key, **kw):
    RESULTS[key] = kw
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "r2_configs.json"), "w") as fh:
        json.dump(RESULTS, fh, indent=1)
    print(key, json.dumps(kw))


def _mbm(robot_name):
    return scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, f"mbm_{robot_name}.json.gz"))


def test_m1_all_panda_mbm_problems(vb, ref):
    """configs[0]: every MotionBenchMaker Panda problem (7 scenarios x 100): start/goal verdicts, the straight
    start->goal edge, 2000 random configurations and 500 RRTC-shaped edges per scene; verdicts must match the
    reference on ALL of them outside the 1e-5 m band (north_star)."""
    robot = vb.Robot("panda")
    mbm = _mbm("panda")
    n_q = n_e = hard = inband = 0
    per_scenario = {}
    for name, problems in mbm["problems"].items():
        got, got0, want, clr, e_got, e_want, e_clr = [], [], [], [], [], [], []
        for p in problems:
            spec = scenes.problem_to_spec(p)
            env = scenes.apply(spec, vb.Environment())
            renv = apply_ref(spec, ref.Env())
            q = np.concatenate([np.asarray([p["start"]] + p["goals"], np.float32), _cfgs(robot, p["index"], 2000)])
            got.append(robot.fkcc_batch(q, env, margin=True)[0])
            got0.append(robot.fkcc_batch(q, env))
            w, c = ref.fkcc_clearance("panda_0_0", renv, q, threads=8)
            want.append(w)
            clr.append(c)
            a, b = _edges(robot, 1000 + p["index"], 500)
            a = np.concatenate([np.asarray([p["start"]], np.float32), a])
            b = np.concatenate([np.asarray([p["goals"][0]], np.float32), b])
            e_got.append(robot.validate_motion_batch(a, b, env))
            w, c = ref.motion_clearance("panda_0_0", renv, a, b, threads=8)
            e_want.append(w)
            e_clr.append(c)
        want, clr, e_want, e_clr = (np.concatenate(v) for v in (want, clr, e_want, e_clr))
        recs = [judge(f"M1 {name} fkcc lane-serial", np.concatenate(got), want, clr),
                judge(f"M1 {name} fkcc warp walk", np.concatenate(got0), want, clr),
                judge(f"M1 {name} edges", np.concatenate(e_got), e_want, e_clr)]
        per_scenario[name] = {"configurations": int(want.size), "edges": int(e_want.size),
                              "within_band": sum(r["within_band"] for r in recs),
                              "mismatch_hard": sum(r["mismatch_hard"] for r in recs)}
        hard += per_scenario[name]["mismatch_hard"]
        inband += per_scenario[name]["within_band"]
        n_q += int(want.size)
        n_e += int(e_want.size)
    _record("m1_panda_mbm", problems=700, configurations=n_q, edges=n_e, hard_mismatches=hard, within_band=inband,
            band_judge="oracle clearance", scenarios=per_scenario)
    assert hard == 0


def _table_pick_cloud(vb, n_total=100000):
    """Surfaces of table_pick scene 1 sampled as src/vamp/pointcloud.py:29-126 does (8334 points per object)."""
    p = _mbm("panda")["problems"]["table_pick"][0]
    rng = np.random.default_rng(0)
    pts = []
    for o in p["box"]:
        c = vb.Cuboid(o["position"], o["orientation_euler_xyz"], o["half_extents"]).raw()
        ctr, axes, half = c[:3], c[3:12].reshape(3, 3), c[12:15]
        u = rng.uniform(-1, 1, (8334, 3)).astype(np.float32)
        face = rng.integers(0, 3, 8334)
        u[np.arange(8334), face] = np.sign(u[np.arange(8334), face])
        pts.append(ctr + (u * half) @ axes)
    for o in p["cylinder"]:
        c = vb.Cylinder(o["position"], o["orientation_euler_xyz"], o["radius"], o["length"]).raw()
        t = rng.uniform(0, 1, 8334)[:, None]
        ang = rng.uniform(0, 2 * np.pi, 8334)
        v = c[3:6] / np.linalg.norm(c[3:6])
        e1 = np.cross(v, [1.0, 0.0, 0.0])
        e1 /= np.linalg.norm(e1)
        e2 = np.cross(v, e1)
        pts.append(c[:3] + t * c[3:6] + c[6] * (np.cos(ang)[:, None] * e1 + np.sin(ang)[:, None] * e2))
    return np.concatenate(pts).astype(np.float32)[:n_total]


@pytest.mark.parametrize("cloud", ["unfiltered", "filtered"])
def test_m3_capt_100k_points_edge_validation(vb, ref, gpu_ctx, cloud):
    """configs[2]: Panda vs a CAPT built (by the reference) from 100k points sampled on the surfaces of table_pick
    scene 1; batched RRTC-shaped edge validation.  `unfiltered`: the raw 100k-point cloud (745 MB of affordance
    vectors: HBM resident); `filtered`: the reference's real pipeline, filter_pointcloud(min_dist 0.02, cull at 1.19 m)
    first (src/vamp/pointcloud.py:150-160, scripts/evaluate_mbm.py:24-26): a few thousand points, L2 resident."""
    import ctypes as C

    robot = vb.Robot("panda")
    pts = _table_pick_cloud(vb)
    n_raw = int(pts.shape[0])
    if cloud == "filtered":
        pts = ref.filter_pointcloud(pts, 0.02, 1.19, [0, 0, 0], [-1.19] * 3, [1.19] * 3, True)
    renv = ref.Env()
    build_ns = renv.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    c = renv.capt(0)
    env = vb.Environment()
    env.add_capt_buffers(c["tests"], c["aff_starts"], c["aff_x"], c["aff_y"], c["aff_z"], c["aabbs"], c["aabb_top"],
                         c["r_point"], c["nlog2"])
    n = 1_000_000  # 1e6 RRTC-shaped edges
    a, b = robot.random_edges(7, 0, n, 0.05, 1.0)
    ok = robot.validate_motion_batch(a, b, env)
    m = 4000
    want, clr = ref.motion_clearance("panda_0_0", renv, a[:m], b[:m], threads=8)
    rec = judge(f"M3 CAPT {cloud} cloud ({pts.shape[0]} points), edges", ok[:m], want, clr)
    hard = rec["mismatch_hard"]
    # device-resident rate (CUDA events on the launch stream) and the host-buffer rate
    L, h = vb.lib(), gpu_ctx.handle
    ptrs = [C.c_void_p() for _ in range(3)]
    for p, nbytes in zip(ptrs, (a.nbytes, b.nbytes, n // 8 + 64)):
        vb._check(L.vb_dev_alloc(h, 0, nbytes, C.byref(p)))
    vb._check(L.vb_dev_upload(h, 0, ptrs[0], a.ctypes.data_as(C.c_void_p), a.nbytes))
    vb._check(L.vb_dev_upload(h, 0, ptrs[1], b.ctypes.data_as(C.c_void_p), b.nbytes))
    ms, best = C.c_double(), 1e30
    for rep in range(4):
        vb._check(L.vb_timer_start(h, 0))
        vb._check(L.vb_validate_motion_d(h, 0, robot.robot_id, None, env.handle, ptrs[0], ptrs[1], n, 0, ptrs[2], None))
        vb._check(L.vb_timer_stop(h, 0, C.byref(ms)))
        best = min(best, ms.value) if rep else best
    for p in ptrs:
        L.vb_dev_free(h, 0, p)
    t_gpu = _timed(lambda: robot.validate_motion_batch(a, b, env))
    threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    ref.validate_motion("panda_0_0", renv, a[:20000], b[:20000], threads=threads)
    t_cpu = time.perf_counter() - t0
    _record(f"m3_capt_{cloud}_edges", points_sampled=n_raw, points_in_capt=int(pts.shape[0]),
            affordance_mb=round(c["aff_x"].nbytes * 3 / 1e6, 1), capt_build_ms=round(build_ns / 1e6, 1), edges=n,
            valid_fraction=float(ok.mean()), hard_mismatches=hard, within_band=rec["within_band"], parity_sample=m,
            gpu_edges_per_s_device=n / (best * 1e-3), gpu_edges_per_s_host_buffers=n / t_gpu, cpu_edges_per_s=20000 / t_cpu,
            cpu_threads=threads)
    assert hard == 0


@pytest.mark.parametrize("name,problem", [("ur5", "table_pick"), ("fetch", "table_pick"), ("baxter", None)])
def test_m4_other_robots_sweep(vb, ref, name, problem):
    """configs[3]: UR5 / Fetch / Baxter fkcc + motion validation over batch sizes 1e3..1e6 (host buffers)."""
    robot = vb.Robot(name)
    mbm = _mbm(name)
    probs = mbm["problems"][problem] if problem else next(iter(mbm["problems"].values()))
    spec = scenes.problem_to_spec(probs[0])
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    q = _cfgs(robot, 3, 1_000_000)
    want, clr = ref.fkcc_clearance(name, renv, q[:20000], threads=8)
    rec_q = judge(f"M4 {name} fkcc", robot.fkcc_batch(q[:20000], env), want, clr)
    sweep = {}
    for n in (1_000, 10_000, 100_000, 1_000_000):
        sweep[str(n)] = n / _timed(lambda: robot.fkcc_batch(q[:n], env))
    a, b = _edges(robot, 4, 100_000, hi={"ur5": 1.5, "fetch": 1.0, "baxter": 0.5}[name])
    e_ok = robot.validate_motion_batch(a, b, env)
    e_want, e_clr = ref.motion_clearance(name, renv, a[:3000], b[:3000], threads=8)
    rec_e = judge(f"M4 {name} edges", e_ok[:3000], e_want, e_clr)
    t_e = _timed(lambda: robot.validate_motion_batch(a, b, env))
    threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    ref.fkcc_packed(name, renv, q[:800_000], threads=threads)
    cpu = 800_000 / (time.perf_counter() - t0)
    _record(f"m4_{name}", valid_fraction=float(want.mean()), hard_mismatches=rec_q["mismatch_hard"] + rec_e["mismatch_hard"],
            within_band=rec_q["within_band"] + rec_e["within_band"], fkcc_configs_per_s_host_buffers=sweep,
            edges_per_s_host_buffers=100_000 / t_e, edge_valid_fraction=float(e_ok.mean()),
            cpu_packed_configs_per_s=cpu, cpu_threads=threads)


def test_m5_two_pandas(vb, ref, gpu_ctx):
    """configs[4]: two Pandas 1 m apart, cage duplicated at both bases, composite validation (N = 1e7).
    One fused launch per call (pairs_fused_kernel); device-resident and host-buffer rates, and the algorithmic-flop
    fraction with W(composite) = W(q_a) + W(q_b) measured by the work-count kernel + the exhaustive 59 x 59 x 11
    robot-robot term the fork's plan describes (MR_DEV_PLAN.md:176-198) for composites whose robots are both valid."""
    import ctypes as C

    ra, rb = vb.robot("panda_0_0"), vb.robot("panda_1_0")
    spec = scenes.cage(offset=ra.base) + scenes.cage(offset=rb.base)
    env = scenes.apply(spec, vb.Environment())
    L, h = vb.lib(), gpu_ctx.handle
    n = 10_000_000
    ptrs = [C.c_void_p() for _ in range(4)]
    for p, nbytes in zip(ptrs, (28 * n, 28 * n, n // 8 + 64, 4 * (1 << 20))):
        vb._check(L.vb_dev_alloc(h, 0, nbytes, C.byref(p)))
    qa_d, qb_d, bits_d, work_d = ptrs
    vb._check(L.vb_random_configs_d(h, 0, ra.robot_id, 61, 0, n, qa_d))
    vb._check(L.vb_random_configs_d(h, 0, rb.robot_id, 62, 0, n, qb_d))
    ms = C.c_double()
    best = 1e30
    for rep in range(4):
        vb._check(L.vb_timer_start(h, 0))
        vb._check(L.vb_validate_pairs_d(h, 0, ra.robot_id, vb._ptr(ra.base), vb._ptr(rb.base), env.handle, qa_d, qb_d, n, bits_d))
        vb._check(L.vb_timer_stop(h, 0, C.byref(ms)))
        best = min(best, ms.value) if rep else best
    valid = C.c_uint64()
    vb._check(L.vb_popcount_d(h, 0, bits_d, n, C.byref(valid)))
    # algorithmic flops on the first 2^20 composites
    m = 1 << 20
    w = np.zeros(m, np.float32)
    w_sum, both = 0.0, None
    for robot, q_d in ((ra, qa_d), (rb, qb_d)):
        vb._check(L.vb_fkcc_work_d(h, 0, robot.robot_id, vb._ptr(robot.base), env.handle, q_d, m, bits_d, work_d))
        vb._check(L.vb_dev_download(h, 0, w.ctypes.data_as(C.c_void_p), work_d, w.nbytes))
        words = np.zeros(m // 32, np.uint32)
        vb._check(L.vb_dev_download(h, 0, words.ctypes.data_as(C.c_void_p), bits_d, words.nbytes))
        ok = vb.unpack_bits(words, m)
        both = ok if both is None else (both & ok)
        w_sum += float(w.astype(np.float64).mean())
    w_comp = w_sum + float(both.mean()) * 59 * 59 * 11
    rate = n / (best * 1e-3)
    # host buffers, and the same answer
    k = 1_000_000
    qa, qb = ra.random_configs(61, 0, k), rb.random_configs(62, 0, k)
    ok_h = vb.validate_pairs_batch(ra, qa, rb, qb, env)
    t = _timed(lambda: vb.validate_pairs_batch(ra, qa, rb, qb, env), reps=2)
    for p in ptrs:
        L.vb_dev_free(h, 0, p)
    _record("m5_two_pandas", composites=n, valid_fraction=valid.value / n, composites_per_s_device=rate, launches_per_call=1,
            flops_per_composite=w_comp, fp32_fraction=w_comp * rate / 74.45e12, composites_per_s_host_buffers=k / t,
            host_valid_fraction=float(ok_h.mean()),
            parity="unpinned (no reference implementation); checked against a CPU restatement in tests/test_pairs_gpu.py")
    assert abs(ok_h.mean() - valid.value / n) < 2e-3
