"""GPU parity of the fkcc path against the reference compiled as an oracle (oracle/_ref).

Bars (BASELINE.json north_star): sphere centres within 1e-5 m; verdicts identical except where the
decisive clearance is within 1e-5 m of contact (those are counted and reported, not hidden).
"""
import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref, judge, BAND

pytestmark = pytest.mark.gpu

ROBOTS = ["sphere", "ur5", "panda", "fetch", "baxter"]
REF_NAME = {"panda": "panda_0_0"}


def _cfgs(robot, seed, n):
    rng = np.random.default_rng(seed)
    return (robot.lower_bounds + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)


@pytest.mark.parametrize("name", ROBOTS)
def test_sphere_fk_centres(vb, ref, name):
    robot = vb.Robot(name)
    q = _cfgs(robot, 1, 4096)
    got = robot.fk_batch(q)
    want = ref.sphere_fk(REF_NAME.get(name, name), q)
    err = np.abs(got[..., :3] - want[..., :3]).max()
    print(f"{name}: max centre error {err:.3e} m")
    assert err <= BAND
    assert np.array_equal(got[..., 3], want[..., 3])


@pytest.mark.parametrize("base", ["panda_0_0", "panda_1_0", "panda_2_2"])
def test_panda_base_variants(vb, ref, base):
    robot = vb.robot(base)
    q = _cfgs(robot, 2, 1024)
    err = np.abs(robot.fk_batch(q) - ref.sphere_fk(base, q)).max()
    assert err <= BAND
    off = robot.base
    spec = scenes.cage(offset=off)
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    ok, mg = robot.fkcc_batch(q, env, margin=True)
    want, clr = ref.fkcc_clearance(base, renv, q)
    judge(f"{base} cage", ok, want, clr, mg)


@pytest.mark.parametrize("name", ROBOTS)
@pytest.mark.parametrize("scene", ["empty", "cage", "clutter"])
def test_fkcc_verdicts(vb, ref, name, scene):
    robot = vb.Robot(name)
    spec = {"empty": [], "cage": scenes.cage(), "clutter": scenes.random_scene(7)}[scene]
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    q = _cfgs(robot, 3, 20000)
    ok, mg = robot.fkcc_batch(q, env, margin=True)
    ok2 = robot.fkcc_batch(q, env)
    want, clr = ref.fkcc_clearance(REF_NAME.get(name, name), renv, q, threads=8)
    judge(f"{name}/{scene} lane-serial", ok, want, clr, mg)
    judge(f"{name}/{scene} warp walk", ok2, want, clr)


@pytest.mark.parametrize("name", ROBOTS)
def test_many_grid_stride_steps(vb, ref, name):
    """Several grid-stride steps per CTA (the batches above fit in one): the staged-configuration pipeline (TMA bulk copy
    one step ahead, vb_kernels.cuh) must hand every step ITS configurations.  A missing proxy fence between the reads of
    the staging buffer and the next bulk copy once let a warp walk the next step's configurations (non-deterministic,
    ~0.1 % of the steps with 2 CTAs per SM); the lane-serial kernel reads its configurations directly."""
    robot = vb.Robot(name)
    spec = scenes.random_scene(7) if name in ("sphere", "ur5", "panda") else []  # (the clutter sits inside Fetch's / Baxter's base)
    env = scenes.apply(spec, vb.Environment())
    n = 1_500_000 if name != "sphere" else 6_000_000
    q = _cfgs(robot, 17, n)
    want, clr = ref.fkcc_clearance(REF_NAME.get(name, name), apply_ref(spec, ref.Env()), q, threads=8)
    for rep in range(3):
        judge(f"{name}/many steps (pass {rep})", robot.fkcc_batch(q, env), want, clr)


def test_cage_known_answers(vb):
    # scripts/sphere_cage_example.py:10-13,67: start and goal are valid in the nominal cage
    robot = vb.Robot("panda")
    env = scenes.apply(scenes.cage(), vb.Environment())
    assert robot.validate(scenes.CAGE_START, env) and robot.validate(scenes.CAGE_GOAL, env)
    # scripts/cpp/test_panda_fk.cc:81-83,121-126: a sphere at the base of the origin robot collides with it
    # and not with the robot at (2, 2, 0), and vice versa (evaluated at the cage start configuration;
    # the all-zero configuration of that script is in self collision for every base)
    q = np.asarray(scenes.CAGE_START, dtype=np.float32)
    origin, offset = robot, vb.robot("panda_2_2")
    e2 = vb.Environment()
    e2.add_sphere(vb.Sphere([0.05, 0.05, 0.05], 0.1))
    assert not origin.fkcc_batch(q, e2)[0] and offset.fkcc_batch(q, e2)[0]
    e3 = vb.Environment()
    e3.add_sphere(vb.Sphere([2.05, 2.05, 0.05], 0.1))
    assert origin.fkcc_batch(q, e3)[0] and not offset.fkcc_batch(q, e3)[0]
    e4 = vb.Environment()
    e4.add_sphere(vb.Sphere([5.0, 5.0, 0.5], 0.3))
    assert origin.fkcc_batch(q, e4)[0] and offset.fkcc_batch(q, e4)[0]
    assert not origin.fkcc_batch(np.zeros(7, dtype=np.float32), e4)[0]  # self collision at q = 0


@pytest.mark.parametrize("name", ROBOTS)
@pytest.mark.parametrize("scene", ["empty", "cage", "clutter"])
def test_validate_motion(vb, ref, name, scene):
    robot = vb.Robot(name)
    spec = {"empty": [], "cage": scenes.cage(), "clutter": scenes.random_scene(11)}[scene]
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    n = 3000
    a = _cfgs(robot, 5, n)
    rng = np.random.default_rng(6)
    step = rng.normal(size=(n, robot.dimension)).astype(np.float32)
    step *= (rng.uniform(0.02, 1.5, (n, 1)) / np.linalg.norm(step, axis=1, keepdims=True)).astype(np.float32)
    b = np.clip(a + step, robot.lower_bounds, robot.lower_bounds + robot.ranges).astype(np.float32)
    b[: n // 10] = _cfgs(robot, 8, n // 10)  # some long (PRM-shaped) edges
    ok, mg = robot.validate_motion_batch(a, b, env, margin=True)
    ok2 = robot.validate_motion_batch(a, b, env)
    want, clr = ref.motion_clearance(REF_NAME.get(name, name), renv, a, b, threads=8)
    # (the margin output is not cross-checked on edges: the reference walks 8 interpolants as one block, whose joint
    # pruning and guard coupling execute tests a per-interpolant walk does not, and its approximate-sqrt pruning skips
    # obstacles the exact pruning visits: the two figures are minima over different sets of tests)
    judge(f"{name}/{scene} edges lane-serial", ok, want, clr)
    judge(f"{name}/{scene} edges warp walk", ok2, want, clr)
    # the (start, vector, distance) form the planners call (planning/validate.hh:23-65), with the distance as a
    # planner passes it (its own float, not recomputed from the vector)
    v = (b - a).astype(np.float32)
    d = (np.linalg.norm(v.astype(np.float64), axis=1) * rng.uniform(0.97, 1.03, n)).astype(np.float32)
    want_v, clr_v = ref.motion_clearance(REF_NAME.get(name, name), renv, a, v, d, threads=8)
    judge(f"{name}/{scene} validate_vector", robot.validate_vector_batch(a, v, d, env), want_v, clr_v)


def test_validate_joint_limits(vb, ref):
    robot = vb.Robot("panda")
    q = _cfgs(robot, 9, 2000)
    q[::3] += np.float32(0.5)  # push a third of them around/over the limits
    env = vb.Environment()
    got = robot.validate_batch(q, env)
    want = ref.validate("panda_0_0", ref.Env(), q)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name", ["panda", "ur5", "fetch"])
def test_empty_and_ragged_batches(vb, ref, name):
    """Batch sizes around the 32-configuration verdict words (0, 1, 31, 32, 33, 95 ...), zero-length and very long
    edges: same verdicts as the reference, and an empty batch is not an error."""
    robot = vb.Robot(name)
    rn = REF_NAME.get(name, name)
    spec = scenes.cage() if name == "panda" else scenes.random_scene(5)
    env, renv = scenes.apply(spec, vb.Environment()), apply_ref(spec, ref.Env())
    q_all = _cfgs(robot, 77, 200)
    want_all, clr_all = ref.fkcc_clearance(rn, renv, q_all)
    for n in (0, 1, 31, 32, 33, 95, 200):
        q = q_all[:n]
        got = robot.fkcc_batch(q, env)
        assert got.shape == (n,)
        if n:
            judge(f"{name} fkcc n={n}", got, want_all[:n], clr_all[:n])
            assert robot.fk_batch(q).shape == (n, robot.n_spheres, 4)
            assert np.array_equal(robot.validate_batch(q, env), ref.validate(rn, renv, q))
        else:
            assert robot.validate_batch(q, env).shape == (0,)
            assert robot.validate_motion_batch(q, q, env).shape == (0,)
            assert robot.fk_batch(q).shape[0] == 0
    # edges: zero length (one pass of 8 copies of the start), short, and long enough for hundreds of rake passes
    a = q_all[:66].copy()
    b = q_all[66:132].copy()
    b[:8] = a[:8]  # zero-length edges
    b[8:16] = (a[8:16] + np.float32(1e-4)).astype(np.float32)
    lo, hi = robot.lower_bounds, robot.lower_bounds + robot.ranges
    a[16:20], b[16:20] = lo, hi  # across the whole joint box: the longest edges inside the limits
    b[20:24] = (a[20:24] + np.float32(40.0)).astype(np.float32)  # far beyond the limits: > 400 passes
    for n in (1, 7, 33, 66):
        got = robot.validate_motion_batch(a[:n], b[:n], env)
        want, clr = ref.motion_clearance(rn, renv, a[:n], b[:n])
        judge(f"{name} edges n={n}", got, want, clr)


@pytest.mark.parametrize("name", ["panda", "ur5", "fetch", "baxter"])
def test_non_finite_and_far_out_of_range_inputs(vb, ref, name):
    """NaN / Inf / 1e30 joint values.  `validate` (joint limits && fkcc, bindings/common.hh:172-182) rejects all of them
    in the reference and here.  `fkcc` alone: a NaN INPUT makes every clearance NaN and the configuration "free" in
    both (the reference reads sign bits -- FloatVector::test_zero, vector/avx.hh:386-389 -- and the input NaN is
    positive; the kernels compare, and a NaN compares false).  An Inf input is where the two may part: x86 turns
    inf - inf into its NEGATIVE default NaN, whose sign bit the reference reads as a collision, while the GPU's
    canonical NaN is positive; such configurations are outside every joint range, `validate` never lets them reach
    fkcc, and DESIGN.md section 5 records the difference instead of emulating x86 NaN signs.  Edges with a non-finite
    endpoint are reported invalid (the reference's pass count is undefined for them and its loop does not end)."""
    robot = vb.Robot(name)
    rn = REF_NAME.get(name, name)
    env, renv = vb.Environment(), ref.Env()
    mid = (robot.lower_bounds + np.float32(0.5) * robot.ranges).astype(np.float32)
    q = np.tile(mid, (6, 1)).astype(np.float32)
    q[1, 0] = np.nan
    q[2, min(3, robot.dimension - 1)] = np.inf
    q[3, robot.dimension - 1] = -np.inf
    q[4, :] = np.nan
    q[5, 1] = np.float32(1e30)
    assert np.array_equal(robot.validate_batch(q, env), ref.validate(rn, renv, q))
    assert not robot.validate_batch(q, env)[1:].any()
    got, want = robot.fkcc_batch(q, env), ref.fkcc(rn, renv, q)
    assert got[0] == want[0]
    assert got[1] and want[1] and got[4] and want[4]  # NaN inputs: free in both
    a = np.tile(mid, (4, 1)).astype(np.float32)
    b = a.copy()
    b[1, 0] = np.nan
    b[2, 1] = np.inf
    a[3, 0] = -np.inf
    ok = robot.validate_motion_batch(a, b, env)
    assert ok[0] == ref.validate_motion(rn, renv, a[:1], b[:1])[0]
    assert not ok[1:].any()


@pytest.mark.parametrize("name", ["panda", "ur5", "fetch"])
@pytest.mark.parametrize("scene", ["65 primitives", "clipped grid"])
def test_large_scenes_exercise_the_grid_fallbacks(vb, ref, name, scene):
    """Branches of the acceleration structure the small scenes never reach: more than 31 obstacles (bit 31 of a candidate
    mask = 'some obstacle with index >= 31': vb_sdf.cuh next_candidate / sdf_build_kernel) and a scene that does not fit
    the +-4 m grid box (clipped grid, cover = 0, spheres outside the grid: grid_class with in == false)."""
    from mr_vamp_b200 import api as _api

    robot = vb.Robot(name)
    if scene == "65 primitives":
        spec = scenes.random_scene(23, n_spheres=15, n_cuboids=15, n_zcuboids=15, n_capsules=10, n_zcapsules=10, reach=1.1)
        # thin the clutter around the base so that a fair share of the configurations stays valid
        spec = [it for it in spec if it[0] != "sphere" or np.hypot(it[1], it[2]) > 0.45 or it[3] > 0.9]
    else:
        floor = _api.Cuboid([0.0, 0.0, -0.3], [0.0, 0.0, 0.3], [6.0, 6.0, 0.05]).raw()  # half extent 6 m > the 4 m box
        far = _api.Cuboid([5.0, 0.5, 0.5], [0.3, 0.2, 0.1], [0.4, 0.4, 0.4]).raw()
        spec = [("cuboid", floor), ("cuboid", far), ("sphere", 0.5, 0.3, 0.6, 0.15), ("sphere", -0.4, 0.5, 0.9, 0.2),
                ("capsule", _api.Cylinder([0.6, -0.4, 0.4], [0.0, 0.0, 0.0], 0.06, 0.5).raw())]
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    rn = REF_NAME.get(name, name)
    q = _cfgs(robot, 33, 30000)
    want, clr = ref.fkcc_clearance(rn, renv, q, threads=8)
    ok, mg = robot.fkcc_batch(q, env, margin=True)
    judge(f"{name}/{scene} lane-serial", ok, want, clr)
    judge(f"{name}/{scene} warp walk", robot.fkcc_batch(q, env), want, clr)
    a, b = q[:3000], np.clip(q[:3000] + np.float32(0.2) * np.sign(q[3000:6000]), robot.lower_bounds, robot.lower_bounds + robot.ranges)
    e_want, e_clr = ref.motion_clearance(rn, renv, a, b, threads=8)
    judge(f"{name}/{scene} edges", robot.validate_motion_batch(a, b, env), e_want, e_clr)
    assert name == "fetch" or 0.02 < want.mean() < 0.98
