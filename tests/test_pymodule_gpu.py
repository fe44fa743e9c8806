"""The compiled Python module mr_vamp_b200._core_ext (pybind11 over the C ABI) on the surface of the reference's
nanobind module vamp._core_ext (bindings/common.hh:685-710, bindings/environment.cc:107-163): single-query calls with the
reference's names, the batch calls, and sphere_validity (collision/validity.hh:152-249) against the compiled reference."""
import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref, judge

pytestmark = pytest.mark.gpu


def _module():
    import mr_vamp_b200._core_ext as m

    return m


def _env(m, spec):
    env = m.Environment()
    for i, item in enumerate(spec):
        if item[0] == "sphere":
            s = m.Sphere([float(v) for v in item[1:4]], float(item[4]))
            s.name = str(i)
            env.add_sphere(s)
    return env


def test_module_single_queries_and_batches_match_the_reference(ref):
    m = _module()
    spec = scenes.cage()
    env = _env(m, spec)
    renv = apply_ref(spec, ref.Env())
    p = m.panda_0_0
    assert (p.dimension(), p.n_spheres(), p.resolution()) == ref.robot_info("panda_0_0")
    # scripts/sphere_cage_example.py:10-13,67
    assert p.validate(scenes.CAGE_START, env) and p.validate(scenes.CAGE_GOAL, env)
    assert p.validate(np.asarray(scenes.CAGE_START, np.float32))  # default: empty environment
    rng = np.random.default_rng(3)
    lo, hi = np.asarray(p.lower_bounds(), np.float32), np.asarray(p.upper_bounds(), np.float32)
    q = (lo + rng.random((5000, 7), dtype=np.float32) * (hi - lo)).astype(np.float32)
    want, clr = ref.fkcc_clearance("panda_0_0", renv, q, threads=8)
    judge("pymodule panda fkcc_batch", p.fkcc_batch(q, env), want, clr)
    assert np.array_equal(p.validate_batch(q, env), ref.validate("panda_0_0", renv, q)) or \
        not ((p.validate_batch(q, env) != ref.validate("panda_0_0", renv, q)) & (clr > 1e-5)).any()
    a, b = q[:1000], q[1000:2000]
    e_want, e_clr = ref.motion_clearance("panda_0_0", renv, a, b, threads=8)
    judge("pymodule panda validate_motion_batch", p.validate_motion_batch(a, b, env), e_want, e_clr)
    v = (b - a).astype(np.float32)
    d = np.linalg.norm(v, axis=1).astype(np.float32)
    v_want, v_clr = ref.motion_clearance("panda_0_0", renv, a, v, d, threads=8)
    judge("pymodule panda validate_vector_batch", p.validate_vector_batch(a, v, d, env), v_want, v_clr)
    fk = p.fk_batch(q[:64])
    assert np.abs(fk - ref.sphere_fk("panda_0_0", q[:64])).max() <= 1e-5
    one = p.fk(q[0])
    assert len(one) == 59 and abs(one[58].x - fk[0, 58, 0]) < 1e-7 and one[3].r == fk[0, 3, 3]
    # the fork's default `panda` is PandaBase<200, 200, 0> (robots/panda_grid.hh:39-41)
    assert np.abs(m.panda.fk_batch(q[:8]) - ref.sphere_fk("panda_2_2", q[:8])).max() <= 1e-5


@pytest.mark.parametrize("name", ["panda_0_0", "ur5"])
def test_sphere_validity_names_match_the_reference(ref, name):
    m = _module()
    rng = np.random.default_rng(8)
    # many overlapping spheres around the arm so that most robot spheres report something
    spec = [("sphere", *rng.uniform(-0.6, 0.6, 2), rng.uniform(0.0, 1.0), rng.uniform(0.08, 0.25)) for _ in range(40)]
    env = _env(m, spec)
    renv = apply_ref(spec, ref.Env())
    robot = getattr(m, name)
    lo, hi = np.asarray(robot.lower_bounds(), np.float32), np.asarray(robot.upper_bounds(), np.float32)
    reported = 0
    for k in range(25):
        q = (lo + rng.random(robot.dimension(), dtype=np.float32) * (hi - lo)).astype(np.float32)
        got = [sorted(int(s) for s in names) for names in robot.sphere_validity(q, env)]
        want = ref.sphere_validity(name, renv, q, cap=64)
        if got != want:
            # only a contact within the band may differ: compare against the spheres' clearances
            s = ref.sphere_fk(name, q[None, :])[0]
            for i, (g, w) in enumerate(zip(got, want)):
                for o in set(g) ^ set(w):
                    c = np.asarray(spec[o][1:4], np.float32)
                    gap = abs(float(np.linalg.norm(s[i, :3] - c)) - (s[i, 3] + np.float32(spec[o][4])))
                    assert gap <= 1e-5, f"{name}: sphere {i} vs obstacle {o}: {gap}"
        reported += sum(len(w) for w in want)
    assert reported > 50


def test_module_has_no_cpu_path():
    m = _module()
    assert "sm_100a" in m.version()
    with pytest.raises((ValueError, RuntimeError)):
        m.panda.validate(np.zeros(6, np.float32))  # wrong dimension is an error, not a guess


def test_problem_dict_to_vamp_in_cpp_reproduces_the_mbm_verdicts(ref):
    """problem_dict_to_vamp (src/vamp/__init__.py:135-181) implemented in the compiled module: MotionBenchMaker problem
    dictionaries -> Environment; start / goal verdicts against the reference over every Panda problem (the reference's own
    count: 699 of 700 valid, resources/README.md:146)."""
    import os

    m = _module()
    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, "mbm_panda.json.gz"))
    valid = total = 0
    for name, problems in mbm["problems"].items():
        for p in problems:
            d = dict(p)
            d.setdefault("problem", name)
            for key in ("sphere", "cylinder", "box"):
                d[key] = [dict(o, name=o.get("name", f"{key}{i}")) for i, o in enumerate(p.get(key, []))]
            env = m.problem_dict_to_vamp(d)
            q = np.asarray([p["start"]] + p["goals"], np.float32)
            got = m.panda_0_0.validate_batch(q, env)
            renv = apply_ref(scenes.problem_to_spec(p), ref.Env())
            want, clr = ref.fkcc_clearance("panda_0_0", renv, q)
            assert not ((got != want) & (clr > 1e-5)).any(), (name, p["index"])
            valid += bool(got.all())
            total += 1
    assert total == 700 and valid == 699
    # ignore_names drops obstacles
    p = mbm["problems"]["table_pick"][0]
    d = dict(p, problem="table_pick")
    for key in ("sphere", "cylinder", "box"):
        d[key] = [dict(o, name=f"{key}{i}") for i, o in enumerate(p.get(key, []))]
    q = np.asarray(_cfgs_for(m.panda_0_0, 5, 4000))
    full = m.panda_0_0.fkcc_batch(q, m.problem_dict_to_vamp(d))
    fewer = m.panda_0_0.fkcc_batch(q, m.problem_dict_to_vamp(d, [f"box{i}" for i in range(len(d["box"]))]))
    assert fewer.sum() > full.sum()


def _cfgs_for(robot, seed, n):
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(robot.lower_bounds(), np.float32), np.asarray(robot.upper_bounds(), np.float32)
    return (lo + rng.random((n, robot.dimension()), dtype=np.float32) * (hi - lo)).astype(np.float32)


def test_multi_robot_layer_over_the_pair_kernels(vb, ref):
    """RobotFactory / MultiRobotEnvironment (robots/robot_factory.hh:23-189, MR_DEV_PLAN.md:176-198) as the API home of
    the robot-robot check: composite validity equals vb_validate_pairs, the inter-robot term equals a restatement from the
    reference's sphere_fk (PARITY UNPINNED: the reference has no implementation)."""
    m = _module()
    assert len(m.RobotFactory.get_available_variants()) == 9
    r0, r1 = m.RobotFactory.create_panda_grid("panda_0_0"), m.RobotFactory.create_panda(1.0, 0.0, 0.0)
    assert r1.get_name() == "panda_1_0" and list(r1.get_base_position()) == [1.0, 0.0, 0.0]
    with pytest.raises(RuntimeError):
        m.RobotFactory.create_panda(0.5, 0.0, 0.0)
    with pytest.raises(RuntimeError):
        m.RobotFactory.create_panda_grid("panda_3_0")
    menv = m.MultiRobotEnvironment()
    menv.add_robot(r0)
    menv.add_robot(r1)
    spec = scenes.cage(radius=0.12)[::2]
    for i, it in enumerate(spec):
        s = m.Sphere([float(v) for v in it[1:4]], float(it[4]))
        menv.add_obstacle(s)
    n = 20000
    qa, qb = _cfgs_for(m.panda_0_0, 81, n), _cfgs_for(m.panda_0_0, 82, n)
    comp = np.stack([qa, qb], axis=1)
    got = menv.validate_batch(comp)
    ra, rb = vb.robot("panda_0_0"), vb.robot("panda_1_0")
    env = scenes.apply(spec, vb.Environment())
    assert np.array_equal(got, vb.validate_pairs_batch(ra, qa, rb, qb, env))
    clear = menv.inter_robot_clear_batch(comp)
    sa, sb = ref.sphere_fk("panda_0_0", qa), ref.sphere_fk("panda_1_0", qb)
    d = sa[:, :, None, :3] - sb[:, None, :, :3]
    rs = sa[:, :, None, 3] + sb[:, None, :, 3]
    v = ((d * d).sum(axis=-1, dtype=np.float32) - rs * rs).reshape(n, -1).min(axis=1)
    assert not ((clear != (v >= 0)) & (np.abs(v) > 2e-6)).any()
    assert 0 < (~clear).sum() < n
    i = int(np.flatnonzero(~clear)[0])
    assert menv.check_inter_robot_collisions([qa[i].tolist(), qb[i].tolist()])
    j = int(np.flatnonzero(clear)[0])
    assert not menv.check_inter_robot_collisions([qa[j].tolist(), qb[j].tolist()])
    # three robots: every pair is checked
    menv.add_robot(m.RobotFactory.create_panda_grid("panda_0_1"))
    qc = _cfgs_for(m.panda_0_0, 83, 2000)
    c3 = np.stack([qa[:2000], qb[:2000], qc], axis=1)
    three = menv.inter_robot_clear_batch(c3)
    assert (three <= clear[:2000]).all() and three.sum() < clear[:2000].sum()
    menv.remove_robot(2)
    assert len(menv) == 2
