"""Two-robot composite validation (BASELINE.json configs[4]).  The reference has no
implementation of the robot-robot check (only the plan in MR_DEV_PLAN.md:176-198), so PARITY IS UNPINNED:
the checker is a CPU restatement built from the reference's own pieces -- Robot::fkcc per robot and
sphere_sphere_sql2 over the world-frame spheres of Robot::sphere_fk of each base variant."""
import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref

pytestmark = pytest.mark.gpu


def _cfgs(robot, seed, n):
    rng = np.random.default_rng(seed)
    return (robot.lower_bounds + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)


def _pairs_restated(sa, sb):
    """min over sphere pairs of d^2 - (ra + rb)^2 (collision/sphere_sphere.hh:9-23), float32 -> [n]"""
    d = sa[:, :, None, :3] - sb[:, None, :, :3]
    rs = sa[:, :, None, 3] + sb[:, None, :, 3]
    v = (d * d).sum(axis=-1, dtype=np.float32) - rs * rs
    return v.reshape(v.shape[0], -1).min(axis=1)


@pytest.mark.parametrize("names", [("panda_0_0", "panda_1_0"), ("panda_1_0", "panda_0_0"), ("ur5", "panda_1_0")])
def test_two_robot_composite_validity(vb, ref, names):
    ra, rb = vb.robot(names[0]), vb.robot(names[1])
    # the 14-sphere cage duplicated at both bases, thinned so that some composites are valid
    spec = scenes.cage(radius=0.12, offset=ra.base)[::2] + scenes.cage(radius=0.12, offset=rb.base)[1::2]
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    n = 20000
    qa, qb = _cfgs(ra, 51, n), _cfgs(rb, 52, n)
    got = vb.validate_pairs_batch(ra, qa, rb, qb, env)
    (va, ca), (vb_, cb) = ref.fkcc_clearance(names[0], renv, qa, threads=8), ref.fkcc_clearance(names[1], renv, qb, threads=8)
    pair_value = _pairs_restated(ref.sphere_fk(names[0], qa), ref.sphere_fk(names[1], qb))
    want = va & vb_ & ~(pair_value < 0)
    # clearance of the decisive pair in metres: sqrt(v + R^2) - R is within |v| / (2 R) ~ 10 |v| for R ~ 0.05
    near = np.abs(pair_value) < 2e-6
    hard = (got != want) & ~near & (ca > 1e-5) & (cb > 1e-5)  # each robot's own band from the oracle's clearance
    print(f"{names}: valid={want.mean():.3f} each={va.mean():.3f}/{vb_.mean():.3f} pair hits={(pair_value < 0).mean():.3f} "
          f"differ={int((got != want).sum())} hard={int(hard.sum())}")
    assert (pair_value < 0).any() and (want != (va & vb_)).any(), "the robot-robot term must matter"
    assert hard.sum() == 0


@pytest.mark.parametrize("names", [("panda_0_0", "panda_1_0"), ("panda_1_0", "panda_0_0")])
def test_fused_kernel_equals_the_three_launch_path(vb, names, monkeypatch):
    """Two robots of one kind run as one fused kernel (pairs_fused_kernel); the sphere-buffer path (fkcc twice, sphere_fk
    twice, pair kernel) that robots of different kinds use must give the same words."""
    ra, rb = vb.robot(names[0]), vb.robot(names[1])
    spec = scenes.cage(radius=0.12, offset=ra.base)[::2]
    env = scenes.apply(spec, vb.Environment())
    n = 100_003
    qa, qb = _cfgs(ra, 71, n), _cfgs(rb, 72, n)
    fused = vb.validate_pairs_batch(ra, qa, rb, qb, env)
    monkeypatch.setenv("VB_PAIRS_UNFUSED", "1")
    unfused = vb.validate_pairs_batch(ra, qa, rb, qb, env)
    monkeypatch.delenv("VB_PAIRS_UNFUSED")
    each = ra.fkcc_batch(qa, env) & rb.fkcc_batch(qb, env)
    differ = int((fused != unfused).sum())
    print(f"{names}: valid {fused.mean():.4f} (each robot alone {each.mean():.4f}), fused vs unfused differ on {differ}")
    assert (fused != each).any(), "the robot-robot term must matter"
    # the two paths round the sphere centres the same way (same chain); a handful of contacts within 1e-6 m^2 may differ
    assert differ <= 2


def test_far_apart_robots_reduce_to_the_and_of_both(vb, ref):
    ra, rb = vb.robot("panda_0_0"), vb.robot("panda_2_2")
    env = scenes.apply(scenes.cage(), vb.Environment())
    q = _cfgs(ra, 53, 5000)
    got = vb.validate_pairs_batch(ra, q, rb, q[::-1].copy(), env)
    assert np.array_equal(got, ra.fkcc_batch(q, env) & rb.fkcc_batch(q[::-1].copy(), env))
