"""CAPT construction on the device (vb_env_add_pointcloud / mr_vamp_b200/csrc/vb_capt_build.cu) against
the reference's constructor (collision/capt.hh:137-398, through the oracle): same split values, same leaf boxes, the
same SET of affordance points in every leaf -- and therefore the same answers to every query."""
import numpy as np
import pytest

from conftest import judge

pytestmark = pytest.mark.gpu


def _cloud(seed, n):
    """Points on a table top, a wall and a post around the robot, plus some scattered points."""
    rng = np.random.default_rng(seed)
    k = n // 4
    table = np.stack([rng.uniform(0.25, 0.9, k), rng.uniform(-0.5, 0.5, k), np.full(k, 0.3) + rng.normal(0, 0.002, k)], axis=1)
    wall = np.stack([np.full(k, -0.45) + rng.normal(0, 0.002, k), rng.uniform(-0.6, 0.6, k), rng.uniform(0.0, 1.0, k)], axis=1)
    post = np.stack([0.1 * np.cos(rng.uniform(0, 6.28, k)) + 0.3, 0.1 * np.sin(rng.uniform(0, 6.28, k)) - 0.55, rng.uniform(0.0, 0.9, k)], axis=1)
    loose = rng.uniform(-0.8, 0.8, (n - 3 * k, 3)) + [0, 0, 0.8]
    return np.concatenate([table, wall, post, loose]).astype(np.float32)


def _leaf_sets(c):
    """per leaf: sorted tuple of its affordance points (infinite padding dropped)"""
    x, y, z, s = c["aff_x"].reshape(-1, 8), c["aff_y"].reshape(-1, 8), c["aff_z"].reshape(-1, 8), c["aff_starts"]
    out = []
    for leaf in range(len(s) - 1):
        px, py, pz = x[s[leaf]:s[leaf + 1]].ravel(), y[s[leaf]:s[leaf + 1]].ravel(), z[s[leaf]:s[leaf + 1]].ravel()
        keep = np.isfinite(px)
        out.append(sorted(zip(px[keep].tolist(), py[keep].tolist(), pz[keep].tolist())))
    return out


@pytest.mark.parametrize("n_points", [1, 7, 300, 5000, 40000])
def test_device_built_capt_equals_the_reference(vb, ref, n_points):
    pts = _cloud(n_points, max(n_points, 4))[:n_points]
    renv = ref.Env()
    renv.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    want = renv.capt(0)
    env = vb.Environment()
    ns = env.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    got = env.capt(0)
    assert got["nlog2"] == want["nlog2"]
    assert np.array_equal(got["tests"], want["tests"], equal_nan=True), "split values differ"
    assert np.array_equal(got["aabb_top"], want["aabb_top"])
    assert np.array_equal(got["aff_starts"], want["aff_starts"]), "vector counts per leaf differ"
    assert np.array_equal(got["aabbs"], want["aabbs"]), "leaf boxes differ"
    a, b = _leaf_sets(got), _leaf_sets(want)
    bad = [i for i in range(len(a)) if a[i] != b[i]]
    assert not bad, f"{len(bad)} of {len(a)} leaves hold different affordance sets (first: leaf {bad[0]})"
    # the cell's own point comes first in its list (capt.hh:173-177)
    s = got["aff_starts"]
    first = np.flatnonzero(s[1:] > s[:-1])
    assert np.array_equal(got["aff_x"].reshape(-1, 8)[s[first], 0], want["aff_x"].reshape(-1, 8)[s[first], 0])
    print(f"{n_points} points: {len(a)} leaves, {int(s[-1])} affordance vectors, device build {ns / 1e6:.2f} ms")


def test_queries_against_a_device_built_capt_match_the_reference(vb, ref):
    """End to end without the oracle on the product side: raw points in, verdicts out."""
    pts = _cloud(3, 40000)[:30000]  # table, wall and post only (the scattered points would leave nothing valid)
    renv = ref.Env()
    renv.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    env = vb.Environment()
    env.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    robot = vb.Robot("panda")
    rng = np.random.default_rng(5)
    q = (robot.lower_bounds + rng.random((20000, 7), dtype=np.float32) * robot.ranges).astype(np.float32)
    want, clr = ref.fkcc_clearance("panda_0_0", renv, q, threads=8)
    judge("device-built CAPT fkcc", robot.fkcc_batch(q, env), want, clr)
    a, b = q[:3000], np.clip(q[:3000] + np.float32(0.15), robot.lower_bounds, robot.lower_bounds + robot.ranges)
    e_want, e_clr = ref.motion_clearance("panda_0_0", renv, a, b, threads=8)
    judge("device-built CAPT edges", robot.validate_motion_batch(a, b, env), e_want, e_clr)
    assert 0.05 < want.mean() < 0.95


def test_device_build_of_the_100k_point_cloud_is_timed(vb, ref):
    """The SURVEY M3 cloud (100 000 points on the faces of table_pick scene 1).  Thousands of its points share a
    coordinate exactly (a face of an axis-aligned box), and which of several equal coordinates lands left of a median is
    up to the sort (the reference uses pdqsort, which is not stable; capt.hh:137-153) -- so the two trees are different
    but equally valid CAPTs of the same cloud: the buffers are compared in what does not depend on ties, and the queries
    by their verdicts."""
    from test_configs_gpu import _table_pick_cloud, _record

    pts = _table_pick_cloud(vb)
    env = vb.Environment()
    ns = env.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    got = env.capt(0)
    renv = ref.Env()
    ref_ns = renv.add_pointcloud(pts, 0.012, 0.06, 0.0025)
    want = renv.capt(0)
    assert got["nlog2"] == want["nlog2"] and np.array_equal(got["aabb_top"], want["aabb_top"])
    assert got["tests"][0] == want["tests"][0]  # the root split: the median of all x coordinates
    # every point is the first entry of exactly one leaf, in both trees
    def reps(c):
        s = c["aff_starts"]
        first = np.flatnonzero(s[1:] > s[:-1])
        x, y, z = (c[k].reshape(-1, 8)[s[first], 0] for k in ("aff_x", "aff_y", "aff_z"))
        return np.unique(np.stack([x, y, z], axis=1), axis=0)
    cloud = np.unique(pts, axis=0)
    assert np.array_equal(reps(got), cloud) and np.array_equal(reps(want), cloud)
    n_got, n_want = int(got["aff_starts"][-1]), int(want["aff_starts"][-1])
    assert abs(n_got - n_want) < 0.02 * n_want, (n_got, n_want)
    robot = vb.Robot("panda")
    a, b = robot.random_edges(7, 0, 4000, 0.05, 1.0)
    e_want, e_clr = ref.motion_clearance("panda_0_0", renv, a, b, threads=8)
    rec = judge("device-built CAPT of the 100k-point cloud, edges", robot.validate_motion_batch(a, b, env), e_want, e_clr)
    _record("n3_capt_build_100k", points=int(pts.shape[0]), affordance_vectors=n_got, reference_affordance_vectors=n_want,
            device_build_ms=ns / 1e6, reference_cpu_build_ms=ref_ns / 1e6, edge_verdicts_differ=rec["differ"],
            includes="upload of the points, the build, download of the buffers")
