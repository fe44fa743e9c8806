"""PRM / FCIT* batching: the sampling loop of planning/prm.hh:111-150 replayed with the compiled
reference (sample -> fkcc -> k nearest -> validate_motion per neighbour) against ONE vb_sample_valid call for the
vertices and ONE vb_validate_motion_indexed call for all neighbour edges."""
import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref, BAND

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,rn", [("panda", "panda_0_0"), ("ur5", "ur5")])
def test_roadmap_vertices_and_edges_match_the_reference_loop(vb, ref, name, rn):
    robot = vb.Robot(name)
    spec = scenes.random_scene(7)
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    k_vertices, k_nn, seed = 600, 6, 1234
    # ---- vertices: first 600 valid samples of the library stream
    got_q, got_idx, consumed = robot.sample_valid(k_vertices, env, seed=seed, first=0, max_candidates=20000)
    stream = robot.random_configs(seed, 0, 20000)
    ok_ref, clr = ref.fkcc_clearance(rn, renv, stream, threads=8)
    want_idx = np.flatnonzero(ok_ref)[:k_vertices]
    assert got_q.shape[0] == k_vertices
    if not np.array_equal(got_idx, want_idx):
        # a sample decided inside the band may be kept by one side only
        odd = set(got_idx.tolist()) ^ set(want_idx.tolist())
        assert all(clr[i] <= BAND for i in odd if i < max(want_idx[-1], got_idx[-1])), "vertex sets differ outside the band"
    else:
        assert consumed == want_idx[-1] + 1  # where the reference's loop stands after its 600th valid sample
    assert np.array_equal(got_q, stream[got_idx.astype(np.int64)])
    # the same with caller-supplied candidates
    q2, idx2, _ = robot.sample_valid(50, env, candidates=stream[:3000])
    assert np.array_equal(idx2, got_idx[:50]) and np.array_equal(q2, got_q[:50])
    # fewer valid candidates than asked for: everything consumed
    q3, idx3, c3 = robot.sample_valid(10_000, env, candidates=stream[:500])
    assert c3 == 500 and idx3.size == int(robot.fkcc_batch(stream[:500], env).sum())
    # ---- edges: every new vertex against its k nearest earlier vertices (brute force), as prm.hh:134-147 does
    pairs = []
    for i in range(1, k_vertices):
        d = np.linalg.norm(got_q[:i] - got_q[i], axis=1)
        for j in np.argsort(d)[:k_nn]:
            pairs.append((int(j), i))  # validate_motion(neighbor, temp)
    pairs = np.asarray(pairs, np.uint32)
    table = vb.DeviceTable(robot, got_q)
    got_e = table.validate_edges(pairs, env)
    want_e, e_clr = ref.motion_clearance(rn, renv, got_q[pairs[:, 0]], got_q[pairs[:, 1]], threads=8)
    hard = (got_e != want_e) & (e_clr > BAND)
    print(f"{name}: {k_vertices} vertices from {consumed} samples, {len(pairs)} neighbour edges, {int(want_e.sum())} valid, "
          f"differ {int((got_e != want_e).sum())} hard {int(hard.sum())}")
    assert not hard.any()
    assert 0 < want_e.sum() < len(pairs)
    # and the same verdicts as the two-array form
    assert np.array_equal(got_e, robot.validate_motion_batch(got_q[pairs[:, 0]], got_q[pairs[:, 1]], env))
    # an out-of-range row is an invalid edge, not a wild read
    bad = np.asarray([[0, k_vertices], [1, 2]], np.uint32)
    assert not table.validate_edges(bad, env)[0]
    table.close()
