"""GPU parity of the batch edge callers: Path::validate and planning::shortcut_path.

The expected result is the reference's own loop (planning/plan.hh:155-168, planning/simplify.hh:115-140)
replayed in Python with one call of the compiled reference's validate_motion (oracle/_ref) per candidate,
exactly as the reference makes them.
"""
import numpy as np
import pytest

from mr_vamp_b200 import scenes
from conftest import apply_ref

pytestmark = pytest.mark.gpu

BAND = 1e-5
REF_NAME = {"panda": "panda_0_0"}


def _paths(robot, seed, n_paths, n_way):
    """Random-walk paths: short steps (mostly valid edges) from uniformly drawn starts."""
    rng = np.random.default_rng(seed)
    lo, hi = robot.lower_bounds, robot.lower_bounds + robot.ranges
    out = []
    for _ in range(n_paths):
        w = [(lo + rng.random(robot.dimension, dtype=np.float32) * robot.ranges).astype(np.float32)]
        for _ in range(n_way - 1):
            step = rng.normal(size=robot.dimension).astype(np.float32)
            step *= np.float32(rng.uniform(0.05, 0.6) / np.linalg.norm(step))
            w.append(np.clip(w[-1] + step, lo, hi).astype(np.float32))
        out.append(np.stack(w))
    return out


def _ref_edge(ref, name, renv, a, b):
    return bool(ref.validate_motion(REF_NAME.get(name, name), renv, a[None, :], b[None, :], threads=1)[0])


def _ref_shortcut(ref, name, renv, w):
    """planning/simplify.hh:115-140, one reference validate_motion call per candidate."""
    path = list(range(len(w)))
    if len(path) < 3:
        return path
    i = 0
    while i + 2 < len(path):
        j = len(path) - 1
        while j > i + 1:
            if _ref_edge(ref, name, renv, w[path[i]], w[path[j]]):
                del path[i + 1:j]
                break
            j -= 1
        i += 1
    return path


def _band_edges(ref, name, renv, w):
    """Candidate edges whose verdict is decided within the 1e-5 m band (either answer is acceptable there), as judged
    by the oracle's clearance (oracle/clear)."""
    n = len(w)
    ia, ib = np.triu_indices(n, k=1)
    _, clr = ref.motion_clearance(REF_NAME.get(name, name), renv, w[ia], w[ib], threads=8)
    return {(int(a), int(b)) for a, b, m in zip(ia, ib, clr) if m <= BAND}


@pytest.mark.parametrize("name", ["panda", "ur5"])
@pytest.mark.parametrize("scene", ["empty", "cage", "clutter"])
def test_shortcut_path_matches_reference(vb, ref, name, scene):
    robot = vb.Robot(name)
    spec = {"empty": [], "cage": scenes.cage(), "clutter": scenes.random_scene(11)}[scene]
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    checked = 0
    for w in _paths(robot, 21, 6, 14) + _paths(robot, 22, 2, 3) + _paths(robot, 23, 1, 2):
        got = robot.shortcut_path(w, env)
        want = w[_ref_shortcut(ref, name, renv, w)]
        if got.shape != want.shape or not np.array_equal(got, want):
            assert _band_edges(ref, name, renv, w), f"{name}/{scene}: shortcut_path differs from the reference outside the band"
            continue
        checked += 1
    assert checked >= 7
    print(f"{name}/{scene}: {checked} paths identical to the reference's shortcut_path")


@pytest.mark.parametrize("name", ["panda", "fetch"])
def test_first_invalid_edge_matches_reference(vb, ref, name):
    robot = vb.Robot(name)
    spec = scenes.random_scene(5)
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    seen_invalid = 0
    for w in _paths(robot, 31, 12, 20) + _paths(robot, 32, 1, 1):
        got = robot.first_invalid_edge(w, env)
        want = -1
        for e in range(len(w) - 1):
            if not _ref_edge(ref, name, renv, w[e], w[e + 1]):
                want = e
                break
        if got != want:
            assert _band_edges(ref, name, renv, w), f"{name}: Path::validate differs from the reference outside the band"
            continue
        seen_invalid += want >= 0
        assert robot.validate_path(w, env) == (want < 0)
    assert seen_invalid >= 1


def _ref_connect(ref, rn, renv, a, b, dist, rng_range):
    """planning/rrtc.hh:172-194 replayed with one reference validate_vector call per increment (float32 arithmetic
    as the reference's vector ops: increment = vector * (1 / n_ext), prior += increment)."""
    n_ext = int(np.ceil(np.float32(dist) / np.float32(rng_range)))
    inc_len = np.float32(dist) / np.float32(n_ext)
    inc = ((b - a).astype(np.float32) * (np.float32(1.0) / np.float32(n_ext))).astype(np.float32)
    prior = a.astype(np.float32).copy()
    i = 0
    while i < n_ext and ref.validate_vector(rn, renv, prior[None, :], inc[None, :], np.array([inc_len], np.float32))[0]:
        prior = (prior + inc).astype(np.float32)
        i += 1
    return i, n_ext


@pytest.mark.parametrize("name,rng_range", [("panda", 1.0), ("ur5", 1.5)])
def test_connect_increments_match_the_rrtc_loop(vb, ref, name, rng_range):
    """vb_connect_increments: all n_extensions increments of many connect attempts in one batch; per attempt the number
    of increments the reference's loop (planning/rrtc.hh:172-194) validates before it stops."""
    robot = vb.Robot(name)
    rn = REF_NAME.get(name, name)
    spec = scenes.random_scene(11)
    env = scenes.apply(spec, vb.Environment())
    renv = apply_ref(spec, ref.Env())
    rng = np.random.default_rng(77)
    n = 300
    lo = robot.lower_bounds
    a = (lo + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)
    b = (lo + rng.random((n, robot.dimension), dtype=np.float32) * robot.ranges).astype(np.float32)
    b[: n // 2] = np.clip(a[: n // 2] + rng.normal(0, 0.25, (n // 2, robot.dimension)), lo, lo + robot.ranges).astype(np.float32)
    dist = np.linalg.norm((b - a).astype(np.float32), axis=1).astype(np.float32)  # what the NN query returns
    n_valid, n_ext = robot.connect_increments(a, b, dist, rng_range, env)
    want = [_ref_connect(ref, rn, renv, a[i], b[i], dist[i], rng_range) for i in range(n)]
    assert np.array_equal(n_ext, [w[1] for w in want])
    differ = [i for i in range(n) if n_valid[i] != want[i][0]]
    for i in differ:
        # either answer is acceptable only when an increment's verdict is decided inside the band
        ne = int(n_ext[i])
        inc = ((b[i] - a[i]) * (np.float32(1.0) / np.float32(ne))).astype(np.float32)
        priors = np.stack([a[i]] * ne).astype(np.float32)
        for k in range(1, ne):
            priors[k] = (priors[k - 1] + inc).astype(np.float32)
        _, clr = ref.motion_clearance(rn, renv, priors, np.stack([inc] * ne), np.full(ne, dist[i] / np.float32(ne), np.float32))
        assert (clr <= BAND).any(), f"{name}: attempt {i}: {n_valid[i]} increments against the reference's {want[i][0]}"
    connected = int((n_valid == n_ext).sum())
    print(f"{name}: {n} connect attempts, {connected} connect, {len(differ)} differ inside the band")
    assert 0 < connected < n
