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


def _band_edges(robot, env, w):
    """Candidate edges whose verdict is decided within the 1e-5 m band (either answer is acceptable there)."""
    n = len(w)
    ia, ib = np.triu_indices(n, k=1)
    _, mg = robot.validate_motion_batch(w[ia], w[ib], env, margin=True)
    return {(int(a), int(b)) for a, b, m in zip(ia, ib, mg) if m <= BAND}


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
            assert _band_edges(robot, env, w), f"{name}/{scene}: shortcut_path differs from the reference outside the band"
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
            assert _band_edges(robot, env, w), f"{name}: Path::validate differs from the reference outside the band"
            continue
        seen_invalid += want >= 0
        assert robot.validate_path(w, env) == (want < 0)
    assert seen_invalid >= 1
