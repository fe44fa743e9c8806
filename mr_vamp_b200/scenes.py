"""Scene descriptions for the fkcc path: neutral "spec" lists that can be applied to an
``api.Environment`` (and, in the tests, to the reference oracle) so both sides receive the SAME raw
shape floats — Euler-angle conversion (Eigen in the reference, collision/factory.hh) never enters
the comparison.

A spec is a list of tuples:
  ("sphere", x, y, z, r) | ("cuboid", raw15) | ("capsule", raw8) | ("heightfield", off_scale6, data2d)
"""
from __future__ import annotations

import gzip
import json
import os
from typing import Any, Sequence

import numpy as np

from . import api

# scripts/sphere_cage_example.py:10-31
CAGE_START = [0.0, -0.785, 0.0, -2.356, 0.0, 1.571, 0.785]
CAGE_GOAL = [2.35, 1.0, 0.0, -0.8, 0.0, 2.5, 0.785]
CAGE_CENTERS = [
    [0.55, 0, 0.25], [0.35, 0.35, 0.25], [0, 0.55, 0.25], [-0.55, 0, 0.25], [-0.35, -0.35, 0.25],
    [0, -0.55, 0.25], [0.35, -0.35, 0.25], [0.35, 0.35, 0.8], [0, 0.55, 0.8], [-0.35, 0.35, 0.8],
    [-0.55, 0, 0.8], [-0.35, -0.35, 0.8], [0, -0.55, 0.8], [0.35, -0.35, 0.8],
]


def cage(radius: float = 0.2, offset: Sequence[float] = (0.0, 0.0, 0.0)) -> list[tuple]:
    return [("sphere", c[0] + offset[0], c[1] + offset[1], c[2] + offset[2], radius) for c in CAGE_CENTERS]


def problem_to_spec(problem: dict[str, Any], offset: Sequence[float] = (0.0, 0.0, 0.0)) -> list[tuple]:
    """MotionBenchMaker problem dict -> spec, as vamp.problem_dict_to_vamp does (src/vamp/__init__.py:135-181):
    spheres, cylinders -> capsules (boxes for the "box" problem), boxes -> cuboids."""
    off = np.asarray(offset, dtype=np.float64)
    spec: list[tuple] = []
    for o in problem.get("sphere", []):
        p = np.asarray(o["position"]) + off
        spec.append(("sphere", p[0], p[1], p[2], o["radius"]))
    for o in problem.get("cylinder", []):
        p = (np.asarray(o["position"]) + off).tolist()
        if problem.get("problem") == "box":
            c = api.Cuboid(p, o["orientation_euler_xyz"], [o["radius"], o["radius"], o["length"] / 2])
            spec.append(("cuboid", c.raw()))
        else:
            c = api.Cylinder(p, o["orientation_euler_xyz"], o["radius"], o["length"])
            spec.append(("capsule", c.raw()))
    for o in problem.get("box", []):
        p = (np.asarray(o["position"]) + off).tolist()
        spec.append(("cuboid", api.Cuboid(p, o["orientation_euler_xyz"], o["half_extents"]).raw()))
    return spec


def random_scene(seed: int, n_spheres=4, n_cuboids=4, n_zcuboids=6, n_capsules=3, n_zcapsules=3, reach=1.0):
    """Seeded clutter around the origin covering all five primitive lists."""
    rng = np.random.default_rng(seed)
    spec: list[tuple] = []

    def pos():
        while True:
            p = rng.uniform(-reach, reach, 3)
            p[2] = abs(p[2])
            if np.linalg.norm(p[:2]) > 0.25:
                return p

    for _ in range(n_spheres):
        p = pos()
        spec.append(("sphere", p[0], p[1], p[2], rng.uniform(0.03, 0.2)))
    for _ in range(n_cuboids):
        spec.append(("cuboid", api.Cuboid(pos(), rng.uniform(-np.pi, np.pi, 3), rng.uniform(0.02, 0.25, 3)).raw()))
    for _ in range(n_zcuboids):
        spec.append(("cuboid", api.Cuboid(pos(), [0, 0, rng.uniform(-np.pi, np.pi)], rng.uniform(0.02, 0.3, 3)).raw()))
    for _ in range(n_capsules):
        spec.append(("capsule", api.Cylinder(pos(), rng.uniform(-np.pi, np.pi, 3), rng.uniform(0.02, 0.1), rng.uniform(0.1, 0.6)).raw()))
    for _ in range(n_zcapsules):
        spec.append(("capsule", api.Cylinder(pos(), [0, 0, rng.uniform(-np.pi, np.pi)], rng.uniform(0.02, 0.1), rng.uniform(0.1, 0.6)).raw()))
    return spec


def apply(spec: list[tuple], env: "api.Environment") -> "api.Environment":
    for item in spec:
        kind = item[0]
        if kind == "sphere":
            env.add_sphere(api.Sphere(item[1:4], item[4]))
        elif kind == "cuboid":
            env.add_cuboid(api.Cuboid.from_raw(item[1]))
        elif kind == "capsule":
            env.add_capsule(api.Cylinder.from_raw(item[1]))
        elif kind == "heightfield":
            d = np.asarray(item[2], dtype=np.float32)
            env.add_heightfield(api.HeightField.from_raw(item[1], d.shape[1], d.shape[0], d))
        else:
            raise ValueError(kind)
    return env


def load_mbm(path: str) -> dict[str, Any]:
    """Problem sets converted from resources/<robot>/problems.tar.bz2 (tests/golden/make_mbm_fixture.py)."""
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rt") as fh:
        return json.load(fh)


GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
