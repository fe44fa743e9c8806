"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

numpy restatement of the fkcc path of mr-vamp, independent of both the CUDA library and the
compiled reference (oracle/_ref).  It exists so that the CPU test-suite can check the *algorithm*
three ways (reference == restatement == CUDA) and so that a box without the reference checkout still
has a checker.  Pinned by tests/test_oracle_cpu.py against oracle/_ref (the reference's own code
compiled here) on sphere centres, primitive values, verdicts and edge verdicts.

What is restated, with the reference lines it follows (paths under src/impl/vamp/):
  * polynomial sin/cos                     vector/interface.hh:438-469
  * sphere_fk / interleaved_sphere_fk      robots/<robot>/fk.hh, interpreted from the robot IR that
                                           mr_vamp_b200.codegen.trace_parser reads from the checkout
                                           (cached as oracle/_ref/ir_<robot>.json so it travels)
  * sphere-sphere / capsule / cuboid       collision/sphere_sphere.hh:9-23, sphere_capsule.hh:8-44,
                                           sphere_cuboid.hh:8-52 (general forms; the z-aligned variants
                                           are the same functions with zero axis components)
  * sphere_environment_in_collision        collision/validity.hh:46-150 WITHOUT the min_distance pruning
                                           (pruning only skips obstacles that cannot collide, up to the
                                           approximate-sqrt band documented in DESIGN.md)
  * validate_motion                        planning/validate.hh:23-75 (same interpolants: a + v*(k+1)/8
                                           - i*v/(8n), n = max(ceil(|v|/8*res), 1)); order-free AND
All arithmetic is float32; FMA contraction of the compiled code is not modelled (ulp-level).
"""
from __future__ import annotations

import json
import os
from typing import Any, Sequence

import numpy as np

F = np.float32
_HERE = os.path.dirname(os.path.abspath(__file__))
_IR_DIR = os.path.join(_HERE, "_ref")


def vamp_sin(v):
    v = np.asarray(v, F)
    p = F(-0.478637850138) * (v * np.abs(v)) + F(1.503684069359) * v
    ap = np.abs(p)
    p2 = p * ap
    p3 = p2 * ap
    return ((F(0.011596870476) * p3 + F(0.140024078368) * p2) + F(0.665200679751) * p).astype(F)


def vamp_cos(v):
    pi = F(3.14159265359)
    w = np.asarray(v, F) + F(float(pi) / 2.0)
    w = np.where(w >= pi, w - F(2) * pi, w).astype(F)
    return vamp_sin(w)


def load_ir(robot: str) -> dict[str, Any]:
    """IR of robots/<robot>/fk.hh: from the cache written by save_ir(), else parsed from $VAMP_SRC."""
    path = os.path.join(_IR_DIR, f"ir_{robot}.json")
    if os.path.exists(path):
        with open(path) as fh:
            return json.load(fh)
    from mr_vamp_b200.codegen import trace_parser  # the parser is a build tool; it computes nothing

    return trace_parser.parse_robot(robot)


def save_ir(robots: Sequence[str] = ("sphere", "ur5", "panda", "fetch", "baxter")) -> None:
    from mr_vamp_b200.codegen import trace_parser

    os.makedirs(_IR_DIR, exist_ok=True)
    for r in robots:
        trace_parser.save_ir(trace_parser.parse_robot(r), os.path.join(_IR_DIR, f"ir_{r}.json"))


def _evaluate(fn, q):
    vals = []

    def opnd(o):
        return F(o["c"]) if isinstance(o, dict) else vals[o]

    zero = np.zeros(q.shape[0], F)
    for op, a, b in fn["values"]:
        if op == "in":
            vals.append(q[:, a].astype(F))
        elif op == "mul":
            vals.append((opnd(a) * opnd(b)).astype(F) + zero)
        elif op == "add":
            vals.append((opnd(a) + opnd(b)).astype(F) + zero)
        elif op == "sub":
            vals.append((opnd(a) - opnd(b)).astype(F) + zero)
        elif op == "neg":
            vals.append(-opnd(a))
        elif op == "sin":
            vals.append(vamp_sin(opnd(a)))
        elif op == "cos":
            vals.append(vamp_cos(opnd(a)))
        else:
            raise ValueError(op)
    return vals


def _coord(vals, a, n, base, radius):
    if a.get("sym") == "radius":
        return np.full(n, radius, F)
    if a["v"] is not None:
        v = vals[a["v"]]
        return (v + F(base["xyz".index(a["base"])])).astype(F) if a.get("base") else v
    if a.get("base"):  # literal + base: evaluated in double, narrowed (C++ semantics)
        return np.full(n, F(float(a["c"]) + float(F(base["xyz".index(a["base"])]))), F)
    return np.full(n, F(a["c"]), F)


def _sphere(vals, s, n, base, radius):
    return [_coord(vals, a, n, base, radius) for a in s]


def sphere_fk(robot: str, q, base=(0.0, 0.0, 0.0), radius: float = 0.2) -> np.ndarray:
    """[n, dim] -> [n, n_spheres, 4] (Robot::sphere_fk)."""
    ir = load_ir(robot)
    q = np.ascontiguousarray(q, F).reshape(-1, ir["dimension"])
    n = q.shape[0]
    fn = ir["functions"]["sphere_fk"]
    vals = _evaluate(fn, q)
    out = np.zeros((n, ir["n_spheres"], 4), F)
    for it in fn["program"]:
        if it["k"] == "out":
            out[:, it["idx"], "xyzr".index(it["field"])] = _coord(vals, it["arg"], n, base, radius)
    return out


# ---- collision primitives: value < 0 <=> collision ------------------------------------------------
def sphere_sphere(ax, ay, az, ar, bx, by, bz, br):
    dx, dy, dz, rs = ax - bx, ay - by, az - bz, ar + br
    return (dx * dx + dy * dy + dz * dz - rs * rs).astype(F)


def sphere_capsule(c8, x, y, z, r):
    x1, y1, z1, xv, yv, zv, cr, rdv = (F(v) for v in c8)
    dot = (x - x1) * xv + (y - y1) * yv + (z - z1) * zv
    t = np.clip(dot * rdv, F(0), F(1)).astype(F)
    dx, dy, dz, rs = x - (x1 + xv * t), y - (y1 + yv * t), z - (z1 + zv * t), r + cr
    return (dx * dx + dy * dy + dz * dz - rs * rs).astype(F)


def sphere_cuboid(c15, x, y, z, r):
    c = [F(v) for v in c15]
    xs, ys, zs = x - c[0], y - c[1], z - c[2]
    acc = np.zeros_like(x, dtype=F)
    for k in range(3):
        a = c[3 + 3 * k : 6 + 3 * k]
        d = np.maximum(np.abs(a[0] * xs + a[1] * ys + a[2] * zs) - c[12 + k], F(0)).astype(F)
        acc = acc + d * d
    return (acc - r * r).astype(F)


class Env:
    """Scene spec (mr_vamp_b200.scenes format) evaluated with the restated primitives."""

    def __init__(self, spec=()):
        self.spec = list(spec)
        for it in self.spec:
            if it[0] not in ("sphere", "cuboid", "capsule"):
                raise ValueError("oracle.port.Env supports spheres, cuboids and capsules")

    def hit(self, x, y, z, r):
        out = np.zeros(np.shape(x), bool)
        for it in self.spec:
            if it[0] == "sphere":
                v = sphere_sphere(x, y, z, r, *(F(t) for t in it[1:5]))
            elif it[0] == "cuboid":
                v = sphere_cuboid(it[1], x, y, z, r)
            else:
                v = sphere_capsule(it[1], x, y, z, r)
            out |= v < 0
        return out


def fkcc(robot: str, env: Env, q, base=(0.0, 0.0, 0.0), radius: float = 0.2) -> np.ndarray:
    """Per-configuration verdict of interleaved_sphere_fk (True = collision free): a configuration
    collides iff some group has a colliding guard AND a colliding fine test."""
    ir = load_ir(robot)
    q = np.ascontiguousarray(q, F).reshape(-1, ir["dimension"])
    n = q.shape[0]
    fn = ir["functions"]["interleaved_sphere_fk"]
    vals = _evaluate(fn, q)
    bad = np.zeros(n, bool)
    for it in fn["program"]:
        if it["k"] == "env":
            g = env.hit(*_sphere(vals, it["guard"], n, base, radius))
            fine = it["fine"]
            if not fine or it.get("leaf"):
                bad |= g
                continue
            f = np.zeros(n, bool)
            for s in fine:
                f |= env.hit(*_sphere(vals, s, n, base, radius))
            bad |= g & f
        elif it["k"] == "self":
            zero = (0.0, 0.0, 0.0)
            a, b = it["guard"]
            g = sphere_sphere(*_sphere(vals, a, n, zero, radius), *_sphere(vals, b, n, zero, radius)) < 0
            if not it["fine"]:
                bad |= g
                continue
            f = np.zeros(n, bool)
            for a, b in it["fine"]:
                f |= sphere_sphere(*_sphere(vals, a, n, zero, radius), *_sphere(vals, b, n, zero, radius)) < 0
            bad |= g & f
    return ~bad


def motion_points(a, b, resolution: int) -> np.ndarray:
    """Interpolants of validate_vector for ONE edge: [8n, dim] (planning/validate.hh:23-65)."""
    a, b = np.asarray(a, F), np.asarray(b, F)
    v = (b - a).astype(F)
    dist = F(np.sqrt(F(np.sum((v * v).astype(F), dtype=F))))
    n = int(max(np.ceil(dist / F(8) * F(resolution)), 1))
    pts = []
    back = (v / F(8 * n)).astype(F)
    block = [(a + v * (F(k + 1) / F(8))).astype(F) for k in range(8)]
    for _ in range(n):
        pts.extend(block)
        block = [(p - back).astype(F) for p in block]
    return np.stack(pts)


def validate_motion(robot: str, env: Env, a, b, resolution: int | None = None, base=(0.0, 0.0, 0.0)) -> np.ndarray:
    ir = load_ir(robot)
    res = resolution or ir["resolution"]
    a = np.ascontiguousarray(a, F).reshape(-1, ir["dimension"])
    b = np.ascontiguousarray(b, F).reshape(-1, ir["dimension"])
    out = np.zeros(a.shape[0], bool)
    for i in range(a.shape[0]):
        out[i] = bool(fkcc(robot, env, motion_points(a[i], b[i], res), base).all())
    return out
