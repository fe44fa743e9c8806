"""numpy interpreter of the robot IR (trace_parser): evaluates every SSA value of a traced function
for a batch of configurations in float32, with the reference's polynomial sin/cos
(vector/interface.hh:438-469).  Used at code-generation time to establish static geometric facts
about a robot (which fine spheres lie inside their group's bounding sphere) that the kernels use
as conservative filters; it is not part of any query path."""
from __future__ import annotations

from typing import Any

import numpy as np

F = np.float32


def vamp_sin(v):
    v = v.astype(F)
    p = F(-0.478637850138) * (v * np.abs(v)) + F(1.503684069359) * v
    ap = np.abs(p)
    p2 = p * ap
    p3 = p2 * ap
    return (F(0.011596870476) * p3 + F(0.140024078368) * p2) + F(0.665200679751) * p


def vamp_cos(v):
    pi = F(3.14159265359)
    w = v.astype(F) + F(float(pi) / 2.0)
    w = np.where(w >= pi, w - F(2) * pi, w)
    return vamp_sin(w)


def evaluate(fn: dict[str, Any], q: np.ndarray) -> list[np.ndarray]:
    """q: [n, dim] float32 -> list of value arrays (one [n] array per SSA value)."""
    vals: list[np.ndarray] = []

    def opnd(o):
        return F(o["c"]) if isinstance(o, dict) else vals[o]

    for op, a, b in fn["values"]:
        if op == "in":
            vals.append(q[:, a].astype(F))
        elif op == "mul":
            vals.append((opnd(a) * opnd(b)).astype(F) + np.zeros(q.shape[0], F))
        elif op == "add":
            vals.append((opnd(a) + opnd(b)).astype(F) + np.zeros(q.shape[0], F))
        elif op == "sub":
            vals.append((opnd(a) - opnd(b)).astype(F) + np.zeros(q.shape[0], F))
        elif op == "neg":
            vals.append(-opnd(a))
        elif op == "sin":
            vals.append(vamp_sin(opnd(a)))
        elif op == "cos":
            vals.append(vamp_cos(opnd(a)))
        else:
            raise ValueError(op)
    return vals


def centre(vals: list[np.ndarray], sphere, n: int) -> np.ndarray:
    """[n, 3] local-frame centre of a sphere argument list [x, y, z, r]."""
    cols = []
    for a in sphere[:3]:
        cols.append(vals[a["v"]] if a["v"] is not None else np.full(n, a["c"], F))
    return np.stack(cols, axis=1).astype(np.float64)


def inside(vals, fine, guard, n: int, tol: float = 1e-6) -> bool:
    """True iff sphere `fine` lies within sphere `guard` (same frame) for every sampled configuration."""
    if fine[3].get("sym") or guard[3].get("sym"):
        return False
    if [a.get("base") for a in fine[:3]] != [a.get("base") for a in guard[:3]]:
        return False
    d = np.linalg.norm(centre(vals, fine, n) - centre(vals, guard, n), axis=1)
    return bool(np.all(d + fine[3]["c"] <= guard[3]["c"] + tol))
