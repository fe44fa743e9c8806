"""Joint-pair tables for self-collision groups whose geometry depends on two joints only.

A self-collision group of the reference (robots/<robot>/fk.hh: `if (sphere_sphere_self_collision(ga, gb))
{ a_i x b_j ... }`) tests spheres of two links against each other.  When the links are separated by one or
two joints, every distance in the group is a function of those joints alone (e.g. Panda link5 against
link7 / hand / fingers: joints 6 and 7).  For such groups the verdict over the whole (qa, qb) rectangle
of the joint ranges is tabulated at code-generation time:

    free    no pair of the group overlaps anywhere in the cell (the group is skipped),
    hit     the guard pair and one fine pair overlap everywhere in the cell (the configuration collides),
    unsure  anything else: the kernel runs the group's tests exactly as the reference does.

`free` and `hit` are decided from the clearances at the four cell corners with a margin that bounds the
variation inside the cell (a quarter of the corner spread + MARGIN metres, which covers the bilinear
interpolation error of these smooth functions at the grid pitch used and the float32 rounding differences
between this interpreter and the kernels); cells the margin cannot decide stay `unsure`, so the table never
changes a verdict, only who computes it.  tests/test_oracle_cpu.py re-checks random points against the cells.
"""
from __future__ import annotations

from typing import Any

import numpy as np

from . import ir_eval

N_CELLS = 256  # cells per joint
MARGIN = 1.0e-3  # metres
MAX_GROUPS = 16  # per table: cell = free bits [0, 16) | hit bits [16, 32)
F = np.float32


def _self_items(fn):
    return [it for it in fn["program"] if it["k"] == "self"]


def _pairs(it):
    """[(sphere a, sphere b)] of the guard pair and the fine pairs of a self item (numeric radii only)."""
    ga, gb = it["guard"]
    fine = [] if (it.get("leaf") or not it["fine"]) else list(it["fine"])
    return (ga, gb), fine


def _numeric(s) -> bool:
    return not s[3].get("sym")


def _clear(vals, a, b, n):
    d = np.linalg.norm(ir_eval.centre(vals, a, n) - ir_eval.centre(vals, b, n), axis=1)
    return d - (float(F(a[3]["c"])) + float(F(b[3]["c"])))


def dependencies(ir: dict[str, Any], fn, bases: int = 32, variations: int = 8, tol: float = MARGIN / 4) -> list[set[int]]:
    """Per self group: the joints that change any of its distances (numerically, over sampled configurations).

    The reference's polynomial sin/cos (vector/interface.hh:438-469) do not satisfy sin^2 + cos^2 = 1 exactly,
    so the traced link frames are not perfectly rigid: a distance between two links moves by up to ~3e-5 m
    with joints that are not between them.  Joints whose influence stays below `tol` (a quarter of the
    table margin) count as independent; MARGIN absorbs the residue."""
    dim = ir["dimension"]
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    rng = np.random.default_rng(12345)
    base = (lo + rng.random((bases, dim)).astype(F) * rg).astype(F)
    q = np.repeat(base[None, :, None, :], dim, axis=0).repeat(variations, axis=2)  # [joint][base][variation][dim]
    for j in range(dim):
        q[j, :, :, j] = lo[j] + rng.random((bases, variations)).astype(F) * rg[j]
    flat = q.reshape(-1, dim)
    n = flat.shape[0]
    vals = ir_eval.evaluate(fn, flat)
    out = []
    for it in _self_items(fn):
        guard, fine = _pairs(it)
        deps: set[int] = set()
        pairs = [guard] + fine
        if not all(_numeric(a) and _numeric(b) for a, b in pairs):
            out.append(set(range(dim)))
            continue
        for a, b in pairs:
            c = _clear(vals, a, b, n).reshape(dim, bases, variations)
            spread = (c.max(axis=2) - c.min(axis=2)).max(axis=1)
            deps |= {j for j in range(dim) if spread[j] > tol}
        out.append(deps)
    return out


def _cells(c: np.ndarray):
    """corner values [N+1, N+1] -> (min, max) over the four corners of every cell, [N, N] each."""
    c4 = np.stack([c[:-1, :-1], c[1:, :-1], c[:-1, 1:], c[1:, 1:]])
    return c4.min(axis=0), c4.max(axis=0)


def build(ir: dict[str, Any], fn, n_cells: int = N_CELLS) -> list[dict[str, Any]]:
    """[{joints: (ja, jb), groups: [self group index per table bit], cells: uint32 [n_cells * n_cells]}].
    cells[ib * n_cells + ia]: bit k = group k free in the cell, bit 16 + k = group k certainly colliding."""
    dim = ir["dimension"]
    if dim < 2:
        return []
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    items = _self_items(fn)
    by_pair: dict[tuple[int, int], list[int]] = {}
    for gi, deps in enumerate(dependencies(ir, fn)):
        if len(deps) == 2:
            key = tuple(sorted(deps))
        elif len(deps) == 1:
            j = next(iter(deps))
            key = (j, j + 1) if j + 1 < dim else (j - 1, j)
        else:
            continue
        by_pair.setdefault(key, []).append(gi)
    tables = []
    for (ja, jb), groups in sorted(by_pair.items()):
        for first in range(0, len(groups), MAX_GROUPS):
            part = groups[first : first + MAX_GROUPS]
            m = n_cells + 1
            ia, ib = np.meshgrid(np.arange(m), np.arange(m), indexing="xy")  # [ib, ia]
            q = np.tile((lo + F(0.5) * rg)[None, :], (m * m, 1)).astype(F)
            # same expression as the kernel's inverse: corner i of joint j sits at lo + i * (range / n_cells)
            q[:, ja] = (lo[ja] + ia.ravel().astype(F) * F(rg[ja] / F(n_cells))).astype(F)
            q[:, jb] = (lo[jb] + ib.ravel().astype(F) * F(rg[jb] / F(n_cells))).astype(F)
            vals = ir_eval.evaluate(fn, q)
            cells = np.zeros((n_cells, n_cells), np.uint32)
            for k, gi in enumerate(part):
                guard, fine = _pairs(items[gi])
                gmn, gmx = _cells(_clear(vals, guard[0], guard[1], m * m).reshape(m, m))
                g_sp = gmx - gmn
                guard_fires = gmx + 0.25 * g_sp + MARGIN < 0.0
                guard_free = gmn - 0.25 * g_sp - MARGIN > 0.0
                if fine:
                    free_lb = np.full((n_cells, n_cells), np.inf)
                    hit_ub = np.full((n_cells, n_cells), np.inf)
                    for a, b in fine:
                        mn, mx = _cells(_clear(vals, a, b, m * m).reshape(m, m))
                        sp = mx - mn
                        free_lb = np.minimum(free_lb, mn - 0.25 * sp - MARGIN)
                        hit_ub = np.minimum(hit_ub, mx + 0.25 * sp + MARGIN)
                    free = guard_free | (free_lb > 0.0)
                    hit = guard_fires & (hit_ub < 0.0)
                else:
                    free, hit = guard_free, guard_fires
                assert not np.any(free & hit)
                cells |= (free.astype(np.uint32) << np.uint32(k)) | (hit.astype(np.uint32) << np.uint32(16 + k))
            tables.append({"joints": (ja, jb), "groups": part, "cells": cells.ravel(), "n": n_cells})
    # a table that decides nothing is not worth a load
    return [t for t in tables if np.any(t["cells"])]


def emit(robot: str, ir: dict[str, Any], tables: list[dict[str, Any]], self_words: int) -> tuple[list[str], list[str]]:
    """(file-scope lines: the cell arrays, struct-scope lines: kPairTables + pair_lookup)."""
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    head: list[str] = []
    body: list[str] = []
    for ti, t in enumerate(tables):
        c = t["cells"]
        head.append(f"// joint-pair table {ti}: joints {t['joints']}, self groups {t['groups']} (codegen/pair_tables.py)")
        head.append(f"__device__ const uint32_t k_{robot}_pair_cells{ti}[{c.size}] = {{")
        for i in range(0, c.size, 32):
            head.append("    " + ",".join(str(int(v)) for v in c[i : i + 32]) + ",")
        head.append("};")
    w = body.append
    w(f"    static constexpr int kPairTables = {len(tables)};  // joint-pair tables of self groups that depend on two joints only")
    w("    // bit g of free_[]: self group g cannot collide at q; returns true when a tabulated group certainly collides")
    w("    __device__ __forceinline__ static bool pair_lookup(const float *__restrict__ q, uint32_t *__restrict__ free_)")
    w("    {")
    for wi in range(self_words):
        w(f"        uint32_t w{wi} = 0u;")
    w("        uint32_t hit = 0u;")
    for ti, t in enumerate(tables):
        n = t["n"]
        ja, jb = t["joints"]
        inv_a, inv_b = F(F(n) / rg[ja]), F(F(n) / rg[jb])
        w("        {")
        w(f"            const float fa = (q[{ja}] - {float(lo[ja])!r}f) * {float(inv_a)!r}f, fb = (q[{jb}] - {float(lo[jb])!r}f) * {float(inv_b)!r}f;")
        w(f"            const bool in = fa >= 0.f && fa < {n}.f && fb >= 0.f && fb < {n}.f;")
        w(f"            const int ia = min(max(static_cast<int>(fa), 0), {n - 1}), ib = min(max(static_cast<int>(fb), 0), {n - 1});")
        w(f"            const uint32_t cell = in ? __ldg(k_{robot}_pair_cells{ti} + ib * {n} + ia) : 0u;")
        w("            hit |= cell >> 16;")
        for k, g in enumerate(t["groups"]):
            w(f"            w{g // 32} |= ((cell >> {k}) & 1u) << {g % 32};")
        w("        }")
    for wi in range(self_words):
        w(f"        free_[{wi}] = w{wi};")
    w("        return hit != 0u;")
    w("    }")
    return head, body


def classify(table: dict[str, Any], ir: dict[str, Any], q: np.ndarray) -> np.ndarray:
    """uint32 cell of every configuration (0 outside the tabulated rectangle): the kernel's lookup, in numpy."""
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    n = table["n"]
    ja, jb = table["joints"]
    fa = ((q[:, ja].astype(F) - lo[ja]) * F(F(n) / rg[ja])).astype(F)
    fb = ((q[:, jb].astype(F) - lo[jb]) * F(F(n) / rg[jb])).astype(F)
    inside = (fa >= 0) & (fa < n) & (fb >= 0) & (fb < n)
    ia = np.clip(np.where(inside, fa, 0).astype(np.int64), 0, n - 1)
    ib = np.clip(np.where(inside, fb, 0).astype(np.int64), 0, n - 1)
    return np.where(inside, table["cells"][ib * n + ia], np.uint32(0)).astype(np.uint32)
