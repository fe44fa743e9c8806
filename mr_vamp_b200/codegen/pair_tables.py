"""Joint-pair tables for self-collision groups whose geometry depends on two joints only.

A self-collision group of the reference (robots/<robot>/fk.hh: `if (sphere_sphere_self_collision(ga, gb))
{ a_i x b_j ... }`) tests spheres of two links against each other.  When the links are separated by one or
two joints, every distance in the group is a function of those joints alone (e.g. Panda link5 against
link7 / hand / fingers: joints 6 and 7).  For such groups the verdict over the whole (qa, qb) rectangle
of the joint ranges is tabulated at code-generation time:

    free    no pair of the group overlaps anywhere in the cell (the group is skipped),
    hit     the guard pair and one fine pair overlap everywhere in the cell (the configuration collides),
    unsure  anything else: the kernel runs the group's tests exactly as the reference does.

`free` and `hit` are decided from the clearances f = |p| - R (p = c_a - c_b, R = r_a + r_b) at the four cell
corners with a margin; cells the margin cannot decide stay `unsure`, so the table never changes a verdict, only
who computes it.

The margin, derived (interpolation_bound below, asserted for every table at generation time).  Inside a cell the
bilinear interpolant of the corner values lies between the smallest and the largest corner value, and f differs
from it by at most

    eps = (h_a^2 M_aa + h_b^2 M_bb) / 8 + h_a h_b M_ab / 4,     M_ij >= sup |d2 f / dq_i dq_j|

(the standard error bound of bilinear interpolation on an h_a x h_b cell).  For f = |p| - R,
d2f/dq_i dq_j = (p_i . p_j + p . p_ij) / |p| - (p . p_i)(p . p_j) / |p|^3, hence |d2f| <= |p_ij| + |p_i| |p_j| / |p|.
The derivative maxima |p_i|, |p_ij| are taken by finite differences over the (N_CELLS + 1)^2 corner grid the table
is built from anyway (smooth trigonometric polynomials sampled at the cell pitch; inflated by 10 %), and |p| is
bounded below by R_eff = max(R - 1 cm, R / 2): the sign of f is only at stake where |p| is within the margin of R
(where the corners have |p| < R_eff, |p| stays below R inside the cell by the Lipschitz bound |p_i| h_i << R / 2).
Every pair is then classified with margin_g = max(MARGIN, 1.1 eps_g + RESIDUE) (+ a quarter of the corner spread,
kept from the first version as slack), RESIDUE covering what the model leaves out: the traced link frames are not
perfectly rigid (polynomial sin / cos: distances move by up to ~3e-5 m with joints that are not between the two
links, see dependencies()) and this interpreter and the kernels round differently (~1e-6 m).  Bounds of the tables built today:
Panda (5, 6) 2.1e-4 m, UR5 (1, 2) 1.2e-3 (margin 1.35e-3) / (2, 3) 2.1e-4, Fetch 4.2e-5 / 6.9e-5, Baxter 1.3e-4.
tests/test_oracle_cpu.py re-checks random points against the cells.
"""
from __future__ import annotations

from typing import Any

import numpy as np

from . import ir_eval

N_CELLS = 256  # cells per joint
MARGIN = 1.0e-3  # metres (floor of the per-pair margin)
RESIDUE = 5.0e-5  # metres: non-rigidity of the traced frames + rounding differences (see the module docstring)
MAX_MARGIN = 5.0e-3  # a table whose derived margin exceeds this is not worth building: assert
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


def interpolation_bound(vals, a, b, m: int, ha: float, hb: float) -> float:
    """Upper bound (metres) of |f - bilinear interpolant of the corner values| inside any cell, for the pair (a, b);
    `vals` = the trace evaluated on the m x m corner grid ([ib, ia] order).  See the module docstring."""
    p = (ir_eval.centre(vals, a, m * m) - ir_eval.centre(vals, b, m * m)).reshape(m, m, 3)
    R = float(F(a[3]["c"])) + float(F(b[3]["c"]))
    r_eff = max(R - 1.0e-2, 0.5 * R)
    nrm = lambda d: float(np.linalg.norm(d, axis=2).max())
    da, db = 1.1 * nrm(np.diff(p, axis=1)) / ha, 1.1 * nrm(np.diff(p, axis=0)) / hb
    daa, dbb = 1.1 * nrm(np.diff(p, 2, axis=1)) / ha**2, 1.1 * nrm(np.diff(p, 2, axis=0)) / hb**2
    dab = 1.1 * nrm(np.diff(np.diff(p, axis=0), axis=1)) / (ha * hb)
    maa, mbb, mab = daa + da * da / r_eff, dbb + db * db / r_eff, dab + da * db / r_eff
    return (ha * ha * maa + hb * hb * mbb) / 8.0 + ha * hb * mab / 4.0


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
            ha, hb = float(rg[ja]) / n_cells, float(rg[jb]) / n_cells
            worst = 0.0
            for k, gi in enumerate(part):
                guard, fine = _pairs(items[gi])

                def margin_of(a, b):
                    nonlocal worst
                    eps = interpolation_bound(vals, a, b, m, ha, hb)
                    worst = max(worst, eps)
                    mg = max(MARGIN, 1.1 * eps + RESIDUE)
                    assert mg <= MAX_MARGIN, (ir["robot"], (ja, jb), gi, eps)
                    return mg

                g_mg = margin_of(guard[0], guard[1])
                gmn, gmx = _cells(_clear(vals, guard[0], guard[1], m * m).reshape(m, m))
                g_sp = gmx - gmn
                guard_fires = gmx + 0.25 * g_sp + g_mg < 0.0
                guard_free = gmn - 0.25 * g_sp - g_mg > 0.0
                if fine:
                    free_lb = np.full((n_cells, n_cells), np.inf)
                    hit_ub = np.full((n_cells, n_cells), np.inf)
                    for a, b in fine:
                        mg = margin_of(a, b)
                        mn, mx = _cells(_clear(vals, a, b, m * m).reshape(m, m))
                        sp = mx - mn
                        free_lb = np.minimum(free_lb, mn - 0.25 * sp - mg)
                        hit_ub = np.minimum(hit_ub, mx + 0.25 * sp + mg)
                    free = guard_free | (free_lb > 0.0)
                    hit = guard_fires & (hit_ub < 0.0)
                else:
                    free, hit = guard_free, guard_fires
                assert not np.any(free & hit)
                cells |= (free.astype(np.uint32) << np.uint32(k)) | (hit.astype(np.uint32) << np.uint32(16 + k))
            tables.append({"joints": (ja, jb), "groups": part, "cells": cells.ravel(), "n": n_cells, "eps": worst})
    # a table that decides nothing is not worth a load
    return [t for t in tables if np.any(t["cells"])]


def emit(robot: str, ir: dict[str, Any], tables: list[dict[str, Any]], self_words: int) -> tuple[list[str], list[str]]:
    """(file-scope lines: the cell arrays, struct-scope lines: kPairTables + pair_lookup)."""
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    head: list[str] = []
    body: list[str] = []
    for ti, t in enumerate(tables):
        c = t["cells"]
        head.append(f"// joint-pair table {ti}: joints {t['joints']}, self groups {t['groups']} (codegen/pair_tables.py); derived "
                    f"interpolation-error bound {t.get('eps', 0.0):.2e} m, margin max({MARGIN:g}, 1.1 eps + {RESIDUE:g}) m")
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
