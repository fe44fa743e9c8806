"""The joint-pair tables of the code generator (codegen/pair_tables.py) only move work: a `free` cell must
never contain a colliding configuration of its group, a `hit` cell never a collision-free one.  Checked
here against the numpy interpretation of the robot trace over random configurations (all joints random:
the joints outside the pair perturb the distances by the ~3e-5 m the table margin has to absorb)."""
import numpy as np
import pytest

from mr_vamp_b200.codegen import ir_eval, pair_tables, trace_parser

F = np.float32


@pytest.mark.parametrize("robot,expect_groups", [("panda", 4), ("ur5", 5)])
def test_pair_table_cells_are_sound(robot, expect_groups):
    ir = trace_parser.parse_robot(robot)
    fn = ir["functions"]["interleaved_sphere_fk"]
    tables = pair_tables.build(ir, fn)
    assert sum(len(t["groups"]) for t in tables) == expect_groups
    items = [it for it in fn["program"] if it["k"] == "self"]
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    rng = np.random.default_rng(7)
    n = 120_000
    q = (lo + rng.random((n, ir["dimension"])).astype(F) * rg).astype(F)
    # a good share of the samples right at the edges of the rectangle and slightly outside it
    q[: n // 20] = np.where(rng.random((n // 20, ir["dimension"])) < 0.5, lo, lo + rg).astype(F)
    q[n // 20 : n // 10] += F(1e-3)
    vals = ir_eval.evaluate(fn, q)
    decided = 0
    for t in tables:
        cell = pair_tables.classify(t, ir, q)
        for k, gi in enumerate(t["groups"]):
            guard, fine = pair_tables._pairs(items[gi])
            g = pair_tables._clear(vals, guard[0], guard[1], n)
            f = np.full(n, -np.inf) if not fine else np.min([pair_tables._clear(vals, a, b, n) for a, b in fine], axis=0)
            collides = (g < 0) & (f < 0)
            free = ((cell >> np.uint32(k)) & 1).astype(bool)
            hit = ((cell >> np.uint32(16 + k)) & 1).astype(bool)
            assert not np.any(free & collides), (robot, items[gi]["name"])
            assert not np.any(hit & ~collides), (robot, items[gi]["name"])
            # the margin is not eaten up: decided cells stay well away from contact
            worst = np.maximum(g, f)
            if np.any(free):
                assert worst[free].min() > 0.25 * pair_tables.MARGIN
            if np.any(hit):
                assert worst[hit].max() < -0.25 * pair_tables.MARGIN
            decided += int(free.sum() + hit.sum())
    assert decided > 0.9 * n * expect_groups * 0.9  # tables decide nearly everything inside the joint ranges


def test_lookup_outside_the_joint_ranges_is_undecided():
    ir = trace_parser.parse_robot("panda")
    t = pair_tables.build(ir, ir["functions"]["interleaved_sphere_fk"])[0]
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    q = np.tile(lo + F(0.5) * rg, (4, 1)).astype(F)
    ja, jb = t["joints"]
    q[0, ja] = lo[ja] - F(0.01)
    q[1, jb] = lo[jb] + rg[jb] + F(0.01)
    q[2, ja] = np.nan
    cell = pair_tables.classify(t, ir, q)
    assert cell[0] == 0 and cell[1] == 0 and cell[2] == 0 and cell[3] != 0


@pytest.mark.parametrize("robot", ["panda", "ur5", "fetch", "baxter"])
def test_derived_interpolation_bound_holds(robot):
    """codegen/pair_tables.py: interpolation_bound -- the margin of a table is derived from the error bound of bilinear
    interpolation with second-derivative maxima taken over the table's corner grid.  Checked here the other way round:
    at random points INSIDE the cells the pair clearance never leaves [min corner - eps, max corner + eps]."""
    ir = trace_parser.parse_robot(robot)
    fn = ir["functions"]["interleaved_sphere_fk"]
    tables = pair_tables.build(ir, fn)
    items = [it for it in fn["program"] if it["k"] == "self"]
    lo, rg = np.array(ir["scale"]["lo"], F), np.array(ir["scale"]["range"], F)
    rng = np.random.default_rng(11)
    n_cells = pair_tables.N_CELLS
    for t in tables:
        assert 0.0 < t["eps"] and max(pair_tables.MARGIN, 1.1 * t["eps"] + pair_tables.RESIDUE) <= pair_tables.MAX_MARGIN
        ja, jb = t["joints"]
        m = 20_000
        ca, cb = rng.integers(0, n_cells, m), rng.integers(0, n_cells, m)
        ua, ub = rng.random(m), rng.random(m)

        def configs(fa, fb):
            q = np.tile((lo + F(0.5) * rg)[None, :], (m, 1)).astype(F)
            q[:, ja] = (lo[ja] + (ca + fa).astype(F) * F(rg[ja] / F(n_cells))).astype(F)
            q[:, jb] = (lo[jb] + (cb + fb).astype(F) * F(rg[jb] / F(n_cells))).astype(F)
            return q

        inner = ir_eval.evaluate(fn, configs(ua, ub))
        corners = [ir_eval.evaluate(fn, configs(np.full(m, fa), np.full(m, fb))) for fa in (0.0, 1.0) for fb in (0.0, 1.0)]
        for gi in t["groups"]:
            guard, fine = pair_tables._pairs(items[gi])
            for a, b in [guard] + fine:
                f = pair_tables._clear(inner, a, b, m)
                c = np.stack([pair_tables._clear(v, a, b, m) for v in corners])
                over = np.maximum(f - c.max(axis=0), c.min(axis=0) - f)
                # the bound is derived for |p| >= R_eff = max(R - 1 cm, R / 2), i.e. f >= -min(1 cm, R / 2): where the sign
                # of f is at stake.  Deeper inside (both spheres far into each other) only the sign is asserted.
                R = float(F(a[3]["c"])) + float(F(b[3]["c"]))
                near = c.min(axis=0) >= -min(1.0e-2, 0.5 * R)
                if near.any():
                    assert over[near].max() <= t["eps"] + 1e-6, (robot, t["joints"], gi, over[near].max(), t["eps"])
                deep = c.max(axis=0) < -min(1.0e-2, 0.5 * R)
                assert np.all(f[deep] < 0.0)
