"""The rigid-link reduction of the code generator (codegen/emit_cuda.py: plan_links) replaces stored sphere
centres by constant affine combinations of <= 4 stored centres of the same link, and the pair stage's
distances by the Gram form over the basis edges.  Both are checked here, on FRESH random configurations
(the generator fitted them on its own 1 024 samples), against the numpy interpretation of the robot trace
(the reference's arithmetic, robots/<robot>/fk.hh): rebuilt centres within the generator's tolerance (1e-6 m for one
link, 2.5e-6 m for a body of several links joined by fixed joints, whose traced centres come out of different chains
of float32 operations -- a quarter of the 1e-5 m the centres are allowed), Gram-form squared distances within the
band the kernels hand to the direct formula (vb_walk.cuh: kGramBand)."""
import numpy as np
import pytest

from mr_vamp_b200.codegen import emit_cuda, ir_eval, trace_parser

F = np.float32
GRAM_BAND = 4e-5  # vb_walk.cuh: kGramBand (m^2)


def _emitter(robot):
    ir = trace_parser.parse_robot(robot)
    fn = ir["functions"]["interleaved_sphere_fk"]
    lo, rg = np.array(ir["scale"]["lo"]), np.array(ir["scale"]["range"])
    rng = np.random.default_rng(0)
    q = (lo + rng.random((1024, ir["dimension"])) * rg).astype(F)
    em = emit_cuda._Emitter(fn, (ir_eval.evaluate(fn, q), 1024)).run_checks()
    return ir, fn, em


def _fresh(ir, fn, n, seed):
    lo, rg = np.array(ir["scale"]["lo"]), np.array(ir["scale"]["range"])
    rng = np.random.default_rng(seed)
    q = (lo + rng.random((n, ir["dimension"])) * rg).astype(F)
    return ir_eval.evaluate(fn, q)


@pytest.mark.parametrize("robot,min_derived", [("panda", 20), ("ur5", 4), ("baxter", 8)])
def test_rebuilt_centres_match_the_trace(robot, min_derived):
    ir, fn, em = _emitter(robot)
    assert len(em.derived) >= min_derived
    n = 4000
    vals = _fresh(ir, fn, n, 99)
    by_key = {}
    for it in fn["program"]:
        if it.get("k") == "env":
            for s in it["fine"] + [it["guard"]]:
                by_key[em.key(s)] = s
    slot_key = {slot: key for key, slot in em.slots.items()}
    worst = 0.0
    for d, rec in enumerate(em.derived):
        want = ir_eval.centre(vals, by_key[slot_key[emit_cuda.DERIVED_SLOT | d]], n)
        got = np.zeros((n, 3), F)
        assert all(not (b & emit_cuda.DERIVED_SLOT) for b in rec["basis"])
        assert abs(sum(rec["w"]) - 1.0) < 1e-6
        for w, b in zip(rec["w"], rec["basis"]):
            got = (got + F(w) * ir_eval.centre(vals, by_key[slot_key[b]], n).astype(F)).astype(F)
        worst = max(worst, float(np.abs(got.astype(np.float64) - want).max()))
    print(f"{robot}: {len(em.derived)} rebuilt centres, {em.n_rows} stored rows, worst error {worst:.2e} m")
    assert worst < 1.2 * emit_cuda.BODY_REBUILD_TOL


@pytest.mark.parametrize("robot", ["panda", "fetch"])
def test_gram_form_of_the_pair_distance(robot):
    ir, fn, em = _emitter(robot)
    n = 4000
    vals = _fresh(ir, fn, n, 123)
    by_key = {}
    for it in fn["program"]:
        if it.get("k") == "env":
            for s in it["fine"] + [it["guard"]]:
                by_key[em.key(s)] = s
    slot_key = {slot: key for key, slot in em.slots.items()}
    checked, worst = 0, 0.0
    for g in em.selfg:
        b_slots = [em.refs[g["b"][0] + j][0] for j in range(g["b"][1])]
        der = [sl & (emit_cuda.DERIVED_SLOT - 1) for sl in b_slots if sl & emit_cuda.DERIVED_SLOT]
        if not der:
            continue
        rec0 = em.derived[der[0]]
        P = [ir_eval.centre(vals, by_key[slot_key[b]], n).astype(F) for b in rec0["basis"]]
        E = [(P[k + 1] - P[0]).astype(F) for k in range(3)]
        gram = np.array(rec0["gram"])
        for i in range(g["a"][1]):
            a_slot = em.refs[g["a"][0] + i][0]
            if a_slot & (emit_cuda.DERIVED_SLOT | emit_cuda.CONST_SLOT):
                continue
            A = ir_eval.centre(vals, by_key[slot_key[a_slot]], n).astype(F)
            u = (A - P[0]).astype(F)
            s0 = (u * u).sum(axis=1, dtype=F)
            q = [F(-2.0) * (u * E[k]).sum(axis=1, dtype=F) for k in range(3)]
            for sl in b_slots:
                if not sl & emit_cuda.DERIVED_SLOT:
                    continue
                w = em.derived[sl & (emit_cuda.DERIVED_SLOT - 1)]["w"][1:4]
                gj = F(sum(w[k] * w[l] * gram[k][l] for k in range(3) for l in range(3)))
                d2 = (s0 + F(w[0]) * q[0] + F(w[1]) * q[1] + F(w[2]) * q[2] + gj).astype(F)
                B = ir_eval.centre(vals, by_key[slot_key[sl]], n)
                want = ((A.astype(np.float64) - B) ** 2).sum(axis=1)
                worst = max(worst, float(np.abs(d2 - want).max()))
                checked += 1
    print(f"{robot}: {checked} (a_i, b_j) pairs over reduced links, worst |d2 error| {worst:.2e} m^2")
    assert checked > 0
    # the kernels decide by the direct formula whenever the Gram form lands within GRAM_BAND of contact
    assert worst < GRAM_BAND / 4


def test_translated_constants_match_the_trace():
    """codegen/emit_cuda.py: LIFT_SLOT -- Fetch's torso / head centres ride on the prismatic torso_lift joint: x and y
    are literals and z is a constant offset from ONE stored centre's z.  Checked on fresh configurations against the
    trace: the rebuilt z within 5e-7 m (float32 rounding of the two traced sums), and the rows really are saved."""
    ir = trace_parser.parse_robot("fetch")
    fn = ir["functions"]["interleaved_sphere_fk"]
    lo, rg = np.array(ir["scale"]["lo"]), np.array(ir["scale"]["range"])
    rng = np.random.default_rng(0)
    q = (lo + rng.random((1024, ir["dimension"])) * rg).astype(F)
    em = emit_cuda._Emitter(fn, (ir_eval.evaluate(fn, q), 1024), False, True).run_checks()
    plain = emit_cuda._Emitter(fn, (ir_eval.evaluate(fn, q), 1024), False, False).run_checks()
    assert len(em.lift_xyz) >= 30 and em.lift_ref is not None
    assert em.n_rows == plain.n_rows - len(em.lift_xyz)
    n = 4000
    vals = _fresh(ir, fn, n, 321)
    by_key = {}
    for it in fn["program"]:
        if it.get("k") == "env":
            for s in it["fine"] + [it["guard"]]:
                by_key[em.key(s)] = s
        elif it.get("k") == "self":
            for s in it["guard"]:
                by_key[em.key(s)] = s
            for a, b in it.get("fine", []):
                by_key[em.key(a)] = a
                by_key[em.key(b)] = b
    slot_key = {slot: key for key, slot in em.slots.items()}
    zref = vals[em.lift_ref[1]].astype(F)
    worst = 0.0
    for slot, (x, y, dz) in em.lift_xyz.items():
        s = by_key[slot_key[slot]]
        want = ir_eval.centre(vals, s, n)
        assert s[0]["v"] is None and s[1]["v"] is None and s[2]["v"] is not None
        got_z = (F(float(dz.rstrip("f"))) + zref).astype(F)
        worst = max(worst, float(np.abs(got_z.astype(np.float64) - want[:, 2]).max()))
        assert abs(float(x.rstrip("f")) - s[0]["c"]) < 1e-7 and abs(float(y.rstrip("f")) - s[1]["c"]) < 1e-7
    print(f"fetch: {len(em.lift_xyz)} translated constants, rows {plain.n_rows} -> {em.n_rows}, worst z error {worst:.2e} m")
    assert worst < 5e-7
    # robots without a prismatic joint have none
    for robot in ("panda", "ur5"):
        ir2 = trace_parser.parse_robot(robot)
        fn2 = ir2["functions"]["interleaved_sphere_fk"]
        lo2, rg2 = np.array(ir2["scale"]["lo"]), np.array(ir2["scale"]["range"])
        q2 = (lo2 + rng.random((256, ir2["dimension"])) * rg2).astype(F)
        assert not emit_cuda._Emitter(fn2, (ir_eval.evaluate(fn2, q2), 256), True, True).run_checks().lift_xyz
