"""Back end of the kernel compiler: robot IR (trace_parser) -> CUDA device code for sm_100a.

For every robot one header ``mr_vamp_b200/_gen/<robot>_trace.cuh`` is written.  The reference
interleaves the FK chain with ~800 inlined collision tests in one straight-line function; that
shape is hostile to a GPU (hundreds of KB of SASS, instruction-cache thrash, minutes of ptxas), so
the trace is split in two:

* ``chain(P&, q)`` / ``chain_attach(P&, q)`` — the FK arithmetic only, straight-line, value by
  value as in ``interleaved_sphere_fk[_attachment]`` (literals narrowed to float exactly where the
  C++ expression narrows them: this keeps sphere centres within 1e-5 m of the reference).  Every
  distinct sphere centre that any check uses is handed to ``p.put(slot, x, y, z)`` (robot-local
  frame); the kernels keep those slots in shared memory, one column per thread.
* constant-memory *check tables* (``vb_tables.cuh``): environment groups (guard sphere + fine
  spheres), self-collision groups (guard pair + a cross product of fine spheres) and attachment
  groups, each entry a (slot, base-offset mask, radius) reference.  The generic loops of
  ``vb_kernels.cuh`` walk these tables, so there is ONE copy of the environment query in the
  kernel instead of 72.
* ``sphere_fk(P&, q)`` <- ``sphere_fk`` (all collision spheres, world frame), unchanged.

Verdicts do not depend on the order of the tests (valid = no group has guard-hit and fine-hit), so
de-interleaving changes work, not results.  Scheduling, register allocation and FMA contraction
are left to nvcc/ptxas.

The generated headers are build products (git-ignored): the traces belong to the VAMP checkout the
library is built against (``$VAMP_SRC``), exactly like the reference's own ``fk.hh`` files belong
to its robot-description pipeline.
"""
from __future__ import annotations

import os
import struct
import sys
from typing import Any

from . import ir_eval, pair_tables, trace_parser

_HERE = os.path.dirname(os.path.abspath(__file__))
GEN_DIR = os.path.join(os.path.dirname(_HERE), "_gen")

ROBOTS = ["sphere", "ur5", "panda", "fetch", "baxter"]


def _f32(c: float) -> str:
    """C literal of float(c) (round-to-nearest narrowing of the double literal)."""
    f = struct.unpack("<f", struct.pack("<f", c))[0]
    if f == 0.0:
        return "0.0f"
    return f"{f!r}f" if ("e" in repr(f) or "." in repr(f)) else f"{f!r}.0f"


def _f64(c: float) -> str:
    return repr(float(c))




CONST_SLOT = 0x4000  # slot code of a configuration-independent centre: CONST_SLOT | index


DERIVED_SLOT = 0x2000  # slot code of a centre that is NOT stored: DERIVED_SLOT | index into the derived table
# slot code of a "translated constant": a centre whose x and y are constants and whose z is a constant plus the z of ONE
# stored centre (Fetch: the 38 spheres of the torso and the head ride on the prismatic torso_lift joint).  (x, y, dz) is
# kept once per warp in the rows behind the constants; z = dz + z(reference slot) is added where the centre is read.
# 37 fewer per-thread rows = 7 resident warps instead of 5.
LIFT_SLOT = 0x1000
LIFT_ON = os.environ.get("VB_LIFT_SLOTS", "1") != "0"
# Rigid-link reduction (plan_links) is a per-robot build decision, taken from measurements on B200 (MotionBenchMaker
# table_pick scene 1, 33.5 M uniform configurations, Gcfg/s without -> with): panda 2.44 -> 2.92, ur5 3.02 -> 3.32,
# fetch 0.73 -> 0.46 (its reduced links are on both sides of most of its 465 pair units; rebuilding them costs more
# than the four extra resident warps give back), baxter: no link has more than four spheres.
LINK_REDUCTION_OFF = {"fetch"}
# Two-level guard lookup (coarse 8 cm grid first, fine grid only when the coarse cell cannot call the guard free;
# env_fetch below).  Measured on the MBM scenes: panda 3.29 -> 3.55, ur5 3.33 -> 3.39 Gcfg/s; fetch 0.75 -> 0.70 and
# baxter 0.48 -> 0.46 (instruction-fetch bound: the second level's code costs them more than its loads save).
TWO_LEVEL_GRID_OFF = {"fetch", "baxter"}
# Self-collision a-side filters as QUEUED work (one (configuration, a_i) test per lane, table driven; vb_walk.cuh:
# filter_units) instead of straight-line lane-parallel code per group.  The straight-line form runs a group's filters on
# all 32 lanes whenever one lane's guard fired (Baxter: 174 fired (lane, group) pairs in 81 distinct groups per warp
# step = 2 useful lanes of 32, and 718 units of generated code); the queued form does exactly the fired tests.
QUEUED_FILTERS = set(os.environ.get("VB_QUEUED_FILTERS", "fetch,baxter").split(","))
MIN_LINK_SPHERES = int(os.environ.get("VB_MIN_LINK_SPHERES", "2"))  # links with fewer fine spheres are left alone
DERIVE_GUARDS = os.environ.get("VB_DERIVE_GUARDS", "1") != "0"  # bounding spheres rebuilt from the link basis too
GUARD_BLOCK = int(os.environ.get("VB_GUARD_BLOCK", "0"))  # self-collision guard pairs per scoped block of self_guards (0: one block)
MERGE_BODIES = os.environ.get("VB_MERGE_BODIES", "1") != "0"  # links joined by fixed joints share one basis
# Largest accepted difference between a rebuilt centre and the traced one (float32, over the sampled configurations).
# One link: 1e-6 m.  A body of several links: the traced centres of its links come out of different chains of float32
# operations and are themselves ~1e-6 m off a common affine image (Panda link7 / hand / fingers: 1.9e-6 m); 2.5e-6 m is
# a quarter of the 1e-5 m the centres (and the verdict band) are allowed.
REBUILD_TOL = float(os.environ.get("VB_REBUILD_TOL", "1e-6"))
BODY_REBUILD_TOL = float(os.environ.get("VB_BODY_REBUILD_TOL", "2.5e-6"))


def slot_offset(slot: int, stride: int, const_row: int, lift_row: int = 0) -> int:
    """Float offset of a slot's x coordinate from the warp's first column (y, z follow at +stride each).
    Regular slots add the lane; constant slots (offset >= 3 * const_row * stride) do not; translated constants
    (offset >= 3 * lift_row * stride) add the reference slot's z."""
    if slot & LIFT_SLOT:
        j = slot & (LIFT_SLOT - 1)
        return 3 * (lift_row + j // 32) * stride + j % 32
    if slot & CONST_SLOT:
        j = slot & (CONST_SLOT - 1)
        return 3 * (const_row + j // 32) * stride + j % 32
    return 3 * slot * stride


class _Emitter:
    """Lowers one traced function to (chain lines, check tables)."""

    def __init__(self, fn: dict[str, Any], samples=None, reduce_links: bool = True, use_lift: bool = False):
        self.reduce_links = reduce_links
        self.use_lift = use_lift
        self.values = fn["values"]
        self.program = fn["program"]
        self.samples = samples  # (vals, n) from ir_eval, for the static containment slack
        self.lines: list[str] = []
        self.slots: dict[tuple, int] = {}
        self.const_slots: set[int] = set()
        self.const_xyz: dict[int, list[str]] = {}
        self.n_rows = 0  # per-thread slot rows (configuration-dependent centres)
        self.refs: list[tuple[int, int, float]] = []  # (slot, base mask, radius)
        self.env: list[dict[str, Any]] = []
        self.selfg: list[dict[str, Any]] = []
        self.att: list[dict[str, Any]] = []
        self.att_env = 0
        self.has_pose = False
        # rigid-link reduction: key -> (basis spheres [4 argument lists], weights [4]); see plan_links
        self.plan: dict[tuple, tuple[list, list[float]]] = {}
        self.derived: list[dict[str, Any]] = []  # {"basis": [slot x 4], "w": [float x 4]}
        self.link_weights: dict[tuple, tuple[list, list[float]]] = {}  # stored non-basis spheres of reduced links
        self.slot_weights: dict[int, list[float]] = {}  # their slot -> weights over the link's basis
        self.link_gram: dict[tuple, list[list[float]]] = {}  # basis keys -> Gram matrix of the basis edges
        # translated constants (LIFT_SLOT): slot code -> (x, y, dz) literals; lift_ref = (stored slot, z value id)
        self.lift_xyz: dict[int, list[str]] = {}
        self.lift_ref: tuple[int, int] | None = None
        self.last_vid = -1  # last value the chain has emitted
        self.pending_puts: dict[int, list[str]] = {}  # value id -> stores waiting for it

    # -- operands -----------------------------------------------------------------------------
    def opnd(self, o) -> str:
        if isinstance(o, dict):
            return _f32(o["c"])
        return f"v{o}"

    def coord(self, a: dict[str, Any]) -> str:
        return f"v{a['v']}" if a["v"] is not None else _f32(a["c"])

    def out_arg(self, a: dict[str, Any]) -> str:
        base = a.get("base")
        if a.get("sym") == "radius":
            return "p.radius"
        if a["v"] is not None:
            return f"(v{a['v']} + p.b{base}())" if base else f"v{a['v']}"
        if base:
            return f"p.cb{base}({_f64(a['c'])})"
        return _f32(a["c"])

    def ref(self, s) -> tuple[int, int, float]:
        """Sphere argument list [x, y, z, r] -> (slot, base mask, radius); emits the put on first use."""
        key = self.key(s)
        if key not in self.slots and key in self.plan:
            basis, wts = self.plan[key]
            self.slots[key] = DERIVED_SLOT | len(self.derived)
            self.derived.append({"basis": [self.ref(b)[0] for b in basis], "w": list(wts),
                                 "gram": self.link_gram[tuple(self.key(b) for b in basis)]})
        if key not in self.slots and self.lift_candidate(s):
            zv = s[2]["v"]
            if self.lift_ref is not None:
                import numpy as np

                vals, _ = self.samples
                d = vals[zv].astype(np.float64) - vals[self.lift_ref[1]].astype(np.float64)
                dz = float(d.mean())
                if float(np.abs(d - dz).max()) < 5e-7:  # (float32 rounding of the two traced sums: ~1.2e-7 m on Fetch)
                    j = len(self.lift_xyz)
                    self.slots[key] = LIFT_SLOT | j
                    self.lift_xyz[self.slots[key]] = [self.coord(s[0]), self.coord(s[1]), _f32(dz)]
                    self.lines.append(f"p.put_c<kStride>(kLiftRow + {j // 32}, {j % 32}, {self.coord(s[0])}, {self.coord(s[1])}, {_f32(dz)});")
        if key not in self.slots:
            xyz = [self.coord(a) for a in s[:3]]
            if all(a["v"] is None for a in s[:3]):
                # configuration-independent centre: ONE copy per warp in the constant rows (see vb_sinks.cuh)
                j = len(self.const_slots)
                self.slots[key] = CONST_SLOT | j
                self.const_slots.add(self.slots[key])
                self.const_xyz[self.slots[key]] = xyz
                self.lines.append(f"p.put_c<kStride>(kConstRow + {j // 32}, {j % 32}, {xyz[0]}, {xyz[1]}, {xyz[2]});")
            else:
                self.slots[key] = self.n_rows
                self.n_rows += 1
                if self.lift_ref is None and self.lift_candidate(s):
                    self.lift_ref = (self.slots[key], s[2]["v"])  # the stored centre the translated constants ride on
                if key in self.link_weights:
                    self.slot_weights[self.slots[key]] = list(self.link_weights[key][1])
                put = f"p.put_s<kStride>({self.slots[key]}, {xyz[0]}, {xyz[1]}, {xyz[2]});"
                # a basis centre of a body of several links may be referenced (by a rebuilt centre of an earlier link)
                # before the chain has computed it: its store waits for the last of its values
                need = max(a["v"] for a in s[:3] if a["v"] is not None)
                if need <= self.last_vid:
                    self.lines.append(put)
                else:
                    self.pending_puts.setdefault(need, []).append(put)
        mask = sum((1 << i) for i, a in enumerate(s[:3]) if a.get("base"))
        for i, a in enumerate(s[:3]):
            assert a.get("base") in (None, "xyz"[i]), s
        r = s[3]
        radius = -1.0 if r.get("sym") == "radius" else float(r["c"])
        assert r.get("sym") == "radius" or r["v"] is None
        return (self.slots[key], mask, radius)

    def lift_candidate(self, s) -> bool:
        """x and y literal, z traced, no base offset, not part of a reduced link: may become a translated constant"""
        return (self.use_lift and self.samples is not None and s[0]["v"] is None and s[1]["v"] is None and s[2]["v"] is not None
                and not any(a.get("base") for a in s[:3]) and self.key(s) not in self.plan and self.key(s) not in self.link_weights)

    @staticmethod
    def key(s) -> tuple:
        return tuple((a["v"] if a["v"] is not None else ("c", a["c"])) for a in s[:3])

    def plan_links(self):
        """Rigid-link reduction of the shared-memory columns.  The fine spheres of one environment group sit on
        one rigid link -- and links joined by fixed joints form one rigid BODY (Panda: link7, the hand and both fingers;
        found from the samples, MERGE_BODIES) -- so each centre is a constant affine combination of (at most) four of
        the body's centres.  For links
        with >= MIN_LINK_SPHERES fine spheres only that basis is stored by the chain; the others are rebuilt
        where a fired group needs them (<= 1e-6 m from the traced value, checked here over the sampled
        configurations in float32), and their chain arithmetic becomes dead code.  Fewer slot rows = more
        resident warps per SM, which is what bounds the fkcc kernels."""
        if self.samples is None or not self.reduce_links:
            return self
        import numpy as np

        vals, n = self.samples
        # guard spheres are read by every configuration (straight-line guard phase): always stored
        guards = set()
        for it in self.program:
            if it.get("k") in ("env", "attach_self"):
                guards.add(self.key(it["guard"]))
            elif it.get("k") == "self":
                guards.add(self.key(it["guard"][0]))
                guards.add(self.key(it["guard"][1]))
        def eligible(it) -> bool:
            fine = it["fine"]
            if len(fine) < MIN_LINK_SPHERES or fine == [it["guard"]]:
                return False
            if any(all(a["v"] is None for a in s[:3]) for s in fine):
                return False  # configuration-independent centres live in the constant rows
            if self.use_lift and all(s[0]["v"] is None and s[1]["v"] is None for s in fine):
                return False  # translated constants (LIFT_SLOT) are cheaper than rebuilt centres
            return len({tuple(a.get("base") for a in s[:3]) for s in fine}) == 1

        def plan_set(fine, guard_list, tol=REBUILD_TOL):
            """One rigid body: `fine` = its fine spheres (of one or several environment groups), `guard_list` = the
            bounding spheres of those groups.  Returns (plan, link weights, (basis keys, Gram matrix)) or None when
            some centre cannot be rebuilt to 1e-6 m."""
            C = [ir_eval.centre(vals, s, n) for s in fine]  # float64 copies of the float32 samples
            # greedy basis on the first sample: farthest pair, then farthest from the line, then from the plane
            # (many links carry collinear or coplanar sphere sets: 2 or 3 basis centres then suffice)
            P = np.stack([c[0] for c in C])
            d = np.linalg.norm(P[:, None] - P[None], axis=2)
            i0, i1 = np.unravel_index(np.argmax(d), d.shape)
            basis = [int(i0), int(i1)]
            for _ in range(2):
                if len(basis) >= len(fine):
                    break
                B = P[basis[1:]] - P[basis[0]]
                Q, _ = np.linalg.qr(B.T)
                R = (P - P[basis[0]]) - (P - P[basis[0]]) @ Q @ Q.T
                dist = np.linalg.norm(R, axis=1)
                k = int(np.argmax(dist))
                if dist[k] < 5e-3:
                    break
                basis.append(k)
            m = min(n, 128)
            fine_keys = {self.key(sj) for sj in fine}

            def with_basis(basis):
                A = np.concatenate([np.stack([C[b][t] - C[basis[0]][t] for b in basis[1:]], axis=1) for t in range(m)])

                def weights(Cj):
                    """Affine weights of a centre over the basis and the float32 reconstruction error over all samples."""
                    y = np.concatenate([Cj[t] - C[basis[0]][t] for t in range(m)])
                    wk, *_ = np.linalg.lstsq(A, y, rcond=None)
                    wts = [1.0 - float(wk.sum())] + [float(v) for v in wk]
                    rec = np.zeros((n, 3), np.float32)
                    for wv, b in zip(wts, basis):
                        rec = (rec + np.float32(wv) * C[b].astype(np.float32)).astype(np.float32)
                    err = float(np.abs(rec.astype(np.float64) - Cj).max())
                    return wts + [0.0] * (4 - len(wts)), err

                basis_args = [fine[b] for b in basis] + [fine[basis[0]]] * (4 - len(basis))
                # Gram matrix of the basis edges e_k = c_bk - c_b0 (a rigid-motion invariant): lets the pair stage
                # evaluate |a - b_j|^2 = |a - c_b0|^2 - 2 sum_k w_k (a - c_b0).e_k + w^T G w with w^T G w a constant
                E = [C[b] - C[basis[0]] for b in basis[1:]] + [np.zeros_like(C[0])] * (4 - len(basis))
                gram = [[float((E[k] * E[l]).sum(axis=1).mean()) for l in range(3)] for k in range(3)]
                plan, lw = {}, {}
                for j, sj in enumerate(fine):
                    if j in basis:
                        continue
                    wts, err = weights(C[j])
                    if err > tol:
                        return None
                    if self.key(sj) in guards:
                        lw[self.key(sj)] = (basis_args, wts)  # stays stored; the weights serve the b-side loop of the pair stage
                    else:
                        plan[self.key(sj)] = (basis_args, wts)
                # the links' bounding spheres (guards of these groups and of their self-collision groups) are rigid with
                # the body too: rebuilt from the basis where they lie in its affine span
                for g in guard_list:
                    gkey = self.key(g)
                    if (DERIVE_GUARDS and gkey not in fine_keys and not all(a["v"] is None for a in g[:3])
                            and gkey not in self.plan and gkey not in plan):
                        wts, err = weights(ir_eval.centre(vals, g, n))
                        if err <= tol:
                            plan[gkey] = (basis_args, wts)
                return plan, lw, (tuple(self.key(b) for b in basis_args), gram)

            res = with_basis(basis)
            if res is None and len(fine) > len(basis):
                # the farthest-point basis is not always the best conditioned one for float32: try others of the same size
                pick = np.random.default_rng(len(fine))
                for _ in range(48):
                    cand = [int(v) for v in pick.choice(len(fine), size=len(basis), replace=False)]
                    B = P[cand[1:]] - P[cand[0]]
                    if np.linalg.svd(B, compute_uv=False)[-1] < 2e-2:
                        continue  # (nearly) degenerate simplex
                    res = with_basis(cand)
                    if res is not None:
                        break
            return res

        def commit(res) -> bool:
            if res is None or not res[0]:
                return False
            plan, lw, (bkeys, gram) = res
            self.link_gram[bkeys] = gram
            self.plan.update(plan)
            self.link_weights.update(lw)
            return True

        # Rigid BODIES first: environment groups whose spheres keep their mutual distances over the samples sit on one
        # rigid body (Panda: link7, the hand and the two fingers -- fixed joints between them), and one basis of four
        # centres serves all of them: 15 stored rows become 4.
        groups = [it for it in self.program if it.get("k") == "env" and eligible(it)]

        def union(members):
            fine, seen = [], set()
            for i in members:
                for sj in groups[i]["fine"]:
                    if self.key(sj) not in seen:
                        seen.add(self.key(sj))
                        fine.append(sj)
            return fine

        def rigid(ga, gb) -> bool:
            """candidate test only (plan_set decides): a few centre pairs keep their distance over the samples"""
            pa = [ga["fine"][0], ga["fine"][-1]]
            pb = [gb["fine"][0], gb["fine"][-1]]
            if any(tuple(a.get("base") for a in x[:3]) != tuple(a.get("base") for a in pa[0][:3]) for x in pb):
                return False
            pb.append(gb["guard"])  # (local-frame values: the base offset is a property of the reference to a centre)
            for x in pa:
                cx = ir_eval.centre(vals, x, n)
                for y in pb:
                    dd = np.linalg.norm(cx - ir_eval.centre(vals, y, n), axis=1)
                    if float(dd.max() - dd.min()) > 2e-4:
                        return False
            return True

        bodies: list[list[int]] = []
        for i in range(len(groups)):
            home = None
            if MERGE_BODIES:
                for body in bodies:
                    if rigid(groups[body[0]], groups[i]) and \
                            plan_set(union(body + [i]), [groups[k]["guard"] for k in body + [i]], BODY_REBUILD_TOL) is not None:
                        home = body
                        break
            if home is None:
                bodies.append([i])
            else:
                home.append(i)
        for body in bodies:
            if len(body) > 1 and commit(plan_set(union(body), [groups[k]["guard"] for k in body], BODY_REBUILD_TOL)):
                continue
            if self.reduce_links == "bodies":
                continue  # (development: only bodies of several links are reduced)
            for i in body:
                commit(plan_set(groups[i]["fine"], [groups[i]["guard"]]))
        return self

    def slack(self, fine_spheres, guard) -> float:
        """Largest overhang of the fine spheres beyond the guard sphere (+1e-4 m), or -1 if unknown."""
        if self.samples is None or not fine_spheres:
            return -1.0
        vals, n = self.samples
        if guard[3].get("sym") or any(s[3].get("sym") for s in fine_spheres):
            return -1.0
        worst = 0.0
        cg = ir_eval.centre(vals, guard, n)
        for s in fine_spheres:
            if [a.get("base") for a in s[:3]] != [a.get("base") for a in guard[:3]]:
                return -1.0
            d = ((ir_eval.centre(vals, s, n) - cg) ** 2).sum(axis=1) ** 0.5
            # The bound holds beyond the samples because the two centres sit on ONE rigid link: their distance is a
            # constant of the robot, which the samples show (spread below the ~3e-5 m non-rigidity of the polynomial
            # sin / cos).  A sphere that moves relative to its guard gets no containment bound at all.
            self.rigid_spread = max(getattr(self, "rigid_spread", 0.0), float(d.max() - d.min()))
            if float(d.max() - d.min()) > 5e-5:
                return -1.0
            worst = max(worst, float(d.max()) + s[3]["c"] - guard[3]["c"])
        return max(worst, 0.0) + 1e-4

    def fire_rate(self, ga, gb) -> float:
        """Fraction of the sampled configurations whose guard pair (ga, gb) overlaps."""
        if self.samples is None:
            return 0.0
        vals, n = self.samples
        d = ((ir_eval.centre(vals, ga, n) - ir_eval.centre(vals, gb, n)) ** 2).sum(axis=1) ** 0.5
        return float((d < ga[3]["c"] + gb[3]["c"]).mean())

    def add_refs(self, refs) -> tuple[int, int]:
        begin = len(self.refs)
        self.refs.extend(refs)
        return begin, len(refs)

    # -- statements ---------------------------------------------------------------------------
    def value(self, vid: int):
        op, a, b = self.values[vid]
        if op == "in":
            rhs = f"q[{a}]"
        elif op in ("mul", "add", "sub"):
            rhs = f"{self.opnd(a)} {'*' if op == 'mul' else '+' if op == 'add' else '-'} {self.opnd(b)}"
        elif op == "neg":
            rhs = f"-{self.opnd(a)}"
        elif op in ("sin", "cos"):
            rhs = f"vamp_{op}({self.opnd(a)})"
        else:
            raise ValueError(op)
        self.lines.append(f"const float v{vid} = {rhs};")
        self.last_vid = max(self.last_vid, vid)
        for put in self.pending_puts.pop(vid, []):
            self.lines.append(put)

    def group(self, it: dict[str, Any]):
        k = it["k"]
        if k == "attach_pose":
            self.lines.append("p.attach_pose(" + ", ".join(self.coord(a) for a in it["args"]) + ");")
            self.has_pose = True
            return
        if k == "attach_env":
            self.att_env = 1
            return
        fine = it["fine"]
        leaf = bool(it.get("leaf")) or not fine or fine == [it["guard"]]
        name = it.get("name", "")
        if k in ("env", "attach_self"):
            guard = self.ref(it["guard"])
            begin, count = self.add_refs([] if leaf else [self.ref(s) for s in fine])
            assert count <= 32, name
            (self.env if k == "env" else self.att).append(
                {"guard": guard, "begin": begin, "count": count, "name": name,
                 "slack": -1.0 if leaf or k != "env" else self.slack(fine, it["guard"])}
            )
            return
        if k == "self":
            ga, gb = self.ref(it["guard"][0]), self.ref(it["guard"][1])
            A: list[tuple] = []
            B: list[tuple] = []
            pairs = []
            if not leaf:
                for a, b in fine:
                    ra, rb = self.ref(a), self.ref(b)
                    pairs.append((ra, rb))
                    if ra not in A:
                        A.append(ra)
                    if rb not in B:
                        B.append(rb)
                # every traced self group is the full cross product of two sphere sets
                if sorted(pairs) != sorted((x, y) for x in A for y in B):
                    raise ValueError("self group is not a cross product: " + name)
            ab, ac = self.add_refs(A)
            bb, bc = self.add_refs(B)
            assert ac <= 32 and bc <= 32, name
            slack_b = -1.0 if leaf else self.slack([b for _, b in fine], it["guard"][1])
            dense = (not leaf) and self.fire_rate(it["guard"][0], it["guard"][1]) > 0.4
            self.selfg.append({"ga": ga, "gb": gb, "a": (ab, ac), "b": (bb, bc), "name": name, "slack": slack_b,
                               "dense": dense})
            return
        raise ValueError(k)

    def run_checks(self):
        self.plan_links()
        for it in self.program:
            if it["k"] == "val":
                self.value(it["id"])
            elif it["k"] == "out":
                raise ValueError("out in check trace")
            else:
                self.group(it)
        assert not self.pending_puts, "a stored centre was never computed"
        return self

    def run_spheres(self) -> list[str]:
        for it in self.program:
            if it["k"] == "val":
                self.value(it["id"])
            elif it["k"] == "out":
                self.lines.append(f"p.out_{it['field']}({it['idx']}, {self.out_arg(it['arg'])});")
            else:
                raise ValueError(it["k"])
        return self.lines


def trace_stats(ir: dict[str, Any]) -> dict[str, Any]:
    """Static flop/test counts of the check trace (the W bookkeeping of DESIGN.md)."""
    f = ir["functions"]["interleaved_sphere_fk"]
    ops = {"mul": 0, "add": 0, "sub": 0, "neg": 0, "sin": 0, "cos": 0, "in": 0}
    for v in f["values"]:
        ops[v[0]] += 1
    env_g = [it for it in f["program"] if it["k"] == "env"]
    self_g = [it for it in f["program"] if it["k"] == "self"]
    return {
        "chain_flops": ops["mul"] + ops["add"] + ops["sub"] + 11 * ops["sin"] + 14 * ops["cos"],
        "ops": ops,
        "env_groups": len(env_g),
        "env_fine": sum(len(g["fine"]) for g in env_g),
        "self_groups": len(self_g),
        "self_fine": sum(len(g["fine"]) for g in self_g),
    }


def _ref(r) -> str:
    slot, mask, radius = r
    return f"{{{slot}, {mask}, {_f32(radius) if radius >= 0 else '-1.0f'}}}"


def _slack(v: float) -> str:
    return "-1.0f" if v < 0 else _f32(v)


def _tables(name: str, em: _Emitter, walk: bool = False, stride: int = 0, const_row: int = 0, queued: bool = False,
            lift_row: int = 0) -> list[str]:
    """Constant-memory check tables; with `walk` also the global-memory mirrors and work-unit tables of
    the warp walk (vb_walk.cuh).  Returns (lines, table type, geometry dict)."""
    refs = list(em.refs)
    # work units: environment groups -> their fine refs; a leaf group -> a copy of its guard behind the refs
    ref_group = [0xFFFF] * len(refs)
    for gi, g in enumerate(em.env):
        if g["count"] == 0:
            g["ubegin"], g["ucount"] = len(refs), 1
            refs.append(g["guard"])
            ref_group.append(gi | 0x8000)
        else:
            g["ubegin"], g["ucount"] = g["begin"], g["count"]
            for k in range(g["count"]):
                ref_group[g["begin"] + k] = gi
    import numpy as np

    F = np.float32
    units = []
    # b-side lists of the pair stage, one copy per distinct list (the groups of one link share it): the pair
    # stage is bound by the latency of these reads, and the L1 beside the slot columns is small
    blist_at: dict[tuple, int] = {}
    blist: list[tuple[int, float, list[float]]] = []  # (slot, radius, weights over the link's basis or None)

    def b_list(g) -> int:
        entries = [em.refs[g["b"][0] + j] for j in range(g["b"][1])]
        key = tuple((sl, r) for sl, _, r in entries)
        if key not in blist_at:
            blist_at[key] = len(blist)
            slots = [sl for sl, _, _ in entries]
            der = [sl & (DERIVED_SLOT - 1) for sl in slots if sl & DERIVED_SLOT]
            basis = em.derived[der[0]]["basis"] if der else None
            gram = em.derived[der[0]]["gram"] if der else None
            for sl, _, r in entries:
                wts = None
                if basis is not None:
                    if sl & DERIVED_SLOT:
                        wts = em.derived[sl & (DERIVED_SLOT - 1)]["w"]
                    elif sl in basis:
                        wts = [1.0 if b == sl and basis.index(sl) == k else 0.0 for k, b in enumerate(basis)]
                    else:
                        wts = em.slot_weights[sl]
                blist.append((sl, r, wts, gram))
        return blist_at[key]

    filts = []
    for gi, g in enumerate(em.selfg):
        g["unit0"] = len(units)
        for i in range(g["a"][1]):
            sa, _, ra = em.refs[g["a"][0] + i]
            sb, _, rb = g["gb"]
            filt = float(F(F(ra) + F(rb)) + F(g["slack"])) if g["slack"] >= 0 else -1.0
            R_ = F(F(ra) + F(rb)) + F(g["slack"])
            filts.append((sa, sb, float(F(R_ * R_)) if g["slack"] >= 0 else -1.0))  # the threshold self_filters compares with
            b_slots = [em.refs[g["b"][0] + j][0] for j in range(g["b"][1])]
            b_der = [sl & (DERIVED_SLOT - 1) for sl in b_slots if sl & DERIVED_SLOT]
            if b_der:
                # every b_j sits on the link whose basis the derived ones share
                bases = {tuple(em.derived[d]["basis"]) for d in b_der}
                assert len(bases) == 1, g["name"]
                assert all((sl & DERIVED_SLOT) or sl in em.derived[b_der[0]]["basis"] or sl in em.slot_weights for sl in b_slots), g["name"]
            assert g["b"][1] < 64 and len(em.derived) < 511, g["name"]
            pack = g["b"][1] | (((b_der[0] + 1) if b_der else 0) << 6) | ((1 << 15) if sa & DERIVED_SLOT else 0)
            units.append(((sa & (DERIVED_SLOT - 1)) if sa & DERIVED_SLOT else slot_offset(sa, stride, const_row, lift_row),
                          0 if sb & DERIVED_SLOT else slot_offset(sb, stride, const_row, lift_row), ra, filt, b_list(g), pack))
    assert len(refs) < 2048 and len(units) < 2048, (name, len(refs), len(units))
    ne, ns, na, nr = len(em.env), len(em.selfg), len(em.att), len(refs)
    t = f"Tables<{max(ne, 1)}, {max(ns, 1)}, {max(na, 1)}, {max(nr, 1)}>"

    def env_rows():
        rows = []
        for g in em.env or [{"guard": (0, 0, 0.0), "begin": 0, "count": 0, "name": "-", "slack": -1.0, "ubegin": 0, "ucount": 0}]:
            rows.append(f"    {{{_ref(g['guard'])}, {g['begin']}, {g['count']}, {_slack(g['slack'])}, {g['ubegin']}, {g['ucount']}}},  // {g['name']}")
        return rows

    def self_rows():
        rows = []
        for g in em.selfg or [{"ga": (0, 0, 0.0), "gb": (0, 0, 0.0), "a": (0, 0), "b": (0, 0), "name": "-", "slack": -1.0, "unit0": 0, "dense": False}]:
            rows.append(
                f"    {{{_ref(g['ga'])}, {_ref(g['gb'])}, {g['a'][0]}, {g['a'][1]}, {g['b'][0]}, {g['b'][1]}, {_slack(g['slack'])}, "
                f"{g['unit0']}, {1 if g.get('dense') else 0}}},  // {g['name']}"
            )
        return rows

    def ref_rows(rr):
        return ["    " + ", ".join(_ref(r) for r in rr[i : i + 6]) + "," for i in range(0, len(rr), 6)]

    rr = refs or [(0, 0, 0.0)]
    # centres the chain does not store (plan_links): pre-multiplied basis offsets + weights
    dd = em.derived or [{"basis": [0, 0, 0, 0], "w": [0.0, 0.0, 0.0, 0.0]}]
    assert all(not (b & DERIVED_SLOT) for d in dd for b in d["basis"])
    assert all(slot_offset(b, stride or 1, const_row, lift_row) < 65536 for d in dd for b in d["basis"]), name
    out = [f"__device__ const DerivedRec {name}_ders[{len(dd)}] = {{"]
    for d in dd:
        offs = ", ".join(str(slot_offset(b, stride, const_row, lift_row)) for b in d["basis"])
        out.append(f"    {{{{{offs}}}, {{{', '.join(_f32(v) for v in d['w'])}}}}},")
    out.append("};")
    out.append(f"__constant__ {t} {name} = {{")
    out.append(f"    {em.n_rows}, {ne}, {ns}, {na}, {em.att_env}, {3 * em.lift_ref[0] + 2 if em.lift_xyz else -1},")
    out.append("    {")
    out.extend("    " + r for r in env_rows())
    out.append("    },")
    out.append("    {")
    out.extend("    " + r for r in self_rows())
    out.append("    },")
    out.append("    {")
    for g in em.att or [{"guard": (0, 0, 0.0), "begin": 0, "count": 0, "name": "-"}]:
        out.append(f"        {{{_ref(g['guard'])}, {g['begin']}, {g['count']}, -1.0f, 0, 0}},  // {g['name']}")
    out.append("    },")
    out.append("    {")
    out.extend("    " + r for r in ref_rows(rr))
    out.append("    },")
    out.append("};")
    geom = {}
    if walk:
        # global-memory mirrors for lane-dependent indices (constant memory serialises those)
        out.append(f"__device__ const SlotRef {name}_grefs[{len(rr)}] = {{")
        out.extend(ref_rows(rr))
        out.append("};")
        rg = ref_group or [0xFFFF]
        out.append(f"__device__ const uint16_t {name}_ref_group[{len(rg)}] = {{")
        for i in range(0, len(rg), 16):
            out.append("    " + ", ".join(str(v) for v in rg[i : i + 16]) + ",")
        out.append("};")
        out.append(f"__device__ const RefGroup {name}_genv[{max(ne, 1)}] = {{")
        out.extend(env_rows())
        out.append("};")
        out.append(f"__device__ const SelfGroup {name}_gself[{max(ns, 1)}] = {{")
        out.extend(self_rows())
        out.append("};")
        uu = units or [(0, 0, 0.0, -1.0, 0, 0)]
        assert all(u[0] < 65536 and u[1] < 65536 for u in uu), name
        out.append(f"__device__ const SelfUnit {name}_units[{len(uu)}] = {{")
        for i in range(0, len(uu), 4):
            out.append("    " + ", ".join(f"{{{a}, {g}, {_f32(ra)}, {_f32(fr) if fr >= 0 else '-1.0f'}, {bb}, {bc}}}"
                                          for a, g, ra, fr, bb, bc in uu[i : i + 4]) + ",")
        out.append("};")
        # queued a-side filters (vb_walk.cuh: filter_units): slot codes + squared threshold per unit, units per group
        ff = filts or [(0, 0, -1.0)]
        out.append(f"__device__ const FilterRec {name}_filts[{len(ff)}] = {{")
        for i in range(0, len(ff), 6):
            out.append("    " + ", ".join(f"{{{a}, {g}, {_f32(t2) if t2 >= 0 else '-1.0f'}}}" for a, g, t2 in ff[i : i + 6]) + ",")
        out.append("};")
        gu = [g["unit0"] | (g["a"][1] << 16) for g in em.selfg] or [0]
        out.append(f"__device__ const uint32_t {name}_gunits[{len(gu)}] = {{")
        for i in range(0, len(gu), 12):
            out.append("    " + ", ".join(f"{v}u" for v in gu[i : i + 12]) + ",")
        out.append("};")
        bl = blist or [(0, 0.0, None, None)]
        out.append(f"__device__ const BRef {name}_brefs[{len(bl)}] = {{")
        for i in range(0, len(bl), 8):
            out.append("    " + ", ".join(f"{{{0 if sl & DERIVED_SLOT else slot_offset(sl, stride, const_row, lift_row)}, {_f32(r) if r >= 0 else '-1.0f'}}}" for sl, r, _, _ in bl[i : i + 8]) + ",")
        out.append("};")
        # b-side spheres of reduced links as weights over the link's basis (stored basis spheres: one-hot)
        def edge_form(wts, gram):
            if wts is None:
                return [0.0, 0.0, 0.0, 0.0]
            wk = wts[1:4]
            g = sum(wk[k] * wk[l] * gram[k][l] for k in range(3) for l in range(3))
            return [wk[0], wk[1], wk[2], g]

        bw_rows = [f"{{{{{', '.join(_f32(v) for v in edge_form(wts, gram))}}}, {_f32(r) if r >= 0 else '-1.0f'}, 0u}}" for _, r, wts, gram in bl]
        out.append(f"__device__ const BWeights {name}_bws[{len(bl)}] = {{")
        for i in range(0, len(bw_rows), 4):
            out.append("    " + ", ".join(bw_rows[i : i + 4]) + ",")
        out.append("};")
        max_u = max([g["ucount"] for g in em.env] or [1])
        max_a = max([g["a"][1] for g in em.selfg if not g.get("dense")] or [1])
        geom = {"env_q": 256, "self_q": 256, "pair_q": 256, "env_groups": max(ne, 1),
                # records of the warp walk that may live in shared memory (vb_walk.cuh: WT): (name, count, bytes each)
                "tab": [("Units", len(uu), 16), ("Brefs", len(bl), 8), ("Bws", len(bl), 24), ("Ders", len(dd), 24),
                        ("Env", max(ne, 1), 24), ("Refs", len(rr), 8), ("RefGroup", len(rg), 2)]
                       + ([("Filts", len(ff), 8), ("GUnits", len(gu), 4)] if queued else []),
                "max_a": max([g["a"][1] for g in em.selfg] or [1])}
    return out, t, geom


def _uniform(em: _Emitter, queued: bool = False) -> list[str]:
    """Straight-line guard phase of the warp walk (vb_walk.cuh): every self-collision guard pair, the
    a-side filters of the dense self groups and every environment guard lookup, with immediate
    shared-memory offsets and radii.  No lane-dependent branch, all loads independent: this is the
    part of the walk that every configuration executes."""
    import numpy as np

    F = np.float32
    ns, ne = len(em.selfg), len(em.env)
    sw, ew = max((ns + 31) // 32, 1), max((ne + 31) // 32, 1)
    leaf_mask = [0] * sw
    for gi, g in enumerate(em.selfg):
        if g["a"][1] == 0:
            leaf_mask[gi // 32] |= 1 << (gi % 32)
    out: list[str] = []
    w = out.append
    w(f"    static constexpr int kSelfWords = {sw}, kEnvWords = {ew};")

    def table(name, typ, vals):
        w(f"    VB_D static {typ} {name}(int i)")
        w("    {")
        w(f"        constexpr {typ} t[{len(vals)}] = {{{', '.join(str(v) + ('u' if typ == 'uint32_t' else '') for v in vals)}}};")
        w("        return t[i];")
        w("    }")

    # environment groups whose guard and fine spheres do not depend on the configuration (e.g. the base link):
    # their verdict is a property of (environment, base) and is evaluated once per warp, not per configuration
    const_slots = set(getattr(em, "const_slots", ()))
    const_mask = [0] * ew
    for gi, g in enumerate(em.env):
        slots = [g["guard"][0]] + [em.refs[g["begin"] + k][0] for k in range(g["count"])]
        if all(sl in const_slots for sl in slots) and g["guard"][2] >= 0:
            const_mask[gi // 32] |= 1 << (gi % 32)
    table("env_const_mask", "uint32_t", const_mask)
    w(f"    static constexpr bool kEnvConstAny = {'true' if any(const_mask) else 'false'};")
    table("self_leaf_mask", "uint32_t", leaf_mask)

    def sphere_loads(loaded: set[int], body: list[str]):
        def need(slot: int):
            if slot not in loaded:
                loaded.add(slot)
                if slot & DERIVED_SLOT:
                    d = em.derived[slot & (DERIVED_SLOT - 1)]
                    terms = [(wv, b) for wv, b in zip(d["w"], d["basis"]) if wv != 0.0]
                    for _, b in terms:
                        need(b)
                    for c in "xyz":
                        expr = " + ".join(f"{_f32(wv)} * s{b}{c}" for wv, b in terms)
                        body.append(f"const float s{slot}{c} = {expr};")
                    return
                if slot & CONST_SLOT:
                    x, y, z = em.const_xyz[slot]
                    body.append(f"const float s{slot}x = {x}, s{slot}y = {y}, s{slot}z = {z};")
                    return
                if slot & LIFT_SLOT:
                    x, y, dz = em.lift_xyz[slot]
                    body.append(f"const float s{slot}x = {x}, s{slot}y = {y}, s{slot}z = {dz} + col[{3 * em.lift_ref[0] + 2} * kStride];")
                    return
                body.append(
                    f"const float s{slot}x = col[{3 * slot} * kStride], s{slot}y = col[{3 * slot + 1} * kStride], s{slot}z = col[{3 * slot + 2} * kStride];"
                )

        def d2(a: int, b: int) -> str:
            need(a)
            need(b)
            return f"sqd(s{a}x - s{b}x, s{a}y - s{b}y, s{a}z - s{b}z)"

        d2.need = need
        return d2

    body: list[str] = []
    for wi in range(sw):
        body.append(f"uint32_t f{wi} = 0u;")
    # Robots with hundreds of guard pairs (Baxter: 355 pairs over 82 centres) would keep every centre they ever loaded
    # in registers to the end of the function: the pairs are emitted in scoped blocks, each loading what it needs.
    blk_pairs = GUARD_BLOCK if len(em.selfg) > 2 * GUARD_BLOCK > 0 else max(len(em.selfg), 1)
    for first in range(0, len(em.selfg), blk_pairs):
        scoped = blk_pairs < len(em.selfg)
        d2 = sphere_loads(set(), body)
        if scoped:
            body.append("{")
        for gi in range(first, min(first + blk_pairs, len(em.selfg))):
            g = em.selfg[gi]
            (sa, _, ra), (sb, _, rb) = g["ga"], g["gb"]
            rs = F(ra) + F(rb)
            rs2 = F(rs * rs)
            body.append(f"f{gi // 32} |= ({d2(sa, sb)} < {_f32(float(rs2))}) ? {1 << (gi % 32)}u : 0u;  // {g['name']}")
        if scoped:
            body.append("}")
    for wi in range(sw):
        body.append(f"fired[{wi}] = f{wi};")
    w("    // self-collision guard pairs (bit g of fired[])")
    w("    __device__ __forceinline__ static void self_guards(const float *__restrict__ col, uint32_t *__restrict__ fired)")
    w("    {")
    w("        auto sqd = [](float dx, float dy, float dz) { return dx * dx + dy * dy + dz * dz; };")
    for ln in body:
        w("        " + ln)
    w("    }")

    # a-side filters: unit (g, a_i) survives when a_i touches the other link's bounding sphere inflated by the
    # group's static slack.  A group is evaluated when its guard fired in any lane of the warp (warp-uniform
    # branch), with immediate offsets and radii; bit `unit` of ubits[] = unit to be run by the pair stage.
    n_units = sum(g["a"][1] for g in em.selfg)
    uw = max((n_units + 31) // 32, 1)
    w(f"    static constexpr int kUnitWords = {uw};  // {n_units} (group, a_i) work units")
    w("    // fired[]: this lane's guard hits (collided / tabulated-free groups already cleared)")
    w("    __device__ __forceinline__ static void self_filters(const float *__restrict__ col, const uint32_t *__restrict__ fired, uint32_t *__restrict__ ubits)")
    w("    {")
    w("        auto sqd = [](float dx, float dy, float dz) { return dx * dx + dy * dy + dz * dz; };")
    for wi in range(uw):
        w(f"        uint32_t u{wi} = 0u;")
    if not queued:
        # groups fired somewhere in the warp: one reduction per word, then warp-uniform branches (no vote per group)
        for wi in range(sw):
            w(f"        const uint32_t any{wi} = __reduce_or_sync(0xffffffffu, fired[{wi}]);")
    unit = 0
    for gi, g in enumerate(em.selfg):
        ac = g["a"][1]
        if ac == 0:
            continue
        assert g["unit0"] == unit, (g["name"], g["unit0"], unit)
        sb, rb = g["gb"][0], g["gb"][2]
        blk: list[str] = []
        d2g = sphere_loads(set(), blk)
        for i in range(ac):
            sa, _, ra = em.refs[g["a"][0] + i]
            pos = unit + i
            if g["slack"] < 0:
                blk.append(f"u{pos // 32} |= {1 << (pos % 32)}u;")
            else:
                R = F(F(ra) + F(rb)) + F(g["slack"])
                blk.append(f"u{pos // 32} |= ({d2g(sa, sb)} < {_f32(float(F(R * R)))}) ? {1 << (pos % 32)}u : 0u;")
        bit = f"(fired[{gi // 32}] & {1 << (gi % 32)}u)"
        if not queued:  # (queued filters: the function is never called, keep it out of the compile)
            w(f"        if ((any{gi // 32} & {1 << (gi % 32)}u) != 0u)  // {g['name']}")
            w("        {")
            w(f"            if ({bit} != 0u)")
            w("            {")
            for ln in blk:
                w("                " + ln)
            w("            }")
            w("        }")
        unit += ac
    for wi in range(uw):
        w(f"        ubits[{wi}] = u{wi};")
    w("    }")
    # environment guards: one branch-free grid fetch each, classified after the self-collision phase
    w("    // issues the grid-cell load of every environment guard sphere (bit g of inside[]: centre inside the grid).")
    w("    // Two levels: every guard first looks its cell up in the coarse grid (a few hundred KB: cache resident); only a")
    w("    // guard the coarse cell cannot call free fetches its cell of the fine grid.  A coarse-free guard gets a")
    w("    // far-away cell, which the classification below reads as free.")
    w("    __device__ __forceinline__ static void env_fetch(const float *__restrict__ col, float bx, float by, float bz,")
    w("                                                     const SdfView &grid, const SdfView &coarse, float radius,")
    w("                                                     uint2 *__restrict__ cells, uint32_t *__restrict__ inside)")
    w("    {")
    w("        bool in;")
    for wi in range(ew):
        w(f"        uint32_t i{wi} = 0u;")
    fetch_body: list[str] = []
    fetch_need = sphere_loads(set(), fetch_body).need
    guards = []
    for gi, g in enumerate(em.env):
        slot, mask, r = g["guard"]
        if slot & (DERIVED_SLOT | LIFT_SLOT):
            fetch_need(slot)
            for ln in fetch_body:
                w("        " + ln)
            fetch_body.clear()
        xyz = []
        for c in range(3):
            e = (f"s{slot}{'xyz'[c]}" if slot & (DERIVED_SLOT | LIFT_SLOT) else em.const_xyz[slot][c] if slot & CONST_SLOT
                 else f"col[{3 * slot + c} * kStride]")
            if mask & (1 << c):
                e += f" + b{'xyz'[c]}"
            xyz.append(e)
        guards.append((gi, g, xyz, "radius" if r < 0 else _f32(r)))
    w("        if (kTwoLevelGrid && coarse.cell != nullptr)")
    w("        {")
    for wi in range(ew):
        w(f"            uint32_t ci{wi} = 0u;")
    for gi, g, xyz, rr in guards:
        w(f"            cells[{gi}] = grid_fetch(coarse, {xyz[0]}, {xyz[1]}, {xyz[2]}, in); ci{gi // 32} |= in ? {1 << (gi % 32)}u : 0u;  // {g['name']}")
    for gi, g, xyz, rr in guards:
        bit = f"{1 << (gi % 32)}u"
        w(f"            if (grid_coarse_free(coarse, cells[{gi}], (ci{gi // 32} & {bit}) != 0u, {rr}))")
        w("            {")
        w(f"                cells[{gi}] = make_uint2(0x7f000000u, 0u); i{gi // 32} |= {bit};")
        w("            }")
        w("            else")
        w("            {")
        w(f"                cells[{gi}] = grid_fetch_inside(grid, {xyz[0]}, {xyz[1]}, {xyz[2]}, (ci{gi // 32} & {bit}) != 0u); i{gi // 32} |= ci{gi // 32} & {bit};")
        w("            }")
    w("        }")
    w("        else")
    w("        {")
    for gi, g, xyz, rr in guards:
        w(f"            cells[{gi}] = grid_fetch(grid, {xyz[0]}, {xyz[1]}, {xyz[2]}, in); i{gi // 32} |= in ? {1 << (gi % 32)}u : 0u;  // {g['name']}")
    w("        }")
    for wi in range(ew):
        w(f"        inside[{wi}] = i{wi};")
    w("    }")
    w("    // bit g of maybe[] (guard not certainly free) and cert[] (guard certainly hit); gm[32 * g] = candidate obstacles of group g")
    w("    __device__ __forceinline__ static void env_classify(const uint2 *__restrict__ cells, const uint32_t *__restrict__ inside, float radius,")
    w("                                                        const SdfView &grid, uint32_t *__restrict__ maybe, uint32_t *__restrict__ cert, uint32_t *gm)")
    w("    {")
    w("        bool mb, ct;")
    w("        uint32_t mk;")
    for wi in range(ew):
        w(f"        uint32_t m{wi} = 0u, c{wi} = 0u;")
    for gi, g in enumerate(em.env):
        slot, mask, r = g["guard"]
        rr = "radius" if r < 0 else _f32(r)
        if g["count"] == 0:
            reach = rr
        elif g["slack"] < 0 or r < 0:
            reach = "-1.0f"
        else:
            reach = _f32(r + g["slack"])
        bit = f"{1 << (gi % 32)}u"
        w(f"        grid_class(grid, cells[{gi}], (inside[{gi // 32}] & {bit}) != 0u, {rr}, {reach}, mb, ct, mk); "
          f"m{gi // 32} |= mb ? {bit} : 0u; c{gi // 32} |= ct ? {bit} : 0u; gm[{32 * gi}] = mk;  // {g['name']}")
    for wi in range(ew):
        w(f"        maybe[{wi}] = m{wi}; cert[{wi}] = c{wi};")
    w("    }")
    return out


def _queue_bytes(geom: dict) -> int:
    # item queue (self items, then environment items: never live together) + pair/fine-unit queue + filter-test queue
    # + candidate masks + state
    return 2 * (max(geom["env_q"], geom["self_q"]) + geom["pair_q"] + geom.get("filt_q", 0)) + 128 * geom["env_groups"] + 16


def _geometry(n_slots: int, geom: dict, dim: int = 0) -> tuple[int, int]:
    """(warps per CTA, CTAs per SM) maximising resident warps within 227 KB of shared memory per SM."""
    queue = _queue_bytes(geom)
    per_warp = 3 * n_slots * 4 * 32 + queue + 4 * dim * 32  # columns + queues + staged configurations (vb_kernels.cuh)
    fixed = 3 * n_slots * 4 + 16 + 16  # stride padding, staging mbarrier
    best = (1, 1)
    forced = int(os.environ.get("VB_FORCE_CTAS", "0"))  # development: A/B of the CTA geometry
    # The register file bounds the resident warps as well: the walk of an articulated robot takes ~125 registers per
    # thread, i.e. 65536 / (32 x 125) = 16 warps per SM, and more CTAs than one per SM do not add warps then (they only
    # add instruction streams).  The one-sphere robot (a handful of rows, ~40 registers) may fill all 32 warp slots.
    max_warps = 16 if n_slots > 8 else 32
    for ctas in ((forced,) if forced else (1, 2, 3, 4, 6, 8)):
        # 228 KB per SM, 1 KB of it reserved per resident CTA
        w = min((228 * 1024 // ctas - 1024 - fixed) // per_warp, 16, max_warps // ctas)
        if w >= 1 and w * ctas > best[0] * best[1]:
            best = (int(w), ctas)
    return best


def emit_robot(ir: dict[str, Any]) -> str:
    robot = ir["robot"]
    cls = robot.capitalize() + "Trace"
    dim, S = ir["dimension"], ir["n_spheres"]
    fns = ir["functions"]
    has_attach = "interleaved_sphere_fk_attachment" in fns
    st = trace_stats(ir)
    def sampled(fn):
        import numpy as np

        rng = np.random.default_rng(0)
        n = 1024
        lo, rg = np.array(ir["scale"]["lo"]), np.array(ir["scale"]["range"])
        q = (lo + rng.random((n, dim)) * rg).astype(np.float32)
        return ir_eval.evaluate(fn, q), n

    reduce_links = robot not in LINK_REDUCTION_OFF or os.environ.get("VB_REDUCE_ALL") == "1"
    if not reduce_links and os.environ.get("VB_REDUCE_BODIES") == "1":
        reduce_links = "bodies"
    em = _Emitter(fns["interleaved_sphere_fk"], sampled(fns["interleaved_sphere_fk"]), reduce_links, LIFT_ON).run_checks()
    em_att = (_Emitter(fns["interleaved_sphere_fk_attachment"], sampled(fns["interleaved_sphere_fk_attachment"]), reduce_links, LIFT_ON).run_checks()
              if has_attach else None)
    # self groups decided by a two-joint table are rare work for the walk, whatever their guard's fire rate
    ptabs = pair_tables.build(ir, fns["interleaved_sphere_fk"])
    for t in ptabs:
        for g in t["groups"]:
            em.selfg[g]["dense"] = False
    out: list[str] = []
    w = out.append
    w(f"// GENERATED by mr_vamp_b200/codegen/emit_cuda.py from {ir['source']} of the VAMP checkout")
    w("// this library was built against.  Build product: do not edit, do not commit.")
    w("#pragma once")
    w('#include "../csrc/vb_sinks.cuh"')
    w('#include "../csrc/vb_tables.cuh"')
    w('#include "../csrc/vb_sdf.cuh"')
    w("namespace vb")
    w("{")
    queued = robot in QUEUED_FILTERS
    max_a = max([g["a"][1] for g in em.selfg] or [1])
    geom = {"env_q": 256, "self_q": 256, "pair_q": 256, "env_groups": max(len(em.env), 1),
            "filt_q": (max(int(os.environ.get("VB_FORCE_FILTQ", "256")), (max_a + 31) // 32 * 32) if queued else 0)}
    # shared-memory rows: one per configuration-dependent centre (a column per thread), then the constant rows
    # (one copy per warp, 32 constants per row)
    const_row = max(em.n_rows, em_att.n_rows if em_att else 0)
    n_const = max(len(em.const_slots), len(em_att.const_slots) if em_att else 0)
    lift_row = const_row + (n_const + 31) // 32  # translated constants (LIFT_SLOT) behind the constants
    n_lift = max(len(em.lift_xyz), len(em_att.lift_xyz) if em_att else 0)
    n_slots = lift_row + (n_lift + 31) // 32
    # Window plan: resident warps, queue capacities, and which of {staged configurations, check tables} go behind
    # the columns.  Scored with the effects measured on B200 (Panda, MBM scene): throughput ~ warps^0.4 (8 -> 11
    # warps: +14 %), staged configurations +3 %, tables in shared memory +5 %.  Overflowing queue units run in
    # place (vb_walk.cuh): correct, but slow when it happens often, so the capacities stay >= 192.
    _, _, tgeom0 = _tables(f"k_{robot}_tables", em, walk=True, stride=33, const_row=const_row, queued=queued, lift_row=lift_row)
    # tables in the order in which they earn their place in shared memory (the unit stages' dependent reads first)
    TAB_PRIORITY = ["Units", "Filts", "GUnits", "Brefs", "Bws", "Ders", "Refs", "RefGroup", "Env"]
    tab_sizes = {tname: (cnt * size + 15) & ~15 for tname, cnt, size in tgeom0["tab"]}
    tab_bytes = sum(tab_sizes.values())

    def tables_that_fit(spare: int) -> list[str]:
        got = []
        for tname in TAB_PRIORITY:
            if tname in tab_sizes and tab_sizes[tname] <= spare:
                got.append(tname)
                spare -= tab_sizes[tname]
        return got

    def window_bytes(g, w_, stage):
        st = 32 * w_ + 1
        b = 4 * ((3 * n_slots * st + 3) & ~3) + w_ * _queue_bytes(g) + ((32 * w_ * dim * 4 + 16) if stage else 0)
        return (b + 15) & ~15

    best = None
    force_warps = int(os.environ.get("VB_FORCE_WARPS", "0"))  # development: A/B of warps against tables in shared memory
    for cap in (256, 224, 192):  # (128 was measured on Fetch: queue overflow, 0.73 -> 0.46 Gcfg/s)
        g = dict(geom, pair_q=cap, env_q=cap, self_q=cap)
        for stage in (False, True):
            w_, c_ = _geometry(n_slots, g, dim if stage else 0)
            for drop in ((w_ - force_warps,) if force_warps else (0, 1)):  # one warp less can buy both extras
                ww = w_ - drop
                if ww < 1:
                    continue
                avail = 228 * 1024 // c_ - 1024
                if window_bytes(g, ww, stage) > avail:
                    continue
                got = tables_that_fit(avail - window_bytes(g, ww, stage))
                share = sum(tab_sizes[t_] for t_ in got) / max(tab_bytes, 1)
                score = (ww * c_) ** 0.4 * (1.03 if stage else 1.0) * (1.0 + 0.05 * share)
                if best is None or score > best[0] + 1e-9:
                    best = (score, g, ww, c_, stage, got)
    _, geom, warps, ctas, stage_cfg, shared_tabs = best
    if os.environ.get("VB_FORCE_STAGE") == "1":  # development: staged configurations whatever the plan says
        stage_cfg, shared_tabs = True, []
    if os.environ.get("VB_FORCE_Q"):  # development: "env,self,pair" capacities, to exercise the overflow paths
        e_, s_, p_ = (int(v) for v in os.environ["VB_FORCE_Q"].split(","))
        geom = dict(geom, env_q=e_, self_q=s_, pair_q=p_)
        shared_tabs = []
    tab, ttype, tgeom = _tables(f"k_{robot}_tables", em, walk=True, stride=32 * warps + 1, const_row=const_row, queued=queued, lift_row=lift_row)
    out.extend(tab)
    tab_off, tab_lines = 0, []
    for tname, cnt, size in tgeom["tab"]:
        if tname in shared_tabs:
            tab_lines.append(f"    static constexpr unsigned kTab{tname} = {tab_off}, kTab{tname}Bytes = {cnt * size};")
            tab_off = (tab_off + cnt * size + 15) & ~15
        else:
            tab_lines.append(f"    static constexpr unsigned kTab{tname} = 0xffffffffu, kTab{tname}Bytes = {cnt * size};  // read through L1")
    if not queued:
        tab_lines.append("    static constexpr unsigned kTabFilts = 0xffffffffu, kTabFiltsBytes = 0, kTabGUnits = 0xffffffffu, kTabGUnitsBytes = 0;")
    tables_in_shared = tab_off > 0
    if em_att:
        tab2, ttype2, _ = _tables(f"k_{robot}_tables_attach", em_att, stride=32 * warps + 1, const_row=const_row, lift_row=lift_row)
        out.extend(tab2)
    pt_head, pt_body = pair_tables.emit(robot, ir, ptabs, max((len(em.selfg) + 31) // 32, 1))
    out.extend(pt_head)
    w(f"struct {cls}")
    w("{")
    w(f'    static constexpr const char *kName = "{robot}";')
    w(f"    static constexpr int kDim = {dim};")
    w(f"    static constexpr int kSpheres = {S};")
    w(f"    static constexpr int kResolution = {ir['resolution']};")
    w(f"    static constexpr bool kHasAttach = {'true' if has_attach else 'false'};")
    w(f"    static constexpr int kSlots = {n_slots};  // shared-memory slot rows: {const_row} per-thread + {n_slots - const_row} of per-warp constants")
    w(f"    static constexpr int kConstRow = {const_row};  // first constant row ({n_const} configuration-independent centres)")
    w(f"    static constexpr int kLiftRow = {lift_row};  // first row of the translated constants ({n_lift} centres: x, y constant, z = dz + z of one stored centre)")
    w(f"    static constexpr int kLiftCount = {n_lift};")
    w(f"    static constexpr int kLiftZRow = {3 * em.lift_ref[0] + 2 if em.lift_xyz else 0};  // row of that centre's z (non-attachment trace)")
    w(f"    static constexpr int kBlock = {32 * warps};  // threads per CTA: {warps} warps x {ctas} CTA(s) per SM fill the shared memory")
    w(f"    static constexpr bool kTablesInShared = {'true' if tables_in_shared else 'false'};  // check tables of the warp walk in shared memory ({tab_off} of {tab_bytes} bytes: {', '.join(shared_tabs) or 'none'})")
    w(f"    static constexpr unsigned kTabBytes = {tab_off};")
    for ln in tab_lines:
        w(ln)
    w(f"    static constexpr bool kTwoLevelGrid = {'false' if robot in TWO_LEVEL_GRID_OFF else 'true'};  // coarse grid level in front of the fine one")
    w(f"    static constexpr int kCtasPerSm = {ctas};  // resident CTAs per SM (second argument of __launch_bounds__)")
    w("    static constexpr int kStride = kBlock + 1;  // slot-column stride in floats (odd: conflict free)")
    w(f"    static constexpr bool kStageCfg = {'true' if stage_cfg else 'false'};  // configurations staged in shared memory by bulk copy")
    w(f"    static constexpr int kEnvQ = {geom['env_q']}, kSelfQ = {geom['self_q']}, kPairQ = {geom['pair_q']};  // work-queue capacities (units)")
    w(f"    static constexpr bool kQueuedFilters = {'true' if queued else 'false'};  // a-side filters as queued work (vb_walk.cuh: filter_units)")
    w(f"    static constexpr int kFiltQ = {geom['filt_q']};  // filter-test queue (>= the largest a-side of a self group: {max_a})")
    w(f"    static constexpr int kChainFlops = {st['chain_flops']};")
    w(f"    static constexpr int kEnvGroups = {st['env_groups']}, kEnvFine = {st['env_fine']};")
    w(f"    static constexpr int kSelfGroups = {st['self_groups']}, kSelfFine = {st['self_fine']};")
    w(f"    using TablesT = {ttype};")
    w(f"    __device__ __forceinline__ static const TablesT &tables() {{ return k_{robot}_tables; }}")
    w("    __device__ __forceinline__ static WalkTables walk_tables()")
    w("    {")
    w(f"        return WalkTables{{k_{robot}_tables_grefs, k_{robot}_tables_ref_group, k_{robot}_tables_genv, k_{robot}_tables_gself, k_{robot}_tables_units, k_{robot}_tables_brefs, k_{robot}_tables_ders, k_{robot}_tables_bws, k_{robot}_tables_filts, k_{robot}_tables_gunits}};")
    w("    }")
    w(f"    __device__ __forceinline__ static const DerivedRec *ders() {{ return k_{robot}_tables_ders; }}")
    w(f"    __device__ __forceinline__ static const DerivedRec *ders_attach() {{ return k_{robot}_tables{'_attach' if em_att else ''}_ders; }}")
    w(f"    static constexpr int kDerived = {len(em.derived)};  // centres rebuilt from their link's basis instead of stored")
    if em_att:
        w(f"    using TablesAttachT = {ttype2};")
        w(f"    __device__ __forceinline__ static const TablesAttachT &tables_attach() {{ return k_{robot}_tables_attach; }}")
    else:
        w("    using TablesAttachT = TablesT;")
        w(f"    __device__ __forceinline__ static const TablesAttachT &tables_attach() {{ return k_{robot}_tables; }}")
    lo = ", ".join(_f32(v) for v in ir["scale"]["lo"])
    rg = ", ".join(_f32(v) for v in ir["scale"]["range"])
    w("    VB_HD static float lo(int j)")
    w("    {")
    w(f"        constexpr float t[{dim}] = {{{lo}}};")
    w("        return t[j];")
    w("    }")
    w("    VB_HD static float range(int j)")
    w("    {")
    w(f"        constexpr float t[{dim}] = {{{rg}}};")
    w("        return t[j];")
    w("    }")
    inv = ir["scale"].get("inv")
    w("    // descale_configuration: (q - lo) * inv, or (q - lo) / range where the reference divides")
    w(f"    static constexpr bool kDescaleDivides = {'false' if inv else 'true'};")
    w("    VB_HD static float inv_range(int j)")
    w("    {")
    w(f"        constexpr float t[{dim}] = {{{', '.join(_f32(v) for v in (inv or ir['scale']['range']))}}};")
    w("        return t[j];")
    w("    }")

    def fn(name: str, body: list[str], sink: str = "SlotSink"):
        w(f"    __device__ VB_CHAIN_ATTR static void {name}(const {sink} &p, const float *__restrict__ q)")
        w("    {")
        for ln in body:
            w("        " + ln)
        w("    }")

    out.extend(_uniform(em, queued))
    out.extend(pt_body)
    fn("chain", em.lines)
    fn("chain_attach", em_att.lines if em_att else ["chain(p, q);"])
    fn("sphere_fk", _Emitter(fns["sphere_fk"]).run_spheres(), "SphereSink")
    w("};")
    w("}  // namespace vb")
    return "\n".join(out) + "\n"


def generate(robots=None, out_dir: str = GEN_DIR) -> list[str]:
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for r in robots or ROBOTS:
        ir = trace_parser.parse_robot(r)
        text = emit_robot(ir)
        path = os.path.join(out_dir, f"{r}_trace.cuh")
        old = None
        if os.path.exists(path):
            with open(path) as fh:
                old = fh.read()
        if old != text:
            with open(path, "w") as fh:
                fh.write(text)
        paths.append(path)
    return paths


if __name__ == "__main__":
    for p in generate(sys.argv[1:] or None, os.environ.get("VB_GEN_OUT", GEN_DIR)):
        print(p, os.path.getsize(p))
