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

from . import ir_eval, trace_parser

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




class _Emitter:
    """Lowers one traced function to (chain lines, check tables)."""

    def __init__(self, fn: dict[str, Any], samples=None):
        self.values = fn["values"]
        self.program = fn["program"]
        self.samples = samples  # (vals, n) from ir_eval, for the static containment slack
        self.lines: list[str] = []
        self.slots: dict[tuple, int] = {}
        self.refs: list[tuple[int, int, float]] = []  # (slot, base mask, radius)
        self.env: list[dict[str, Any]] = []
        self.selfg: list[dict[str, Any]] = []
        self.att: list[dict[str, Any]] = []
        self.att_env = 0
        self.has_pose = False

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
        key = tuple((a["v"] if a["v"] is not None else ("c", a["c"])) for a in s[:3])
        if key not in self.slots:
            self.slots[key] = len(self.slots)
            self.lines.append(
                f"p.put_s<kStride>({self.slots[key]}, {self.coord(s[0])}, {self.coord(s[1])}, {self.coord(s[2])});"
            )
        mask = sum((1 << i) for i, a in enumerate(s[:3]) if a.get("base"))
        for i, a in enumerate(s[:3]):
            assert a.get("base") in (None, "xyz"[i]), s
        r = s[3]
        radius = -1.0 if r.get("sym") == "radius" else float(r["c"])
        assert r.get("sym") == "radius" or r["v"] is None
        return (self.slots[key], mask, radius)

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
            worst = max(worst, float(d.max()) + s[3]["c"] - guard[3]["c"])
        return max(worst, 0.0) + 1e-4

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
            self.selfg.append({"ga": ga, "gb": gb, "a": (ab, ac), "b": (bb, bc), "name": name, "slack": slack_b})
            return
        raise ValueError(k)

    def run_checks(self):
        for it in self.program:
            if it["k"] == "val":
                self.value(it["id"])
            elif it["k"] == "out":
                raise ValueError("out in check trace")
            else:
                self.group(it)
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


def _tables(name: str, em: _Emitter) -> list[str]:
    ne, ns, na, nr = len(em.env), len(em.selfg), len(em.att), len(em.refs)
    t = f"Tables<{max(ne, 1)}, {max(ns, 1)}, {max(na, 1)}, {max(nr, 1)}>"
    out = [f"__constant__ {t} {name} = {{"]
    out.append(f"    {len(em.slots)}, {ne}, {ns}, {na}, {em.att_env},")
    out.append("    {")
    for g in em.env or [{"guard": (0, 0, 0.0), "begin": 0, "count": 0, "name": "-", "slack": -1.0}]:
        out.append(f"        {{{_ref(g['guard'])}, {g['begin']}, {g['count']}, {_slack(g['slack'])}}},  // {g['name']}")
    out.append("    },")
    out.append("    {")
    for g in em.selfg or [{"ga": (0, 0, 0.0), "gb": (0, 0, 0.0), "a": (0, 0), "b": (0, 0), "name": "-", "slack": -1.0}]:
        out.append(
            f"        {{{_ref(g['ga'])}, {_ref(g['gb'])}, {g['a'][0]}, {g['a'][1]}, {g['b'][0]}, {g['b'][1]}, {_slack(g['slack'])}}},"
            f"  // {g['name']}"
        )
    out.append("    },")
    out.append("    {")
    for g in em.att or [{"guard": (0, 0, 0.0), "begin": 0, "count": 0, "name": "-"}]:
        out.append(f"        {{{_ref(g['guard'])}, {g['begin']}, {g['count']}, -1.0f}},  // {g['name']}")
    out.append("    },")
    out.append("    {")
    refs = em.refs or [(0, 0, 0.0)]
    for i in range(0, len(refs), 6):
        out.append("        " + ", ".join(_ref(r) for r in refs[i : i + 6]) + ",")
    out.append("    },")
    out.append("};")
    # mirror of the refs in global memory for lane-dependent indices (constant memory serialises those)
    out.append(f"__device__ const SlotRef {name}_grefs[{max(nr, 1)}] = {{")
    for i in range(0, len(refs), 6):
        out.append("    " + ", ".join(_ref(r) for r in refs[i : i + 6]) + ",")
    out.append("};")
    return out, t


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

    em = _Emitter(fns["interleaved_sphere_fk"], sampled(fns["interleaved_sphere_fk"])).run_checks()
    em_att = _Emitter(fns["interleaved_sphere_fk_attachment"]).run_checks() if has_attach else None
    out: list[str] = []
    w = out.append
    w(f"// GENERATED by mr_vamp_b200/codegen/emit_cuda.py from {ir['source']} of the VAMP checkout")
    w("// this library was built against.  Build product: do not edit, do not commit.")
    w("#pragma once")
    w('#include "../csrc/vb_sinks.cuh"')
    w('#include "../csrc/vb_tables.cuh"')
    w("namespace vb")
    w("{")
    tab, ttype = _tables(f"k_{robot}_tables", em)
    out.extend(tab)
    if em_att:
        tab2, ttype2 = _tables(f"k_{robot}_tables_attach", em_att)
        out.extend(tab2)
    w(f"struct {cls}")
    w("{")
    w(f'    static constexpr const char *kName = "{robot}";')
    w(f"    static constexpr int kDim = {dim};")
    w(f"    static constexpr int kSpheres = {S};")
    w(f"    static constexpr int kResolution = {ir['resolution']};")
    w(f"    static constexpr bool kHasAttach = {'true' if has_attach else 'false'};")
    n_slots = max(len(em.slots), len(em_att.slots) if em_att else 0)
    w(f"    static constexpr int kSlots = {n_slots};  // distinct sphere centres used by the checks")
    w(f"    static constexpr int kBlock = {128 if n_slots <= 72 else 64};  // threads per CTA (shared-memory budget)")
    w("    static constexpr int kStride = kBlock + 1;  // slot-column stride in floats (odd: conflict free)")
    w(f"    static constexpr int kChainFlops = {st['chain_flops']};")
    w(f"    static constexpr int kEnvGroups = {st['env_groups']}, kEnvFine = {st['env_fine']};")
    w(f"    static constexpr int kSelfGroups = {st['self_groups']}, kSelfFine = {st['self_fine']};")
    w(f"    using TablesT = {ttype};")
    w(f"    __device__ __forceinline__ static const TablesT &tables() {{ return k_{robot}_tables; }}")
    w(f"    __device__ __forceinline__ static const SlotRef *grefs() {{ return k_{robot}_tables_grefs; }}")
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
        w(f"    __device__ __noinline__ static void {name}(const {sink} &p, const float *__restrict__ q)")
        w("    {")
        for ln in body:
            w("        " + ln)
        w("    }")

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
    for p in generate(sys.argv[1:] or None):
        print(p, os.path.getsize(p))
