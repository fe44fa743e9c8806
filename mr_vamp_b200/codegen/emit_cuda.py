"""Back end of the kernel compiler: robot IR (trace_parser) -> CUDA device code for sm_100a.

For every robot one header ``mr_vamp_b200/_gen/<robot>_trace.cuh`` is written.  It holds a struct
``vb::<Robot>Trace`` with three policy-templated device functions

* ``fkcc(P&, const float* q)``          <- ``interleaved_sphere_fk``           (robots/<r>/fk.hh)
* ``fkcc_attach(P&, const float* q)``   <- ``interleaved_sphere_fk_attachment``
* ``sphere_fk(P&, const float* q)``     <- ``sphere_fk``

The policy supplies the collision tests (``env_test``, ``self_test``, ``attach_*``), the base
offset and the output sink, so the same straight-line trace serves the verdict kernels, the
clearance ("margin") kernels and the work-counting kernels in ``csrc/vb_kernels.cuh``.  The emitted
arithmetic follows the IR value by value (literals narrowed to float exactly where the C++
expression narrows them), which is what keeps sphere centres within 1e-5 m of the reference; the
scheduling, register allocation and FMA contraction are left to nvcc/ptxas.

The generated headers are build products (git-ignored): the traces belong to the VAMP checkout the
library is built against (``$VAMP_SRC``), exactly like the reference's own ``fk.hh`` files belong
to its robot-description pipeline.
"""
from __future__ import annotations

import os
import struct
import sys
from typing import Any

from . import trace_parser

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
    def __init__(self, fn: dict[str, Any]):
        self.values = fn["values"]
        self.program = fn["program"]
        self.lines: list[str] = []
        self.pending_ops = 0

    # -- operands -----------------------------------------------------------------------------
    def opnd(self, o) -> str:
        if isinstance(o, dict):
            return _f32(o["c"])
        return f"v{o}"

    def arg(self, a: dict[str, Any]) -> str:
        """Argument of a check call / output: value | literal [+ base]."""
        base = a.get("base")
        if a.get("sym") == "radius":
            return "p.radius"
        if a["v"] is not None:
            s = f"v{a['v']}"
            return f"({s} + p.b{base}())" if base else s
        if base:
            # `0.122 + base_z`: double + float -> double, narrowed when the check upcasts it
            return f"p.cb{base}({_f64(a['c'])})"
        return _f32(a["c"])

    def sphere(self, s) -> str:
        return ", ".join(self.arg(a) for a in s)

    # -- statements ---------------------------------------------------------------------------
    def value(self, vid: int):
        op, a, b = self.values[vid]
        if op == "in":
            rhs = f"q[{a}]"
        elif op in ("mul", "add", "sub"):
            rhs = f"{self.opnd(a)} {'*' if op == 'mul' else '+' if op == 'add' else '-'} {self.opnd(b)}"
            self.pending_ops += 1
        elif op == "neg":
            rhs = f"-{self.opnd(a)}"
        elif op == "sin":
            rhs = f"vamp_sin({self.opnd(a)})"
            self.pending_ops += 11
        elif op == "cos":
            rhs = f"vamp_cos({self.opnd(a)})"
            self.pending_ops += 14
        else:
            raise ValueError(op)
        self.lines.append(f"const float v{vid} = {rhs};")

    def flush_ops(self):
        if self.pending_ops:
            self.lines.append(f"p.ops({self.pending_ops});")
            self.pending_ops = 0

    def group(self, it: dict[str, Any]):
        k = it["k"]
        self.flush_ops()
        if k == "attach_pose":
            self.lines.append("p.attach_pose(" + ", ".join(self.arg(a) for a in it["args"]) + ");")
            return
        if k == "attach_env":
            self.lines.append("if (p.attach_env()) { return false; }")
            return
        if k == "env":
            test = lambda s: f"p.env_test({self.sphere(s)})"
        elif k == "self":
            test = lambda s: f"p.self_test({self.sphere(s[0])}, {self.sphere(s[1])})"
        elif k == "attach_self":
            test = lambda s: f"p.attach_self({self.sphere(s)})"
        else:
            raise ValueError(k)
        name = it.get("name", "")
        fine = it["fine"]
        if it.get("leaf") or not fine or fine == [it["guard"]]:
            # a bare `if (check) return false;` : the guard is the test
            self.lines.append(f"if ({test(it['guard'])}) {{ return false; }}  // {name}")
            return
        self.lines.append(f"if ({test(it['guard'])})  // {name}")
        self.lines.append("{")
        for s in fine:
            self.lines.append(f"    if ({test(s)}) {{ return false; }}")
        self.lines.append("}")

    def run_checks(self) -> list[str]:
        for it in self.program:
            if it["k"] == "val":
                self.value(it["id"])
            elif it["k"] == "out":
                raise ValueError("out in check trace")
            else:
                self.group(it)
        self.flush_ops()
        self.lines.append("return true;")
        return self.lines

    def run_spheres(self) -> list[str]:
        for it in self.program:
            if it["k"] == "val":
                self.value(it["id"])
            elif it["k"] == "out":
                self.lines.append(f"p.out_{it['field']}({it['idx']}, {self.arg(it['arg'])});")
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
        "env_leaf": sum(1 for g in env_g if g.get("leaf")),
        "self_groups": len(self_g),
        "self_fine": sum(len(g["fine"]) for g in self_g),
    }


def emit_robot(ir: dict[str, Any]) -> str:
    robot = ir["robot"]
    cls = robot.capitalize() + "Trace"
    dim, S = ir["dimension"], ir["n_spheres"]
    fns = ir["functions"]
    has_attach = "interleaved_sphere_fk_attachment" in fns
    st = trace_stats(ir)
    out: list[str] = []
    w = out.append
    w(f"// GENERATED by mr_vamp_b200/codegen/emit_cuda.py from {ir['source']} of the VAMP checkout")
    w("// this library was built against.  Build product: do not edit, do not commit.")
    w("#pragma once")
    w('#include "../csrc/vb_numerics.cuh"')
    w("namespace vb")
    w("{")
    w(f"struct {cls}")
    w("{")
    w(f'    static constexpr const char *kName = "{robot}";')
    w(f"    static constexpr int kDim = {dim};")
    w(f"    static constexpr int kSpheres = {S};")
    w(f"    static constexpr int kResolution = {ir['resolution']};")
    w(f"    static constexpr bool kHasAttach = {'true' if has_attach else 'false'};")
    w(f"    static constexpr int kChainFlops = {st['chain_flops']};")
    w(f"    static constexpr int kEnvGroups = {st['env_groups']}, kEnvFine = {st['env_fine']};")
    w(f"    static constexpr int kSelfGroups = {st['self_groups']}, kSelfFine = {st['self_fine']};")
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

    def fn(name: str, ret: str, body: list[str]):
        w("    template <class P>")
        w(f"    __device__ __forceinline__ static {ret} {name}(P &p, const float *__restrict__ q)")
        w("    {")
        for ln in body:
            w("        " + ln)
        w("    }")

    fn("fkcc", "bool", _Emitter(fns["interleaved_sphere_fk"]).run_checks())
    if has_attach:
        fn("fkcc_attach", "bool", _Emitter(fns["interleaved_sphere_fk_attachment"]).run_checks())
    else:
        fn("fkcc_attach", "bool", ["return fkcc(p, q);"])
    fn("sphere_fk", "void", _Emitter(fns["sphere_fk"]).run_spheres())
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
