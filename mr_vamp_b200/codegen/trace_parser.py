"""Front end of the kernel compiler: VAMP traced-robot programs -> robot IR.

VAMP describes a robot to its collision checker as a machine-generated,
straight-line "trace" of forward kinematics interleaved with collision checks
(``robots/<robot>/fk.hh``: ``sphere_fk``, ``interleaved_sphere_fk`` and
``interleaved_sphere_fk_attachment``; statement grammar in SURVEY.md appendix
A.1).  Those files are the ecosystem's robot interchange format: a maintainer
adds a robot by generating a new trace.  This module reads such a trace from a
VAMP / mr-vamp checkout (``$VAMP_SRC``; nothing from it is stored in this
repository) and lowers it to a small SSA IR that the back ends in this package
re-schedule for sm_100a (``emit_cuda.py``) or interpret with numpy
(``oracle/port.py``).

IR (plain JSON-able dicts)
--------------------------
``values``: every SSA value in definition order, ``[op, a, b]`` with
  op in {"in","mul","add","sub","neg","sin","cos"}; operands are either an int
  (value id) or a ``{"c": <float>, "f": <bool>}`` literal (``f`` = the source
  literal carried an ``f`` suffix, i.e. was a float and not a double).
``program``: the statement stream in source order; items are
  ``{"k":"val","id":i}``                          define value i
  ``{"k":"env","name":..,"guard":S,"fine":[S..]}``  environment group
  ``{"k":"self","name":..,"guard":[S,S],"fine":[[S,S]..]}`` self-collision group
  ``{"k":"out","field":"x|y|z|r","idx":n,"arg":A}`` sphere_fk output
  ``{"k":"attach_pose","args":[A*7]}``, ``{"k":"attach_self","sphere":S}``,
  ``{"k":"attach_env"}``                           attachment statements
where a sphere S = [A,A,A,A] and an argument
A = {"v": id | None, "c": float | None, "f": bool, "base": "x"|"y"|"z"|None}.
"""
from __future__ import annotations

import json
import os
import re
from typing import Any

_NUM = r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?[fF]?"
_ID = r"[A-Za-z_][A-Za-z_0-9]*"
_RE_NUM = re.compile(r"^" + _NUM + r"$")
_RE_ID = re.compile(r"^" + _ID + r"$")


def vamp_src_root() -> str:
    root = os.environ.get("VAMP_SRC", "/root/reference")
    impl = os.path.join(root, "src", "impl", "vamp")
    if not os.path.isdir(impl):
        raise FileNotFoundError(
            f"VAMP checkout not found at {root!r} (set VAMP_SRC); the robot traces "
            "robots/<robot>/fk.hh are the input of the kernel compiler"
        )
    return impl


def _function_body(text: str, name: str) -> str:
    """Return the text between the braces of ``name(...) ... { ... }``."""
    m = re.search(r"\b" + re.escape(name) + r"\s*\(", text)
    if not m:
        raise KeyError(name)
    i = text.index("{", m.end())
    depth = 0
    for j in range(i, len(text)):
        ch = text[j]
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                return text[i + 1 : j]
    raise ValueError("unbalanced braces in " + name)


def _lit(tok: str) -> dict[str, Any]:
    f = tok[-1] in "fF"
    return {"c": float(tok.rstrip("fF")), "f": f}


class _Parser:
    def __init__(self, body: str):
        self.names: dict[str, int] = {}
        self.values: list[list[Any]] = []
        self.program: list[dict[str, Any]] = []
        self.body = body

    # -- operands ---------------------------------------------------------
    def operand(self, tok: str):
        tok = tok.strip()
        if _RE_NUM.match(tok):
            return _lit(tok)
        if _RE_ID.match(tok):
            return self.names[tok]
        raise ValueError("operand? " + tok)

    def arg(self, tok: str) -> dict[str, Any]:
        """Check-call argument: ``X``, ``X + base_a`` with X a value or literal."""
        tok = tok.strip()
        base = None
        m = re.match(r"^(.*?)\s*\+\s*base_([xyz])$", tok)
        if m:
            tok, base = m.group(1).strip(), m.group(2)
        else:
            m = re.match(r"^base_([xyz])\s*\+\s*(.*)$", tok)
            if m:
                base, tok = m.group(1), m.group(2).strip()
        if _RE_NUM.match(tok):
            lit = _lit(tok)
            return {"v": None, "c": lit["c"], "f": lit["f"], "base": base}
        if tok == "radius":  # robots::sphere global radius
            return {"v": None, "c": None, "f": True, "base": base, "sym": "radius"}
        m = re.match(r"^q\[(\d+)\]$", tok)
        if m:
            return {"v": self.define(["in", int(m.group(1)), None], None), "c": None, "f": True, "base": base}
        return {"v": self.names[tok], "c": None, "f": True, "base": base}

    def define(self, node, name):
        vid = len(self.values)
        self.values.append(node)
        if name is not None:
            if name in self.names:
                raise ValueError("redefinition of " + name)
            self.names[name] = vid
        self.program.append({"k": "val", "id": vid})
        return vid

    # -- expressions ------------------------------------------------------
    def expr(self, name: str, rhs: str):
        rhs = rhs.strip()
        m = re.match(r"^q\[(\d+)\]$", rhs)
        if m:
            return self.define(["in", int(m.group(1)), None], name)
        m = re.match(r"^(" + _ID + r")\.(sin|cos)\(\)$", rhs)
        if m:
            return self.define([m.group(2), self.names[m.group(1)], None], name)
        m = re.match(r"^-\s*(" + _ID + r")$", rhs)
        if m:
            return self.define(["neg", self.names[m.group(1)], None], name)
        m = re.match(r"^(" + _ID + "|" + _NUM + r")\s*([*+-])\s*(" + _ID + "|" + _NUM + r")$", rhs)
        if m:
            op = {"*": "mul", "+": "add", "-": "sub"}[m.group(2)]
            return self.define([op, self.operand(m.group(1)), self.operand(m.group(3))], name)
        raise ValueError("expression? " + rhs)

    # -- statements -------------------------------------------------------
    def run(self):
        # join lines, keep /*name*/ comments, drop // comments
        text = re.sub(r"//[^\n]*", "", self.body)
        text = re.sub(r"\s+", " ", text)
        pos = 0
        n = len(text)
        group = None  # currently open group dict
        depth = 0
        re_auto = re.compile(r"\s*auto\s+(" + _ID + r")\s*=\s*([^;]+);")
        re_float = re.compile(r"\s*float\s+base_[xyz]\s*=[^;]+;")
        re_out = re.compile(r"\s*out\.([xyzr])\[(\d+)\]\s*=\s*([^;]+);")
        re_if = re.compile(r"\s*if\s*\(\s*(?:/\*(.*?)\*/)?\s*(" + _ID + r")\s*(?:<[^>]*>)?\s*\(")
        re_call = re.compile(r"\s*(" + _ID + r")\s*(?:<[^>]*>)?\s*\(")
        while pos < n:
            if text[pos:].strip() == "":
                break
            m = re_float.match(text, pos)
            if m:
                pos = m.end()
                continue
            m = re_auto.match(text, pos)
            if m:
                self.expr(m.group(1), m.group(2))
                pos = m.end()
                continue
            m = re_out.match(text, pos)
            if m:
                self.program.append(
                    {"k": "out", "field": m.group(1), "idx": int(m.group(2)), "arg": self.arg(m.group(3))}
                )
                pos = m.end()
                continue
            m = re_if.match(text, pos)
            if m:
                name, fn = m.group(1), m.group(2)
                args, pos = self._call_args(text, m.end())
                # closing paren of if(...)
                mm = re.compile(r"\s*\)\s*\{").match(text, pos)
                if not mm:
                    raise ValueError("if-tail? " + text[pos : pos + 80])
                pos = mm.end()
                depth += 1
                item = self._check(fn, args)
                if depth == 1:
                    if item["k"] == "attach_env":
                        group = {"k": "attach_env", "fine": None}
                    else:
                        group = dict(item, name=(name or "").strip(), fine=[])
                        group["guard"] = group.pop("sphere")
                    self.program.append(group)
                else:
                    if group is None or group.get("fine") is None:
                        raise ValueError("nested check outside group")
                    group["fine"].append(item["sphere"])
                continue
            mm = re.compile(r"\s*return\s+(true|false)\s*;").match(text, pos)
            if mm:
                pos = mm.end()
                continue
            mm = re.compile(r"\s*return\s+not\s+").match(text, pos)
            if mm:  # robots::sphere: `return not sphere_environment_in_collision(...)`
                mc = re_call.match(text, mm.end())
                args, pos = self._call_args(text, mc.end())
                pos = re.compile(r"\s*;").match(text, pos).end()
                item = self._check(mc.group(1), args)
                # a group whose guard is its single fine sphere
                self.program.append({"k": "env", "name": "sphere", "guard": item["sphere"], "fine": [item["sphere"]]})
                continue
            mm = re.compile(r"\s*\}").match(text, pos)
            if mm:
                pos = mm.end()
                depth -= 1
                if depth == 0:
                    # unnamed top-level `if (check) { return false; }` (fetch): a group whose
                    # guard is its only fine test
                    if group is not None and group.get("fine") == [] and group.get("name") == "":
                        group["fine"] = [group["guard"]]
                        group["leaf"] = True
                    group = None
                continue
            mc = re_call.match(text, pos)
            if mc and mc.group(1) == "set_attachment_pose":
                args, pos = self._call_args(text, mc.end())
                pos = re.compile(r"\s*;").match(text, pos).end()
                self.program.append({"k": "attach_pose", "args": [self.arg(a) for a in args[1:]]})
                continue
            raise ValueError("statement? " + text[pos : pos + 120])
        return {"values": self.values, "program": self.program}

    def _call_args(self, text: str, pos: int):
        """Parse a comma separated argument list up to the matching ')'."""
        depth = 1
        start = pos
        args = []
        while depth:
            ch = text[pos]
            if ch in "([":
                depth += 1
            elif ch in ")]":
                depth -= 1
                if depth == 0:
                    args.append(text[start:pos])
                    pos += 1
                    break
            elif ch == "," and depth == 1:
                args.append(text[start:pos])
                start = pos + 1
            pos += 1
        return [a.strip() for a in args if a.strip() != ""], pos

    def _check(self, fn: str, args: list[str]):
        if fn == "sphere_environment_in_collision":
            assert args[0] == "environment" and len(args) == 5, args
            return {"k": "env", "sphere": [self.arg(a) for a in args[1:]]}
        if fn == "sphere_sphere_self_collision":
            assert len(args) == 8, args
            a = [self.arg(x) for x in args]
            return {"k": "self", "sphere": [a[:4], a[4:]]}
        if fn == "attachment_sphere_collision":
            assert args[0] == "environment" and len(args) == 5, args
            return {"k": "attach_self", "sphere": [self.arg(a) for a in args[1:]]}
        if fn == "attachment_environment_collision":
            return {"k": "attach_env", "sphere": None}
        raise ValueError("check? " + fn)


_ROBOT_META = {
    # name: (dimension, resolution)  -- robots/{panda_base,ur5,fetch,baxter,sphere}.hh
    "panda": (7, 32),
    "ur5": (6, 32),
    "fetch": (8, 32),
    "baxter": (14, 64),
    "sphere": (3, 32),
}


def _scaling(text: str, dim: int):
    """Joint ranges from ``scale_configuration_block`` (q[i] = lo + q[i]*range)."""
    try:
        body = _function_body(text, "scale_configuration_block")
    except KeyError:
        return None
    lo, rng = [None] * dim, [None] * dim
    for m in re.finditer(r"q\[(\d+)\]\s*=\s*(" + _NUM + r")\s*\+\s*\(q\[\d+\]\s*\*\s*(" + _NUM + r")\)", body):
        i = int(m.group(1))
        lo[i] = float(m.group(2).rstrip("fF"))
        rng[i] = float(m.group(3).rstrip("fF"))
    if any(v is None for v in lo):
        return None
    out = {"lo": lo, "range": rng}
    # descale_configuration: q' = inv * (q - lo) with `inv` a literal (used by the joint-limit test,
    # bindings/common.hh:176-179)
    try:
        dbody = _function_body(text, "descale_configuration_block")
        inv = [None] * dim
        for m in re.finditer(r"q\[(\d+)\]\s*=\s*(" + _NUM + r")\s*\*\s*\(q\[\d+\]\s*-\s*(" + _NUM + r")\)", dbody):
            inv[int(m.group(1))] = float(m.group(2).rstrip("fF"))
        if all(v is not None for v in inv):
            out["inv"] = inv
    except KeyError:
        pass
    return out


def parse_robot(robot: str) -> dict[str, Any]:
    """Parse every traced function of ``robots/<robot>/fk.hh`` into IR."""
    path = os.path.join(vamp_src_root(), "robots", robot, "fk.hh")
    with open(path) as fh:
        text = fh.read()
    dim, res = _ROBOT_META[robot]
    n_spheres = int(re.search(r"n_spheres\s*=\s*(\d+)", text).group(1))
    ir: dict[str, Any] = {
        "robot": robot,
        "dimension": dim,
        "resolution": res,
        "n_spheres": n_spheres,
        "source": f"robots/{robot}/fk.hh",
        "functions": {},
    }
    sc = _scaling(text, dim)
    if sc is not None:
        ir["scale"] = sc
    elif robot == "sphere":
        ir["scale"] = {"lo": [-10.0, -10.0, 0.0], "range": [20.0, 20.0, 5.0]}
    for fn in ("sphere_fk", "interleaved_sphere_fk", "interleaved_sphere_fk_attachment"):
        try:
            # the function NAME must be followed by '(' directly (skips the forwarder templates)
            body = _function_body(text, "inline void " + fn if fn == "sphere_fk" else "inline bool " + fn)
        except KeyError:
            continue
        ir["functions"][fn] = _Parser(body).run()
    return ir


def save_ir(ir: dict[str, Any], path: str) -> None:
    with open(path, "w") as fh:
        json.dump(ir, fh, separators=(",", ":"))


def load_ir(path: str) -> dict[str, Any]:
    with open(path) as fh:
        return json.load(fh)


if __name__ == "__main__":
    import sys

    for r in sys.argv[1:] or ["sphere", "ur5", "panda", "fetch", "baxter"]:
        ir = parse_robot(r)
        for fn, f in ir["functions"].items():
            kinds: dict[str, int] = {}
            for it in f["program"]:
                kinds[it["k"]] = kinds.get(it["k"], 0) + 1
            ops: dict[str, int] = {}
            for v in f["values"]:
                ops[v[0]] = ops.get(v[0], 0) + 1
            fine_env = sum(len(it["fine"]) for it in f["program"] if it["k"] == "env")
            fine_self = sum(len(it["fine"]) for it in f["program"] if it["k"] == "self")
            print(r, fn, kinds, ops, "fine_env", fine_env, "fine_self", fine_self)
