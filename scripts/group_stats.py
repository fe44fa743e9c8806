"""Dev tool: per-group guard / fine-hit statistics of a robot trace over random configurations
(numpy interpretation of the IR; self groups only + env guards vs an MBM scene)."""
import sys
import numpy as np
from mr_vamp_b200.codegen import trace_parser, ir_eval

robot = sys.argv[1] if len(sys.argv) > 1 else "panda"
n = 20000
ir = trace_parser.parse_robot(robot)
fn = ir["functions"]["interleaved_sphere_fk"]
lo = np.array(ir["scale"]["lo"], np.float32); rg = np.array(ir["scale"]["range"], np.float32)
rng = np.random.default_rng(0)
q = (lo + rng.random((n, ir["dimension"]), dtype=np.float32) * rg).astype(np.float32)
vals = ir_eval.evaluate(fn, q)
C = lambda s: ir_eval.centre(vals, s, n)
collided = np.zeros(n, bool)
tot_pairs = 0
print("group, guard-rate, fine pairs, hit-rate | cumulative collided")
for it in fn["program"]:
    if it["k"] != "self":
        continue
    ga, gb = it["guard"]
    d = np.linalg.norm(C(ga) - C(gb), axis=1)
    fired = d < ga[3]["c"] + gb[3]["c"]
    hit = np.zeros(n, bool)
    # a-side survivors vs bounding gb
    A = {}; B = {}
    for a, b in it["fine"]:
        ka = (a[0]["v"], a[1]["v"], a[2]["v"], a[0]["c"], a[1]["c"], a[2]["c"]); kb = (b[0]["v"], b[1]["v"], b[2]["v"], b[0]["c"], b[1]["c"], b[2]["c"])
        A[ka] = a; B[kb] = b
        dd = np.linalg.norm(C(a) - C(b), axis=1)
        hit |= dd < a[3]["c"] + b[3]["c"]
    a_surv = np.zeros(n)
    for a in A.values():
        a_surv += (np.linalg.norm(C(a) - C(gb), axis=1) < a[3]["c"] + gb[3]["c"] + 6e-4)
    hit &= fired
    collided |= hit
    print(f"{it['name']:40s} guard {fired.mean():.3f}  pairs {len(it['fine']):4d} (A {len(A)} x B {len(B)})  a-surv|fired {a_surv[fired].mean() if fired.any() else 0:.2f}  hit {hit.mean():.3f} | {collided.mean():.3f}")
    tot_pairs += fired.mean() * len(it["fine"])
print("expected fine pairs per config (reference semantics, no early exit):", tot_pairs)
