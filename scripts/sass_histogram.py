"""Dev tool: opcode histogram + the TMA / mbarrier lines of the shipped Panda fkcc kernel -> profiles/<round>_panda_fkcc_sass.txt
usage: python scripts/sass_histogram.py [round-prefix]"""
import collections
import re
import subprocess
import sys

rp = sys.argv[1] if len(sys.argv) > 1 else "r2"
sass = subprocess.run("cuobjdump -sass mr_vamp_b200/lib/libvamp_b200.so", shell=True, capture_output=True, text=True).stdout
m = re.search(r"Function : (_ZN2vb11fkcc_kernelINS_10PandaTraceELi0EEEvNS_10LaunchArgsE)\n(.*?)(?=\n\s*Function :|\Z)", sass, re.S)
body = m.group(2)
ops = collections.Counter()
special = []
for line in body.splitlines():
    mm = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not mm:
        continue
    op = mm.group(2)
    ops[op.split(".")[0]] += 1
    if op.startswith(("UBLKCP", "SYNCS", "FENCE.VIEW", "UTMA")):
        special.append(" " + re.sub(r"\s+", " ", line.strip()))
out = [f"SASS opcode histogram of the shipped vb::fkcc_kernel<vb::PandaTrace, 0> (cuobjdump -sass mr_vamp_b200/lib/libvamp_b200.so, "
       f"sm_100a cubin): {sum(ops.values())} instructions"]
out += [f"{n:7d} {op}" for op, n in ops.most_common(32)]
out += ["", "TMA bulk copy / mbarrier / proxy-fence instructions of the staged-configuration pipeline:"] + special
open(f"profiles/{rp}_panda_fkcc_sass.txt", "w").write("\n".join(out) + "\n")
print("\n".join(out[:12]))
