"""Registers / spills (ptxas log) and static SASS instruction counts (cuobjdump) of the ray-march kernels, per flavour."""
import re
import subprocess
import sys
import os

D = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "volume_renderer_b200", "csrc")
pat = sys.argv[1] if len(sys.argv) > 1 else "Li2EN"  # LAYOUT_TEXHW kernels
for a in (0, 1, 2):
    for g in ("lookup", "compute"):
        log = open(os.path.join(D, "raymarch_%s_a%d.ptxas.log" % (g, a))).read()
        for m in re.finditer(r"Compiling entry function '([^']+)'[^\n]*\n(?:ptxas info[^\n]*\n)*?\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers", log):
            name = m.group(1)
            if pat not in name:
                continue
            obj = os.path.join(D, "raymarch_%s_a%d.o" % (g, a))
            sass = subprocess.run(["cuobjdump", "-sass", "-fun", name, obj], stdout=subprocess.PIPE, text=True).stdout
            ins = [l for l in sass.splitlines() if re.match(r"\s+/\*[0-9a-f]{4}\*/", l)]
            ops = {}
            for l in ins:
                op = l.split("*/")[1].strip().split()[0].rstrip(";")
                if op.startswith("@"):
                    op = l.split("*/")[1].strip().split()[1].rstrip(";")
                op = op.split(".")[0]
                ops[op] = ops.get(op, 0) + 1
            top = sorted(ops.items(), key=lambda kv: -kv[1])[:12]
            print("%s a%d %s: regs %s, spill st/ld %s/%s, SASS %d; %s" % (g, a, name[-28:], m.group(5), m.group(3), m.group(4), len(ins), " ".join("%s=%d" % kv for kv in top)))
