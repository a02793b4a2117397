"""Opcode histogram per kernel of libjrl_lbm.so from `cuobjdump -sass`: static instruction count,
TMA bulk copies (UBLKCP), mbarrier operations (SYNCS), 128-bit loads / stores, local-memory spills (STL / LDL),
reciprocals and division checks (MUFU.RCP, FCHK) and slow-path calls (CALL).  Writes profiles/r2_sass_summary.txt.

    python tools/sass_summary.py [path/to/libjrl_lbm.so] > profiles/r2_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "jrl-lbm_b200", "libjrl_lbm.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
elf = subprocess.run(["cuobjdump", "--list-elf", lib], capture_output=True, text=True).stdout
KEYS = ["UBLKCP", "SYNCS", "LDG.E.128", "STG.E.128", "LDG", "STG", "LDS", "STS", "SHFL", "STL", "LDL", "MUFU.RCP", "FCHK", "CALL",
        "FFMA", "FMUL", "FADD", "DFMA", "DADD", "DMUL", "BRA", "ATOM", "RED"]
stats, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        stats[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        stats[cur]["instr"] += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                stats[cur][k] += 1
print("# cuobjdump -sass opcode histogram of", os.path.relpath(lib, ROOT))
print("# architectures:", sorted(set(re.findall(r"sm_\d+a?", elf))))
demangle = subprocess.run(["c++filt"], input="\n".join(stats), capture_output=True, text=True).stdout.splitlines()
for (f, c), name in zip(stats.items(), demangle):
    cols = " ".join(f"{k}={c[k]}" for k in KEYS if c[k])
    print(f"{name[:110]:<110} instr={c['instr']:<6} {cols}")
