"""Registers / spills / stack per kernel from the build's ptxas -v log -> a table (profiles/r2_ptxas.txt).
    python tools/ptxas_table.py [unity-raytracer_b200/build/ptxas.log]"""
import re, subprocess, sys
path = sys.argv[1] if len(sys.argv) > 1 else "unity-raytracer_b200/build/ptxas.log"
txt = open(path).read()
rows = []
for m in re.finditer(r"Compiling entry function '([^']+)' for 'sm_100a'\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", txt):
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    name = re.sub(r"\(.*", "", name).replace("void ", "")
    if name.startswith("cub::") or "cub::" in name:
        name = "cub::" + re.sub(r"<.*", "", name.split("::")[-1])
    rows.append((name, int(m.group(5)), int(m.group(2)), int(m.group(3)), int(m.group(4))))
print(f"{'kernel':64s} {'regs':>5s} {'stack B':>8s} {'spill st B':>10s} {'spill ld B':>10s}")
for r in rows:
    print(f"{r[0]:64s} {r[1]:5d} {r[2]:8d} {r[3]:10d} {r[4]:10d}")
