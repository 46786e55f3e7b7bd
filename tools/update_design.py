#!/usr/bin/env python
"""update_design.py: regenerates DESIGN.md section 6.1 (between the SWEEP markers) from the committed bench lines."""
import os
import subprocess
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
n1 = os.path.join(ROOT, "profiles", "bench", "r2_bench_n1_final.json")
more = [os.path.join(ROOT, "profiles", "bench", "r2_bench_n%d_final.json" % k) for k in (2, 4, 8)]
args = [sys.executable, os.path.join(ROOT, "tools", "sweep_table.py"), n1] + [m for m in more if os.path.exists(m)]
tab = subprocess.run(args, capture_output=True, text=True, check=True).stdout
p = os.path.join(ROOT, "DESIGN.md")
s = open(p).read()
a, b = s.index("<!-- SWEEP:BEGIN -->") + len("<!-- SWEEP:BEGIN -->"), s.index("<!-- SWEEP:END -->")
open(p, "w").write(s[:a] + "\n" + tab + s[b:])
print("DESIGN.md 6.1 regenerated from", os.path.basename(n1))
