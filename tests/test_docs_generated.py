"""DESIGN.md section 6.1 is GENERATED from the committed bench lines (tools/update_design.py -> tools/sweep_table.py), not typed:
regenerating it from profiles/bench/r2_bench_n{1,2,4,8}_final.json must reproduce the text between the SWEEP markers, and the
bench line must carry what the measurement contract asks for."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "profiles", "bench")


def test_design_6_1_matches_the_bench_lines():
    lines = [os.path.join(BENCH, "r2_bench_n%d_final.json" % k) for k in (1, 2, 4, 8)]
    lines = [p for p in lines if os.path.exists(p)]
    assert lines and lines[0].endswith("n1_final.json")
    tab = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sweep_table.py")] + lines, capture_output=True, text=True,
                         check=True).stdout
    s = open(os.path.join(ROOT, "DESIGN.md")).read()
    a, b = s.index("<!-- SWEEP:BEGIN -->") + len("<!-- SWEEP:BEGIN -->"), s.index("<!-- SWEEP:END -->")
    assert s[a:b].strip() == tab.strip()


def test_bench_line_keys():
    d = json.load(open(os.path.join(BENCH, "r2_bench_n1_final.json")))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "clocks", "gpu_launches", "sweep"):
        assert k in d, k
    assert d["roofline"]["bound"] == "hbm" and 0 < d["roofline"]["frac"] < 1 and d["roofline"]["traffic"] > 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert len(d["sweep"]) >= 40 and all(e["oracle_check"] for e in d["sweep"])
    assert all(r["oracle_check"] for r in d.get("n4", []))
    r = json.load(open(os.path.join(BENCH, "r2_bench_reference_final.json")))
    assert r["impl"] == "reference" and r["metric"] == d["metric"] and r["unit"] == d["unit"]
    assert r["config"]["workload"] == d["config"]["workload"]
