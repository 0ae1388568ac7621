"""Markdown table of the benchmark lines under profiles/matrix (one JSON line per file, written by tools/run_matrix.sh)."""
import glob
import json
import os
import re

rows = {}
for f in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "..", "profiles", "matrix", "*.json"))):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        continue
    m = re.match(r"(.+)_n(\d+)_(.+)_(fp32|bf16)\.json", os.path.basename(f))
    if not m:
        continue
    cfg, n, arm, dt = m.group(1), int(m.group(2)), m.group(3), m.group(4)
    rows[(cfg, dt, n, arm)] = (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"])
print("| config | dtype | GPUs | arm | ms/step (device) | images/s | e2e ms/step |")
print("|---|---|---|---|---|---|---|")
for (cfg, dt, n, arm), (ms, v, e) in sorted(rows.items()):
    print(f"| {cfg} | {dt} | {n} | {arm} | {ms:.4f} | {v:,.0f} | {e:.4f} |")
