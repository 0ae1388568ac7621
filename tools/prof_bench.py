"""Per-kernel attribution of one bench.py run (torch.profiler/CUPTI; numbers are for attribution only)."""
import runpy
import sys

import torch
from torch.profiler import ProfilerActivity, profile

out = sys.argv[1]
sys.argv = ["bench.py"] + sys.argv[2:]
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    try:
        runpy.run_path("bench.py", run_name="__main__")
    except SystemExit:
        pass
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total / max(e.count, 1), e.device_time_total) for e in prof.key_averages()]
rows.sort(key=lambda r: -r[3])
with open(out, "w") as f:
    for k, c, avg, tot in rows[:40]:
        f.write(f"{tot:12.1f} us  n={c:6d}  avg={avg:8.2f} us  {k[:150]}\n")
