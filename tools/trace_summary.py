"""Summarise a kernel timeline written by bench.py (SINGA_B200_TRACE_OUT=<prefix>): total device time per kernel
name over the traced steps, sorted, plus the span of one step.  Usage: python tools/trace_summary.py <file> [steps]"""
import collections
import sys

path = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
tot = collections.defaultdict(float)
cnt = collections.Counter()
t_first, t_last = None, 0.0
for ln in open(path):
    f = ln.split(None, 3)
    if len(f) < 4:
        continue
    start, dur, name = float(f[0]), float(f[1]), f[3].strip()
    name = name.split("(")[0][:90]
    tot[name] += dur
    cnt[name] += 1
    t_first = start if t_first is None else t_first
    t_last = max(t_last, start + dur)
print(f"span {(t_last - t_first) / steps:9.1f} us/step   busy {sum(tot.values()) / steps:9.1f} us/step")
for name, t in sorted(tot.items(), key=lambda kv: -kv[1])[:28]:
    print(f"{t / steps:9.1f} us/step  x{cnt[name] / steps:5.1f}  {name}")
