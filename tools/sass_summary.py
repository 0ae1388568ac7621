"""Per-kernel counts of the SASS mnemonics that show a Blackwell-native kernel (tcgen05 MMA / TMEM loads / TMA incl.
im2col mode / multimem / reductions) in the built library.  Needs only cuobjdump (no GPU).
Usage: python tools/sass_summary.py [libsinga_b200.so] > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

lib = sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).resolve().parent.parent / "singa_b200/lib/libsinga_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()   # noqa: E731
pat = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTCBAR[.\w]*|"
                 r"UTCATOMSWS[.\w]*|LDGMC[.\w]*|STG\.E[.\w]*MC[.\w]*|MULTIMEM[.\w]*|REDG[.\w]*|SYNCS[.\w]*|HMMA[.\w]*|"
                 r"UCGABAR[.\w]*|ACQBULK|MEMBAR[.\w]*)")
cur, counts = None, collections.OrderedDict()
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur:
        for tok in pat.findall(ln):
            counts[cur][tok] += 1
print(f"# {lib}: {len(counts)} kernels; mnemonics per kernel (only kernels that use any of them)")
for name, c in counts.items():
    if not c:
        continue
    d = demangle(name)
    d = re.sub(r"\(CUtensorMap_st.*", "", d)
    print(f"\n{d[:150]}")
    print("   " + ", ".join(f"{k} x{v}" for k, v in sorted(c.items())))
