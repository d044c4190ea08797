"""Per-kernel counts of the SASS mnemonics that show which hardware path a kernel uses (tcgen05 MMA, TMEM loads, TMA
loads / stores, legacy mma.sync), from `cuobjdump -sass` of the built library.

    python scripts/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

lib = Path(__file__).resolve().parent.parent / "woft_b200" / "libwoft_b200.so"
out = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
pat = {"UTCHMMA (tcgen05.mma f16/tf32)": r"\bUTC[A-Z]*MMA", "LDTM (tcgen05.ld)": r"\bLDTM", "UTMALDG (TMA load)": r"\bUTMALDG",
       "UTMASTG (TMA store)": r"\bUTMASTG", "UBLKCP (bulk copy)": r"\bUBLKCP", "SYNCS (mbarrier)": r"\bSYNCS",
       "HMMA (mma.sync)": r"\bHMMA", "MUFU": r"\bMUFU", "FFMA2": r"\bFFMA2", "STG": r"\bSTG", "LDG": r"\bLDG"}
counts, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur:
        for name, rx in pat.items():
            if re.search(rx, line):
                counts[cur][name] += 1
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            counts[cur]["instructions"] += 1
print(f"cuobjdump -sass {lib.name}: {len(counts)} kernels (sm_100a)\n")
cols = list(pat)
print(f"{'kernel':70s} {'instr':>6s} " + " ".join(f"{c.split()[0]:>8s}" for c in cols))
for k, c in counts.items():
    print(f"{k[:70]:70s} {c['instructions']:6d} " + " ".join(f"{c[n]:8d}" for n in cols))
