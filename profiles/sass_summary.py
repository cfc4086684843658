"""Per-kernel SASS opcode counts of libqpolicy.so (cuobjdump -sass): which kernels carry tcgen05 MMAs (UTCHMMA), TMEM
loads (LDTM), TMA loads / stores / reduce-adds (UTMALDG / UTMASTG / UTMAREDG), bulk copies (UBLKCP), cp.async (LDGSTS) ...
    python profiles/sass_summary.py > profiles/r2_sass_opcodes.csv
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "scala_akka_tensorflow_sharetrade_helper_b200", "libqpolicy.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
OPS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "SYNCS", "UCGABAR_ARV", "UCGABAR_WAIT", "HMMA",
       "FFMA", "LDGSTS", "LDS", "STS", "LDG", "STG", "RED", "ATOMG", "MEMBAR", "BAR", "MUFU", "DADD", "DMUL"]
rows = []
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0].strip()
    cnt, n = collections.Counter(), 0
    for line in f.split("\n"):
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            n += 1
            cnt[m.group(1).split(".")[0]] += 1
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    rows.append((re.sub(r"\(.*", "", dem), n, cnt))
print("kernel,instructions," + ",".join(OPS))
for name, n, cnt in sorted(rows, key=lambda r: -r[1]):
    print(f'"{name}",{n},' + ",".join(str(cnt.get(k, 0)) for k in OPS))
