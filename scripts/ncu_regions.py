#!/usr/bin/env python
"""Where do the warps of a kernel spend their time?  Splits the per-instruction stall samples of an .ncu-rep
(`--page source`) into code regions delimited by marker instructions (tensor-memory loads / stores, MMAs, bulk
copies, votes, sleeps), so warp-specialised kernels can be read role by role.

  python scripts/ncu_regions.py gpurun_out/prof.ncu-rep [--top 25]
"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 20
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(src)))
h, data = r[1], r[2:]
iS, iI, iSrc, iA = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source"), h.index("Address")
tot = sum(int(x[iS]) for x in data) or 1
print(f"total stall samples {tot}, {len(data)} SASS instructions\n")
print("regions (rows of the SASS listing up to and including the marker instruction):")
MARK = r"LDTM|STTM|UTCHMMA|UBLKCP|UTMALDG|UTCBAR|USETMAXREG|REDG|ATOM|EXIT|NANOSLEEP|VOTE|HMMA|BAR\.SYNC|LDGMC"
out, last = [], 0
for i, x in enumerate(data):
    s = x[iSrc]
    if re.search(MARK, s):
        op = [t for t in s.split() if not t.startswith("@")][0]
        out.append([last, i, op, sum(int(y[iS]) for y in data[last:i + 1]), sum(int(y[iI]) for y in data[last:i + 1])])
        last = i + 1
out.append([last, len(data) - 1, "END", sum(int(y[iS]) for y in data[last:]), 0])
merged = []
for o in out:
    if merged and merged[-1][2] == o[2] and re.match(r"UTCHMMA|STTM|LDTM|HMMA", o[2]):
        merged[-1][1] = o[1]; merged[-1][3] += o[3]; merged[-1][4] += o[4]
    else:
        merged.append(o)
for a in merged:
    if a[3] > tot * 0.003:
        print(f"  rows {a[0]:5d}-{a[1]:5d}  up to {a[2]:30s} {a[3]:8d} samples {100 * a[3] / tot:5.1f} %   {a[4]:>12d} warp-instructions")
print(f"\ntop {top_n} instructions by samples:")
for x in sorted(data, key=lambda x: -int(x[iS]))[:top_n]:
    print(f"  {int(x[iS]):7d} {100 * int(x[iS]) / tot:5.1f} %  executed {x[iI]:>10s}  {x[iSrc][:100]}")
