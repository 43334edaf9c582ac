#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of the built library (evidence that the tensor-core / TMA paths are real).

  python scripts/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "fedmse_decentralized_b200/_C/libfedmse_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
counts, cur = collections.OrderedDict(), None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        keep_full = op.startswith(("UTC", "UTMA", "UBLK", "LDTM", "STTM", "SYNCS", "RED", "HMMA", "LDSM", "USETMAXREG", "LDS.128", "LDG.E.STRONG"))
        counts[cur][op if keep_full else op.split(".")[0]] += 1
print("SASS evidence per kernel: cuobjdump -sass of libfedmse_b200.so (sm_100a), counts of selected mnemonics.")
print("tcgen05.mma -> UTCHMMA, tcgen05.ld/st -> LDTM/STTM, tcgen05.commit -> UTCBAR, TMA tensor loads -> UTMALDG,")
print("cp.async.bulk -> UBLKCP, mbarrier -> SYNCS.*, red.release.sys -> REDG...SYS, ld.acquire.sys -> LDG.E.STRONG.SYS,")
print("mma.sync tf32 -> HMMA.1688.F32.TF32, ldmatrix -> LDSM, setmaxnreg -> USETMAXREG\n")
for f, c in counts.items():
    name = subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip()
    sel = {k: v for k, v in c.items() if k.startswith(("UTC", "UTMA", "UBLK", "LDTM", "STTM", "SYNCS", "RED", "HMMA", "LDSM", "USETMAXREG", "LDG.E.STRONG"))
           or k in ("FFMA", "FFMA2", "FMUL2", "FADD2", "LDS.128", "BAR", "SHFL", "MUFU")}
    print(f"{name[:120]}\n    {sum(c.values())} instructions: " + ", ".join(f"{k} {v}" for k, v in sorted(sel.items(), key=lambda kv: -kv[1])))
