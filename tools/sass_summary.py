#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-specific SASS mnemonics in the built library (evidence that the hot kernels
are tcgen05 / TMEM / TMA code):   python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "multimodal-fusion_b200", "hg_b200", "libhg_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WANT = OrderedDict([("UTCHMMA", r"\bUTCHMMA"), ("UTCHMMA.2CTA", r"\bUTCHMMA\.2CTA"), ("LDTM (tcgen05.ld)", r"\bLDTM"),
                    ("UTMALDG (TMA load)", r"\bUTMALDG"), ("UTMALDG.MULTICAST", r"\bUTMALDG\.\w+\.MULTICAST|UTMALDG.*MULTICAST"),
                    ("UTCBAR (tcgen05.commit)", r"\bUTCBAR"), ("SYNCS (mbarrier)", r"\bSYNCS"),
                    ("UTCATOMSWS (TMEM alloc)", r"\bUTCATOMSWS"), ("FFMA2 (packed FP32)", r"\bFFMA2"), ("HMMA (legacy mma.sync)", r"\bHMMA")])
cur, rows = None, OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        rows[cur] = OrderedDict((k, 0) for k in WANT)
        rows[cur]["instructions"] = 0
        continue
    if cur is None or "/*" not in line:
        continue
    if re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", line):
        rows[cur]["instructions"] += 1
    for k, pat in WANT.items():
        if re.search(pat, line):
            rows[cur][k] += 1
demangle = subprocess.run(["cu++filt"] + list(rows), capture_output=True, text=True).stdout.splitlines()
print(f"cuobjdump -sass {os.path.relpath(so, ROOT)}: {len(rows)} kernels; Blackwell mnemonics per kernel (kernels without any omitted)\n")
tot = OrderedDict((k, 0) for k in WANT)
for (name, r), pretty in zip(rows.items(), demangle):
    for k in WANT:
        tot[k] += r[k]
    if not any(r[k] for k in WANT if k != "SYNCS (mbarrier)" and k != "FFMA2 (packed FP32)"):
        continue
    short = re.sub(r"\(.*", "", re.sub(r"\((int|bool)\)", "", pretty))[:70]
    print(f"{short:70s} " + "  ".join(f"{k.split(' ')[0]}={r[k]}" for k in WANT if r[k]) + f"  [{r['instructions']} instr]")
print("\ntotals over the library: " + ", ".join(f"{k}: {v}" for k, v in tot.items()))
