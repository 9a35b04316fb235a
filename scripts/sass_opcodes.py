#!/usr/bin/env python
"""Opcode histogram of every kernel in libmixfe.so plus the lines that prove the Blackwell-native paths
(UBLKCP = TMA bulk copy, SYNCS = mbarrier, FADD2 / FFMA2 / FMUL2 = packed f32x2).  No GPU needed:
  python scripts/sass_opcodes.py > profiles/sass_opcodes.txt        (re-run whenever csrc/ changes)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "mixing_style_transfer_b200", "libmixfe.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
print(f"# {os.path.relpath(so, ROOT)}: cuobjdump -sass, architectures {arch}")
kernels = collections.OrderedDict()
cur = None
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = kernels.setdefault(m.group(1), collections.Counter())
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
    if m and cur is not None:
        cur[m.group(1)] += 1
NATIVE = ("UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "FADD2", "FFMA2", "FMUL2", "UTCHMMA", "LDTM", "HMMA")
for name, ops in kernels.items():
    short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    total = sum(ops.values())
    print(f"\n## {short}: {total} instructions")
    print("   " + "  ".join(f"{k}:{v}" for k, v in ops.most_common(14)))
    nat = {k: ops[k] for k in NATIVE if ops.get(k)}
    if nat:
        print("   native: " + "  ".join(f"{k}:{v}" for k, v in nat.items()))
