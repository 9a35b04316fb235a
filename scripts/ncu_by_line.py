#!/usr/bin/env python
"""Per-source-line instruction and stall-sample counts of one kernel, read offline.

  python scripts/ncu_by_line.py <prof.ncu-rep> <lib.so> <kernel substring> [top N]

`ncu --page source --csv` lists SASS instructions with their executed counts; `nvdisasm -gi` of the same cubin gives
the source line (and the inlining chain) of every SASS instruction.  The two lists are in the same order, so they are
joined by position.  Each instruction is attributed to the OUTERMOST line of its inlining chain that lies in the
kernel's own file (so helper code is charged to the call site), and also to its innermost line."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, so, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-sf", cubin], capture_output=True, text=True).stdout

# ---- walk the disassembly of the kernel's .text section ------------------------------------------------------
lines = dis.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith("//") and ".text." in l and kname in l)
sass = []                      # (opcode text, [(file, line), ...] innermost first, outermost last)
chain, fresh = [], True        # fresh: the next //## line starts a new annotation group
for l in lines[start + 1:]:
    if l.startswith("//---------------------"):
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        if fresh:
            chain = [(os.path.basename(m.group(1)), int(m.group(2)))]
            fresh = False
        if m.group(3):
            chain.append((os.path.basename(m.group(3)), int(m.group(4))))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        sass.append((m.group(2).strip(), list(chain)))
        fresh = True

# NCU_LAUNCH=<n>: the n-th captured launch of a report that holds several kernels (default: the first)
sel = ["--launch-skip", os.environ["NCU_LAUNCH"], "--launch-count", "1"] if os.environ.get("NCU_LAUNCH") else []
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + sel, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ia, isrc, isamp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
iwf = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) > ia and r[ia].isdigit()]
if len(body) != len(sass):
    print(f"warning: {len(body)} profiled instructions vs {len(sass)} disassembled; joining by position up to the shorter", file=sys.stderr)

kfile = None
for _, ch in sass:
    if ch:
        kfile = ch[-1][0]
        break
outer = collections.defaultdict(lambda: [0, 0, 0, collections.Counter(), collections.Counter()])
total_i = total_s = total_w = 0
for r, (op, ch) in zip(body, sass):
    n, s = int(r[ia]), int(r[isamp])
    w = int(r[iwf]) if iwf is not None and r[iwf].isdigit() else 0
    total_i += n; total_s += s; total_w += w
    key = ch[-1] if ch else ("?", 0)                     # outermost frame = the line of the kernel body
    o = outer[key]
    o[0] += n; o[1] += s; o[2] += w
    opc = (op.split()[1] if op.startswith("@") else op.split()[0]).split(".")[0]
    o[3][opc] += n
    for i, h in stall_cols:
        if r[i].isdigit() and int(r[i]):
            o[4][h[6:]] += int(r[i])
print(f"kernel {kname}: {total_i / 1e6:.1f} M warp instructions, {total_s} stall samples, {total_w / 1e6:.1f} M shared wavefronts")
print("file:line   inst M  inst%  samp%  wavef M  | top opcodes | top stalls")
for key, o in sorted(outer.items(), key=lambda kv: -kv[1][0])[:top]:
    ops = " ".join(f"{k}:{v / 1e6:.0f}" for k, v in o[3].most_common(5))
    st = " ".join(f"{k}:{100 * v / max(1, o[1]):.0f}%" for k, v in o[4].most_common(3))
    print(f"{key[0]}:{key[1]:<5d} {o[0] / 1e6:7.1f} {100 * o[0] / total_i:5.1f}% {100 * o[1] / max(1, total_s):5.1f}% {o[2] / 1e6:7.1f}  | {ops} | {st}")
