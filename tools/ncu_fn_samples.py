#!/usr/bin/env python3
"""Per-instruction stall samples of ONE device function inside one kernel of an ncu report (SASS source page).

usage: ncu_fn_samples.py <report.ncu-rep> <lib.so> <kernel-substring> <function-substring> [-v]
Prints the totals per stall reason for the function and, with -v, one line per instruction.  Function boundaries come
from nvdisasm's symbol order (the report has none)."""
import csv, io, os, re, subprocess, sys, tempfile

rep, so, kern, fn = sys.argv[1:5]
verbose = len(sys.argv) > 5
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
lay, cur, pending = [], None, None
for line in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m:
        cur = m.group(1); pending = "<body>" if kern in cur else None; continue
    if cur is None or kern not in cur:
        continue
    m = re.match(r"^\$" + re.escape(cur) + r"\$(\S+):", line)
    if m:
        pending = m.group(1); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", line)
    if m and pending is not None:
        lay.append((int(m.group(1), 16), pending)); pending = None
rng = None
for i, (o, f) in enumerate(lay):
    if fn in f:
        rng = (o, lay[i + 1][0] if i + 1 < len(lay) else 1 << 30); break
if rng is None:
    sys.exit("function not found in kernel; have: " + ", ".join(f for _, f in lay))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kernel, hdr, out, seen = None, None, [], False
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == "Kernel Name":
        if seen:
            break                      # first launch of the kernel only
        kernel = r[1]; hdr = None; continue
    if r[0] == "Address":
        hdr = r; continue
    if kernel and kern in kernel and hdr:
        out.append(r); H = hdr; seen = True
base = int(out[0][0], 16)
cols = ["# Samples", "Instructions Executed", "stall_long_sb", "stall_wait", "stall_math", "stall_short_sb", "stall_branch_resolving",
        "stall_no_inst", "stall_dispatch", "stall_not_selected", "stall_barrier", "stall_lg"]
idx = [H.index(c) for c in cols]
print("off    " + " ".join(c.replace("stall_", "")[:8].rjust(8) for c in cols))
tot = [0] * len(cols)
for r in out:
    off = int(r[0], 16) - base
    if rng[0] <= off < rng[1]:
        v = [int(r[i] or 0) for i in idx]
        tot = [a + b for a, b in zip(tot, v)]
        if verbose:
            print("%06x " % off + " ".join(str(x).rjust(8) for x in v) + "  " + r[H.index("Source")][:60])
print("TOTAL  " + " ".join(str(x).rjust(8) for x in tot))
