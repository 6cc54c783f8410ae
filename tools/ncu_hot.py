#!/usr/bin/env python3
"""Per-function dynamic profile from an ncu report's SASS source page: executed warp-instructions and
stall samples aggregated by device function (function boundaries from nvdisasm symbol order are not in
the report, so functions are delimited by RET / the first instruction after it).
usage: ncu_hot.py <report.ncu-rep> <lib.so>"""
import csv, io, subprocess, sys, re, os, tempfile, collections
rep, so = sys.argv[1], sys.argv[2]
# function layout (offset ranges) per kernel from nvdisasm
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
layout = {}   # kernel -> list of (offset, fn)
cur = None
pending = None
for line in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m:
        cur = m.group(1); layout[cur] = []; pending = "<kernel body>"; continue
    if cur is None: continue
    m = re.match(r"^\$" + re.escape(cur) + r"\$(\S+):", line)
    if m:
        pending = m.group(1); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", line)
    if m and pending is not None:
        layout[cur].append((int(m.group(1), 16), pending)); pending = None
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kernel = None; hdr = None; base = None
agg = {}
for row in csv.reader(io.StringIO(raw)):
    if not row: continue
    if row[0] == "Kernel Name":
        kernel = row[1]; hdr = None; base = None; continue
    if row[0] == "Address":
        hdr = row; continue
    if hdr is None or kernel is None: continue
    addr = int(row[0], 16)
    if base is None: base = addr
    off = addr - base
    mangled = [k for k in layout if re.sub(r"\(.*", "", kernel).split("::")[-1].split("<")[0] in k]
    lay = layout[mangled[0]] if mangled else []
    fn = "<kernel body>"
    for o, f in lay:
        if o <= off: fn = f
        else: break
    d = agg.setdefault(kernel, collections.OrderedDict()).setdefault(fn, [0, 0])
    d[0] += int(row[hdr.index("Instructions Executed")] or 0)
    d[1] += int(row[hdr.index("# Samples")] or 0)
for k, fns in agg.items():
    ti = sum(v[0] for v in fns.values()); ts = sum(v[1] for v in fns.values())
    print("== %s: %d warp-instructions, %d samples" % (k.split("(")[0], ti, ts))
    for fn, (i, s) in sorted(fns.items(), key=lambda kv: -kv[1][1]):
        name = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip() if fn.startswith("_Z") else fn
        name = re.sub(r"\(.*", "", name)
        print("  %-44s inst %5.1f%%  samples %5.1f%%  (cycles/inst rel %.2f)" % (name[:44], 100.0 * i / ti, 100.0 * s / ts, (s / ts) / (i / ti) if i else 0))
