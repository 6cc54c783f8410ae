#!/usr/bin/env python3
"""DRAM traffic per unit of every kernel in an `ncu --set full` report, stamped with the hash of the CUDA sources of the build
it was captured on (bench.py refuses to quote the file for any other build).
usage: ncu_traffic.py <report.ncu-rep> <units per launch> <capture label> > profiles/ncu_traffic.json"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (source_sha16)

rep, units, label = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, unit_row, data = rows[0], rows[1], rows[2:]
ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
for d in data:
    name = d[ki].split("(")[0].split("<")[0]
    b = float(d[ri].replace(",", "")) * scale.get(unit_row[ri], 1.0) + float(d[wi].replace(",", "")) * scale.get(unit_row[wi], 1.0)
    out[name] = b / units          # the last launch of a kernel in the report wins (same launch shape every time)
print(json.dumps({"source_sha16": bench.source_sha16(), "capture": label, "units_per_launch": units, "bytes_per_unit": out}, indent=1))
