#!/usr/bin/env python3
"""Static SASS statistics per device function of one kernel (nvdisasm on the cubin inside a .so).

usage: sass_stats.py <lib.so> <kernel-substring>
Prints, for the kernel body and each out-of-line device function it contains, the instruction count
split into IMAD-class (fma pipe: IMAD*, incl. IMAD.MOV/IMAD.IADD), ALU (IADD3, LOP3, SEL, ISETP, SHF, MOV ...),
local/global memory, and control.  Used to watch code size (i-cache) and the add/mul balance."""
import collections
import os
import re
import subprocess
import sys
import tempfile


def main():
    so, pat = sys.argv[1], sys.argv[2]
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
    cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-c", cubin], capture_output=True, text=True).stdout
    cur_kernel, cur_fn = None, None
    stats = collections.OrderedDict()
    for line in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
        if m:
            cur_kernel = m.group(1)
            cur_fn = "<kernel body>"
            continue
        if cur_kernel is None or pat not in cur_kernel:
            continue
        m = re.match(r"^\$" + re.escape(cur_kernel) + r"\$(\S+):", line)
        if m:
            cur_fn = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(2)
        d = stats.setdefault((cur_kernel, cur_fn), collections.Counter())
        d["total"] += 1
        base = op.split(".")[0]
        if op.startswith("IMAD.WIDE") or op.startswith("IMAD.HI"):
            d["imad_wide"] += 1
        elif op in ("IMAD.MOV.U32", "IMAD.MOV", "IMAD.IADD", "IMAD.SHL", "IMAD.X") or op.startswith("IMAD.MOV") or op.startswith("IMAD.IADD"):
            d["imad_alu"] += 1
        elif base == "IMAD":
            d["imad_lo"] += 1
        elif base in ("LDL", "STL"):
            d["local"] += 1
        elif base in ("LDG", "STG", "LD", "ST", "LDS", "STS", "LDC", "LDCU"):
            d["mem"] += 1
        elif base in ("BRA", "CALL", "RET", "EXIT", "BSSY", "BSYNC", "WARPSYNC", "NOP", "BREAK", "JMP", "BRX"):
            d["ctrl"] += 1
        else:
            d["alu"] += 1
    print("%-64s %7s %7s %7s %7s %7s %7s %7s %7s" % ("function", "total", "wide", "imadlo", "imadalu", "alu", "local", "mem", "ctrl"))
    tot = collections.Counter()
    for (kname, fn), d in stats.items():
        short = fn
        try:
            short = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip() or fn
        except Exception:
            pass
        short = re.sub(r"\(.*", "", short)[:64]
        print("%-64s %7d %7d %7d %7d %7d %7d %7d %7d" % (short, d["total"], d["imad_wide"], d["imad_lo"], d["imad_alu"], d["alu"], d["local"], d["mem"], d["ctrl"]))
        tot.update(d)
    print("%-64s %7d %7d %7d %7d %7d %7d %7d %7d  (%.0f KB of SASS)" % ("TOTAL", tot["total"], tot["imad_wide"], tot["imad_lo"], tot["imad_alu"], tot["alu"], tot["local"],
                                                                    tot["mem"], tot["ctrl"], tot["total"] * 16 / 1024.0))


if __name__ == "__main__":
    main()
