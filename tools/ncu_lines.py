#!/usr/bin/env python3
"""Per-source-line instruction counts from an ncu report: joins `ncu --page source --csv` (SASS rows with
'Instructions Executed' / 'Thread Instructions Executed') with nvdisasm's line info of the same cubin.
usage: ncu_lines.py report.ncu-rep object.o kernel-substring [top=40]"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


INCLUSIVE = False
INNER = False       # --inner (with --frame): the INNERMOST frame inside FILE instead of the outermost
FRAME = None        # --frame FILE: attribute an instruction to its outermost inlining frame inside FILE (implies -gi)


def line_map(obj, kernel):
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, capture_output=True)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-gi" if (INCLUSIVE or FRAME) else "-g", "-c", os.path.join(d, cubin)], capture_output=True, text=True).stdout
    out, cur, on, fresh = {}, ("?", 0), False, True
    chain, after_instr = [], True
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            on = kernel in ln
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            # with -gi a chain of "inner inlined at outer" lines precedes the instruction: the LAST line is the
            # outermost frame (the line of the kernel body that is responsible for the instruction)
            if after_instr:
                chain, after_instr = [], False
            chain.append((os.path.basename(m.group(1)), int(m.group(2))))
            cur = chain[-1]
            if FRAME:
                inside = [c for c in chain if c[0] == FRAME]
                cur = (inside[0] if INNER else inside[-1]) if inside else chain[-1]
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur, m.group(2).strip())
            after_instr = True
    return out


def main():
    global INCLUSIVE
    if "--inclusive" in sys.argv:          # attribute inlined code to the outermost call site
        INCLUSIVE = True
        sys.argv.remove("--inclusive")
    global FRAME, INNER
    if "--inner" in sys.argv:
        INNER = True
        sys.argv.remove("--inner")
    if "--frame" in sys.argv:              # e.g. --frame rt_trace.cuh: per line of the walk instead of per line of the kernel body
        i = sys.argv.index("--frame")
        FRAME = sys.argv[i + 1]
        del sys.argv[i:i + 2]
    rep, obj, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    lm = line_map(obj, kernel)
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    H = rows[hdr]
    ia, ie, it, isamp = H.index("Address"), H.index("Instructions Executed"), H.index("Thread Instructions Executed"), H.index("# Samples")
    base = None
    per = defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    for r in rows[hdr + 1:]:
        if len(r) <= it or not r[ia].startswith("0x"):
            continue
        addr = int(r[ia], 16)
        base = addr if base is None else base
        key, _ = lm.get(addr - base, (("?", 0), ""))
        v = (int(r[ie] or 0), int(r[it] or 0), int(r[isamp] or 0))
        for k in range(3):
            per[key][k] += v[k]
            tot[k] += v[k]
    print("total warp-instr %d, thread-instr %d, avg lanes %.2f, samples %d" % (tot[0], tot[1], tot[1] / max(tot[0], 1), tot[2]))
    unit = float(os.environ.get("NCU_LINES_UNITS", "0"))     # e.g. rays in the launch: prints warp-instr per unit
    byfile = defaultdict(lambda: [0, 0, 0])
    for (f, l), v in per.items():
        for k in range(3):
            byfile[f][k] += v[k]
    for f, v in sorted(byfile.items(), key=lambda x: -x[1][0]):
        print("  %-18s %5.1f%% of warp-instr, lanes %.2f, %5.1f%% of samples" % (f, 100.0 * v[0] / tot[0], v[1] / max(v[0], 1), 100.0 * v[2] / max(tot[2], 1)))
    print("top lines:")
    for (f, l), v in sorted(per.items(), key=lambda x: -x[1][0])[:top]:
        print("  %-16s:%-4d %5.2f%% warp-instr  lanes %5.2f  samples %5.2f%%%s" % (f, l, 100.0 * v[0] / tot[0], v[1] / max(v[0], 1), 100.0 * v[2] / max(tot[2], 1),
                                                                                 "  %.2f/unit" % (v[0] / unit) if unit else ""))


if __name__ == "__main__":
    main()
