#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples for one kernel of an .ncu-rep.

    python tools/ncu_lines.py <report.ncu-rep> <kernel-substring> [launch-index] [top]

Joins `ncu --page source --csv` (SASS rows with addresses) with `nvdisasm -g` line markers of the matching
cubin extracted from the in-tree libexo_env.so (built with -lineinfo).
"""
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_rows(rep, kernel, launch):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kernel],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
    rows = rows[starts[launch]:starts[launch + 1]]
    name = rows[0][1]
    h = rows[1]
    cols = {k: h.index(k) for k in ("Address", "Source", "Instructions Executed", "Warp Stall Sampling (All Samples)")}
    data = []
    for r in rows[2:]:
        if len(r) <= max(cols.values()) or not r[cols["Address"]].startswith("0x"):
            continue
        data.append((int(r[cols["Address"]], 16), r[cols["Source"]], float(r[cols["Instructions Executed"]] or 0),
                     float(r[cols["Warp Stall Sampling (All Samples)"]] or 0)))
    base = min(d[0] for d in data)
    return name, [(a - base, s, n, st) for a, s, n, st in data]


def line_map(mangled_sub):
    tmp = tempfile.mkdtemp()
    so = glob.glob(os.path.join(REPO, "*_b200", "libexo_env.so"))[0]
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
    for cubin in glob.glob(os.path.join(tmp, "*sm_100a.cubin")):
        txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
        m = {}
        cur, infn = None, False
        for ln in txt.splitlines():
            if ln.startswith(".text."):
                infn = mangled_sub in ln
                continue
            if not infn:
                continue
            mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if mm:
                cur = (os.path.basename(mm.group(1)), int(mm.group(2)))
                continue
            mm = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
            if mm:
                m[int(mm.group(1), 16)] = cur
        if m:
            return m
    return {}


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    name, rows = sass_rows(rep, kernel, launch)
    # mangled-name fragment: the demangled kernel name up to '<' is enough to find the function, template args disambiguate
    frag = sys.argv[5] if len(sys.argv) > 5 else re.sub(r"\W", "", name.split("(")[0].split("::")[-1].split("<")[0])
    lm = line_map(frag)
    agg = {}
    tot_i = sum(r[2] for r in rows) or 1
    tot_s = sum(r[3] for r in rows) or 1
    for off, sass, n, st in rows:
        key = lm.get(off, ("?", 0))
        a = agg.setdefault(key, [0.0, 0.0])
        a[0] += n
        a[1] += st
    src = {}
    print("%s\n total warp-instructions %.0f, stall samples %.0f" % (name[:100], tot_i, tot_s))
    for (f, l), (n, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        if f not in src:
            p = glob.glob(os.path.join(REPO, "*_b200", "csrc", f))
            src[f] = open(p[0]).read().splitlines() if p else []
        text = src[f][l - 1].strip()[:100] if 0 < l <= len(src[f]) else ""
        print("%5.1f%% inst %5.1f%% stall  %s:%d  %s" % (100 * n / tot_i, 100 * st / tot_s, f, l, text))


if __name__ == "__main__":
    main()
