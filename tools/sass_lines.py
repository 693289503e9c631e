#!/usr/bin/env python3
"""Join ncu's per-SASS-instruction counters with nvdisasm's line info -> per-source-line totals.

  ncu -i X.ncu-rep --page source --csv --kernel-name regex:K --launch-count 1 > sass.csv
  cuobjdump -xelf all lib.so ; nvdisasm -g -c file.cubin > file.sass
  python tools/sass_lines.py sass.csv file.sass <mangled kernel name substring> [top]
"""
import csv
import re
import sys
import collections


def main():
    sass_csv, dis, kname = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(sass_csv)))
    h = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
    hdr = rows[h]
    si, ni, ii = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    prof = []
    for r in rows[h + 1:]:
        if len(r) <= ii:
            break
        prof.append((r[si].strip(), int(r[ii] or 0), int(r[ni] or 0)))
    # nvdisasm: instructions of the kernel in order, with the last seen //## File line
    lines = open(dis).read().split("\n")
    start = next(i for i, l in enumerate(lines) if ".text." in l and kname in l and ".section" in l)
    cur, ins = ("?", 0), []
    for l in lines[start + 1:]:
        if l.startswith("\t.section") or l.startswith("//-----"):
            if ins:
                break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
        if m:
            ins.append((m.group(1).strip(), cur))
    if len(ins) != len(prof):
        print(f"warning: {len(ins)} disassembled vs {len(prof)} profiled instructions (different build?)", file=sys.stderr)
    agg = collections.defaultdict(lambda: [0, 0])
    for (_, loc), (_, n, s) in zip(ins, prof):
        agg[loc][0] += n
        agg[loc][1] += s
    ti, ts = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
    src = {}
    print(f"# {kname}: {ti} warp instructions, {ts} stall samples; per source line (inlined callees keep their own line)")
    for loc, (n, s) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        f, ln = loc
        if f not in src:
            try:
                import glob
                p = glob.glob(f"**/{f}", recursive=True)
                src[f] = open(p[0]).read().split("\n") if p else []
            except OSError:
                src[f] = []
        text = src[f][ln - 1].strip()[:90] if 0 < ln <= len(src[f]) else ""
        print(f"{100 * n / ti:5.1f}% inst {100 * s / max(ts, 1):5.1f}% stall  {f}:{ln:<4d} {text}")


if __name__ == "__main__":
    main()
