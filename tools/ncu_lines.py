#!/usr/bin/env python
"""Warp instructions executed per source line, from the source page of an .ncu-rep captured with
--import-source on (first captured launch):  python tools/ncu_lines.py prof.ncu-rep [top]"""
import csv
import io
import subprocess
import sys


def main(path, top=50):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source=cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg, seen_kernel = None, None, [], 0
    for r in csv.reader(io.StringIO(out)):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1]
            continue
        if len(r) == 2 and r[0] == "Function Name":
            continue
        if r and r[0] == "Kernel Name":
            seen_kernel += 1
            continue
        if seen_kernel > 1:
            break
        if r and r[0] == "Line No":
            hdr = r
            continue
        if hdr and len(r) == len(hdr) and r[0] != "":       # a source line (SASS rows have no line number)
            try:
                inst = int(r[hdr.index("Instructions Executed")])
            except ValueError:
                continue
            if inst:
                agg.append((inst, (cur or "?").split("/")[-1], r[0], r[1].strip()[:100]))
    tot = sum(a[0] for a in agg)
    print(f"{path}: {tot} warp instructions in the first captured launch")
    by_file = {}
    for a in agg:
        by_file[a[1]] = by_file.get(a[1], 0) + a[0]
    for f, v in sorted(by_file.items(), key=lambda kv: -kv[1]):
        print(f"  {f}: {v} ({100 * v / tot:.1f} %)")
    for a in sorted(agg, reverse=True)[:top]:
        print("%9d %5.1f%%  %s:%s  %s" % (a[0], 100 * a[0] / tot, a[1], a[2], a[3]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 50)
