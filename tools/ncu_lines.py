"""Aggregate an `ncu --page source --csv --print-source sass,cuda` dump per CUDA source line:
warp instructions executed and stall samples.  usage: ncu_lines.py dump.csv [top]"""
import csv
import sys
from collections import defaultdict


def main(path, top=40):
    fn, fpath, hdr = None, None, None
    agg = defaultdict(lambda: defaultdict(lambda: [0, 0, ""]))
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            fpath = r[1].split("/")[-1]
        elif r[0] == "Function Name":
            fn = r[1]
        elif r[0] == "Line No":
            hdr = r
            i_s, i_e = hdr.index("# Samples"), hdr.index("Instructions Executed")
        elif hdr and r[0].isdigit() and len(r) > i_e:
            try:
                a = agg[fn][(fpath, int(r[0]))]
                a[0] += int(r[i_e] or 0)
                a[1] += int(r[i_s] or 0)
                a[2] = r[1].strip()[:110]
            except ValueError:
                pass
    for f, lines in agg.items():
        ti = sum(v[0] for v in lines.values()) or 1
        ts = sum(v[1] for v in lines.values()) or 1
        print("==", f[:100], "| warp inst", ti, "| samples", ts)
        for (fp, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
            print(f"  {fp:22s}:{ln:4d} inst {v[0] / ti * 100:5.1f}%  samples {v[1] / ts * 100:5.1f}%  {v[2]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
