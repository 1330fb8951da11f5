"""Summarise an `ncu --page source --csv` dump: per-kernel instruction mix and stall-sample share."""
import csv
import sys
from collections import Counter


def main(path, top=30):
    secs, cur, hdr = [], None, None
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            secs.append(cur)
        elif r[0] == "Address":
            hdr = r
        elif r[0].startswith("0x") and cur is not None:
            cur["rows"].append(r)
    ia, ie, it, isamp = (hdr.index(k) for k in ("Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
    for s in secs:
        body = s["rows"]
        tot = sum(int(r[ie]) for r in body)
        tt = sum(int(r[it]) for r in body)
        print("==", s["name"][:90], "| SASS lines", len(body), "| warp inst", tot, "| thread inst", tt, "| avg active %.2f" % (tt / max(1, tot)))
        c, cs = Counter(), Counter()
        for r in body:
            toks = r[ia].split()
            op = toks[1] if toks[0].startswith("@") else toks[0]
            op = op.split(".")[0]
            c[op] += int(r[ie])
            cs[op] += int(r[isamp])
        ts = max(1, sum(cs.values()))
        for op, v in c.most_common(top):
            print(f"  {op:10s} {v / tot * 100:6.2f}%  stall samples {cs[op] / ts * 100:5.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
