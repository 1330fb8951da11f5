"""Print DRAM bytes and duration per launch from an ncu metrics CSV (dram__bytes_read/write.sum, gpu__time_duration.sum).
usage: ncu_wave.py launches.csv [first_n]"""
import csv
import re
import sys
from collections import defaultdict

SCALE = {'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'Gbyte': 1e9, 'ns': 1, 'us': 1e3, 'ms': 1e6}


def main(path, first=40):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    L = defaultdict(dict)
    for r in rows[1:]:
        k = re.sub(r'\(.*', '', r[4]).replace('void <unnamed>::', '').replace('<unnamed>::', '')
        L[int(r[0])]['k'] = k
        L[int(r[0])][r[12]] = float(r[14].replace(',', '')) * SCALE.get(r[13], 1)
    for i in sorted(L)[:first]:
        l = L[i]
        rd, wr, t = l['dram__bytes_read.sum'], l['dram__bytes_write.sum'], l['gpu__time_duration.sum']
        print(f"{i:4d} {l['k'][:24]:24s} rd {rd * 1e-6:7.1f} MB  wr {wr * 1e-6:7.1f} MB  {t * 1e-3:7.1f} us  {(rd + wr) / t:6.0f} GB/s")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
