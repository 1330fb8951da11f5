"""Per-batch DRAM traffic and instruction counts of the ray-batch workload from an ncu launch list:
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum --csv
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline [--workload c5]   (run with Y4_BENCH_ONLY_RAYBATCH=1)
The last len(batches) k_trace launches of the list are the timed step, in batch order.
usage: ncu_raybatch.py launches.csv out.json "<workload>" batch1,batch2,..."""
import csv
import json
import sys
from collections import defaultdict


def main(path, out, workload, batches):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    i_id, i_k, i_m, i_v = (hdr.index(c) for c in ("ID", "Kernel Name", "Metric Name", "Metric Value"))
    L = defaultdict(dict)
    for r in rows[1:]:
        if len(r) > i_v and r[i_id].isdigit():
            L[int(r[i_id])]["kernel"] = r[i_k]
            L[int(r[i_id])][r[i_m]] = float(r[i_v].replace(",", ""))
    tr = [L[i] for i in sorted(L) if "k_trace<" in L[i]["kernel"]]
    last = tr[-len(batches):]
    res = {}
    for b, l in zip(batches, last):
        res[b] = {"kernel": l["kernel"].split("(")[0], "time_us_under_ncu": l["gpu__time_duration.sum"] * 1e-3,
                  "dram_bytes": l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"], "warp_inst": l["smsp__inst_executed.sum"],
                  "thread_inst": l["smsp__thread_inst_executed.sum"], "active_lanes_per_inst": l["smsp__thread_inst_executed.sum"] / l["smsp__inst_executed.sum"]}
        print(b, {k: (round(v, 2) if isinstance(v, float) else v) for k, v in res[b].items()})
    json.dump({"workload": workload, "source": path.split("/")[-1], "note": "one launch per batch, as ncu replays it (alone, cold caches)", "batches": res}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4].split(","))
