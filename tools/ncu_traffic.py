"""Sum an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum --csv` launch list per
kernel class of a path-tracer pass (the classes bench.py's roofline uses) and write the per-launch DRAM traffic as JSON.
usage: ncu_traffic.py launches.csv out.json "<workload the capture ran>"
"""
import csv
import json
import re
import sys
from collections import defaultdict

CLASSES = [  # (regex on the demangled kernel name, class name as bench.py prints it)
    (r"k_pt_gen\(", "camera rays"),
    (r"k_trace<0", "trace_closest"),
    (r"k_trace<1", "trace_shadow"),
    (r"k_trace2<0", "trace_closest"),
    (r"k_trace2<1", "trace_shadow"),
    (r"k_trace_tshadow", "trace_shadow"),
    (r"k_pt_stage<1", "extra light-sample stages"),
    (r"k_pt_path_compact<", "path stages (camera hit, start, per-bounce)"),
    (r"k_pt_stage<", "path stages (camera hit, start, per-bounce)"),
    (r"k_pt_splat\(", "splat"),
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}


def main(path, out, workload):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    i_id, i_k, i_m, i_u, i_v = (hdr.index(c) for c in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value"))
    launches = defaultdict(dict)
    for r in rows[1:]:
        if len(r) <= i_v or not r[i_id].isdigit():
            continue
        launches[int(r[i_id])]["kernel"] = r[i_k]
        launches[int(r[i_id])][r[i_m]] = float(r[i_v].replace(",", "")) * SCALE.get(r[i_u], 1.0)
    agg = defaultdict(lambda: {"launches": 0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0, "time_ns": 0.0, "warp_inst": 0.0, "thread_inst": 0.0})
    for l in launches.values():
        for rx, name in CLASSES:
            if re.search(rx, l["kernel"]):
                a = agg[name]
                a["launches"] += 1
                a["dram_read_bytes"] += l.get("dram__bytes_read.sum", 0.0)
                a["dram_write_bytes"] += l.get("dram__bytes_write.sum", 0.0)
                a["time_ns"] += l.get("gpu__time_duration.sum", 0.0)
                a["warp_inst"] += l.get("smsp__inst_executed.sum", 0.0)
                a["thread_inst"] += l.get("smsp__thread_inst_executed.sum", 0.0)
                break
    for a in agg.values():
        a["dram_bytes_per_launch"] = (a["dram_read_bytes"] + a["dram_write_bytes"]) / max(1, a["launches"])
        a["dram_GBps_under_ncu"] = (a["dram_read_bytes"] + a["dram_write_bytes"]) / max(1.0, a["time_ns"])
        a["warp_inst_per_launch"] = a["warp_inst"] / max(1, a["launches"])
        a["active_lanes_per_inst"] = a["thread_inst"] / a["warp_inst"] if a["warp_inst"] else None
    json.dump({"workload": workload, "source": path.split("/")[-1],
               "note": "ncu replays every launch alone and cold: bytes are per launch as measured, times are not bench values",
               "classes": dict(agg)}, open(out, "w"), indent=1)
    for k, a in agg.items():
        print(f"{k:48s} launches {a['launches']:4d}  dram {(a['dram_read_bytes'] + a['dram_write_bytes']) * 1e-9:8.3f} GB  "
              f"per launch {a['dram_bytes_per_launch'] * 1e-6:8.2f} MB  time {a['time_ns'] * 1e-6:8.2f} ms")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
