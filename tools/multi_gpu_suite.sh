#!/bin/bash
# N-GPU measurement set of one box: NCCL film test (N = 2 only) + bench lines for C2 / C4 / C5, weak and strong scaling.
# usage: tools/multi_gpu_suite.sh N tag     (run under `gpurun --gpus N`)
N=$1; TAG=$2; OUT=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
if [ "$N" = "2" ]; then python -m pytest tests/test_multi_nccl.py -x -q -m gpu > $OUT/${TAG}_nccl_test.log 2>&1; echo "nccl test rc=$?"; tail -3 $OUT/${TAG}_nccl_test.log; fi
$TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_c2_weak_n$N.json 2> $OUT/${TAG}_n$N.err; echo "c2 weak rc=$?"
$TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --scaling strong > $OUT/${TAG}_c2_strong_n$N.json 2>> $OUT/${TAG}_n$N.err; echo "c2 strong rc=$?"
$TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-raybatch --workload c4 > $OUT/${TAG}_c4_weak_n$N.json 2>> $OUT/${TAG}_n$N.err; echo "c4 weak rc=$?"
$TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-raybatch --workload c4 --scaling strong > $OUT/${TAG}_c4_strong_n$N.json 2>> $OUT/${TAG}_n$N.err; echo "c4 strong rc=$?"
$TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --workload c5 > $OUT/${TAG}_c5_weak_n$N.json 2>> $OUT/${TAG}_n$N.err; echo "c5 weak rc=$?"
$TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --workload c5 --scaling strong > $OUT/${TAG}_c5_strong_n$N.json 2>> $OUT/${TAG}_n$N.err; echo "c5 strong rc=$?"
tail -5 $OUT/${TAG}_n$N.err
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/${TAG}_c*_n$N.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        rb = d.get("raybatch") or {}
        print(f.split("/")[-1], "value %.1f %s" % (d["value"], d["unit"]), "ms/step %.2f" % d["ms_per_step"], "e2e %.1f" % d.get("e2e", {}).get("value", 0), ("raybatch %.1f e2e %.1f" % (rb.get("value", 0), rb.get("e2e", {}).get("value", 0))) if rb else "")
    except Exception as e:
        print(f, "unreadable", e)
PY
