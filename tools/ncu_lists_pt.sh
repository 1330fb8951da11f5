#!/bin/bash
# launch lists of the two path-tracer workloads only (C2, C4): tools/ncu_suite.sh without the ray batches and the full captures.
# The per-launch averages in profiles/pt_traffic*.json are tied to the wavefront size of the build that was captured.
TAG=$1; OUT=gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
NCU="ncu --clock-control none --csv"
$NCU --metrics $M -c 600 --log-file $OUT/${TAG}_c2_launches.csv python bench.py --steps 1 --warmup 1 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > $OUT/${TAG}_ncu_c2.log 2>&1; echo "c2 rc=$?"
python tools/ncu_traffic.py $OUT/${TAG}_c2_launches.csv $OUT/pt_traffic.json "cornell 1920x1080 spp16 bounces8" > $OUT/${TAG}_c2_classes.txt
$NCU --metrics $M --kernel-name regex:"k_pt_|k_trace" -c 700 --log-file $OUT/${TAG}_c4_launches.csv python bench.py --steps 1 --warmup 1 --workload c4 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > $OUT/${TAG}_ncu_c4.log 2>&1; echo "c4 rc=$?"
python tools/ncu_traffic.py $OUT/${TAG}_c4_launches.csv $OUT/pt_traffic_c4.json "city10000000 3840x2160 spp16 bounces8" > $OUT/${TAG}_c4_classes.txt
cat $OUT/${TAG}_c2_classes.txt $OUT/${TAG}_c4_classes.txt
