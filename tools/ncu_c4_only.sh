#!/bin/bash
# the C4 part of tools/ncu_suite.sh on its own.  usage: tools/ncu_c4_only.sh TAG
TAG=$1; OUT=gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
ncu --clock-control none --csv --metrics $M --kernel-name regex:"k_pt_|k_trace" -c 1100 --log-file $OUT/${TAG}_c4_launches.csv python bench.py --steps 1 --warmup 1 --workload c4 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > $OUT/${TAG}_ncu_c4.log 2>&1; echo "c4 rc=$?"
python tools/ncu_traffic.py $OUT/${TAG}_c4_launches.csv $OUT/pt_traffic_c4.json "city10000000 3840x2160 spp16 bounces8" > $OUT/${TAG}_c4_classes.txt
REP=/tmp/ncu_rep; mkdir -p $REP
ncu --set full --import-source on --clock-control none -c 3 --kernel-name regex:k_trace --launch-skip 40 -o $REP/full_c4_trace -f python bench.py --steps 1 --warmup 1 --workload c4 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > /dev/null 2>&1; echo "full c4 trace rc=$?"
python tools/ncu_raw.py $REP/full_c4_trace.ncu-rep > $OUT/${TAG}_full_c4_trace_raw.txt 2>&1
cat $OUT/${TAG}_c4_classes.txt
