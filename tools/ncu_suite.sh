#!/bin/bash
# ncu evidence set of one round (1 GPU): launch lists with DRAM bytes + instruction counts for C2 / C3 / C4 / C5, and --set full
# captures of the dominant kernels.  usage: tools/ncu_suite.sh TAG   (under gpurun); writes gpurun_out/TAG_*.
TAG=$1; OUT=gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum
NCU="ncu --clock-control none --csv"
# C2 path tracer: one warm-up pass + one timed pass
$NCU --metrics $M -c 600 --log-file $OUT/${TAG}_c2_launches.csv python bench.py --steps 1 --warmup 1 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > $OUT/${TAG}_ncu_c2.log 2>&1; echo "c2 rc=$?"
python tools/ncu_traffic.py $OUT/${TAG}_c2_launches.csv $OUT/pt_traffic.json "cornell 1920x1080 spp16 bounces8" > $OUT/${TAG}_c2_classes.txt
# C3 ray batches
$NCU --metrics $M --kernel-name regex:k_trace --log-file $OUT/${TAG}_c3_launches.csv python bench.py --steps 1 --warmup 1 --only-raybatch --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_c3.log 2>&1; echo "c3 rc=$?"
python tools/ncu_raybatch.py $OUT/${TAG}_c3_launches.csv $OUT/rb_classes.json "sphere nu1000 res4096" primary,shadow,diffuse > $OUT/${TAG}_c3_batches.txt
# C5 ray batches
$NCU --metrics $M --kernel-name regex:k_trace --log-file $OUT/${TAG}_c5_launches.csv python bench.py --steps 1 --warmup 1 --workload c5 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_c5.log 2>&1; echo "c5 rc=$?"
python tools/ncu_raybatch.py $OUT/${TAG}_c5_launches.csv $OUT/rb_classes_c5.json "soup 50000000 tris 16777216 rays" incoherent,segment > $OUT/${TAG}_c5_batches.txt
# C4 path tracer: the warm-up pass only is enough for per-launch class averages (one pass = ~1000 launches); the kernel filter
# keeps the 65 bottom-level builds of the two-level scene out of the launch budget
$NCU --metrics $M --kernel-name regex:"k_pt_|k_trace" -c 1100 --log-file $OUT/${TAG}_c4_launches.csv python bench.py --steps 1 --warmup 1 --workload c4 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > $OUT/${TAG}_ncu_c4.log 2>&1; echo "c4 rc=$?"
python tools/ncu_traffic.py $OUT/${TAG}_c4_launches.csv $OUT/pt_traffic_c4.json "city10000000 3840x2160 spp16 bounces8" > $OUT/${TAG}_c4_classes.txt
# --set full: three consecutive launches each (read with tools/ncu_raw.py; the longest one is the one quoted).  The reports stay on
# the box (gpurun brings back at most 64 MiB): only their raw-page summaries are kept.
#   C3 ray batches (primary, shadow, cosine-diffuse), C5 batches (sorted order), C4 closest-hit / shadow launches of early bounces
#   (two-level scene), C2 path kernel of early bounces
REP=/tmp/ncu_rep; mkdir -p $REP
FULL="ncu --set full --import-source on --clock-control none -c 3"
$FULL --kernel-name regex:k_trace --launch-skip 7 -o $REP/full_c3_trace -f python bench.py --steps 1 --warmup 1 --only-raybatch --no-cpu-baseline --no-e2e > /dev/null 2>&1; echo "full c3 rc=$?"
$FULL --kernel-name regex:k_trace --launch-skip 4 -o $REP/full_c5_trace -f python bench.py --steps 1 --warmup 1 --workload c5 --no-cpu-baseline --no-e2e > /dev/null 2>&1; echo "full c5 rc=$?"
$FULL --kernel-name regex:k_trace --launch-skip 40 -o $REP/full_c4_trace -f python bench.py --steps 1 --warmup 1 --workload c4 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > /dev/null 2>&1; echo "full c4 trace rc=$?"
$FULL --kernel-name regex:k_pt_path_compact --launch-skip 0 -o $REP/full_c2_path -f python bench.py --steps 1 --warmup 1 --no-raybatch --no-cpu-baseline --no-e2e --no-profile > /dev/null 2>&1; echo "full c2 path rc=$?"
for r in c3_trace c5_trace c4_trace c2_path; do python tools/ncu_raw.py $REP/full_$r.ncu-rep > $OUT/${TAG}_full_${r}_raw.txt 2>&1; done
# (compute-sanitizer is closed on this pool: no memcheck run can be recorded)
cat $OUT/${TAG}_c2_classes.txt $OUT/${TAG}_c3_batches.txt $OUT/${TAG}_c5_batches.txt $OUT/${TAG}_c4_classes.txt
ls -la $REP; du -sh $OUT
