#!/bin/bash
# pixel-samples per wavefront (2 wavefronts in flight) under the C2 bench
run() { echo "== $*"; env "$@" timeout 300 python bench.py --no-raybatch --no-cpu-baseline --no-e2e --steps 6 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), {k: round(v,2) for k,v in d['roofline']['class_ms_per_pass'].items()})"; }
run Y4_BENCH_WAVE=1048576
run Y4_BENCH_WAVE=2097152
run Y4_BENCH_WAVE=4194304
run Y4_BENCH_WAVE=8388608
run Y4_BENCH_WAVE=16777216
run Y4_BENCH_WAVE=8388608 Y4GPU_PT_LANES=3
run Y4_BENCH_WAVE=2097152 Y4GPU_PT_LANES=3
run Y4_BENCH_WAVE=2097152 Y4GPU_PT_LANES=4
