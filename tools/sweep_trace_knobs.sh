#!/bin/bash
# k_trace warp-scheduling knobs under the path tracer (C2): refill threshold and triangle-postponing divisor.
run() { echo "== $*"; env "$@" timeout 300 python bench.py --no-raybatch --no-cpu-baseline --no-e2e --steps 6 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), {k: round(v,2) for k,v in d['roofline']['class_ms_per_pass'].items()})"; }
run Y4GPU_FETCH_THRESHOLD=20 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=12 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=16 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=24 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=28 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=32 Y4GPU_POSTPONE_DIV=5
run Y4GPU_FETCH_THRESHOLD=24 Y4GPU_POSTPONE_DIV=0
run Y4GPU_FETCH_THRESHOLD=24 Y4GPU_POSTPONE_DIV=3
run Y4GPU_FETCH_THRESHOLD=24 Y4GPU_POSTPONE_DIV=8
run Y4GPU_FETCH_THRESHOLD=20 Y4GPU_MIN_BLOCKS=4
run Y4GPU_FETCH_THRESHOLD=20 Y4GPU_PT_TRACE_BLOCKS=1
run Y4GPU_FETCH_THRESHOLD=20 Y4GPU_PT_TRACE_BLOCKS=3
