#!/bin/bash
# remaining path-tracer knobs at the 8 Mi wavefront (C2)
run() { echo "== $*"; env "$@" timeout 300 python bench.py --no-raybatch --no-cpu-baseline --no-e2e --steps 6 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3), {k: round(v,2) for k,v in d['roofline']['class_ms_per_pass'].items()})"; }
run Y4GPU_PT_MIN_BLOCKS=3
run Y4GPU_PT_MIN_BLOCKS=2
run Y4GPU_PT_MIN_BLOCKS=4
run Y4GPU_PT_COMPACT_MINB=4
run Y4GPU_PT_COMPACT_MINB=6
run Y4GPU_PT_FETCH_THRESHOLD=12
run Y4GPU_PT_FETCH_THRESHOLD=20
run Y4GPU_PT_TRACE_BLOCKS=3
run Y4GPU_PT_TRACE_BLOCKS=1
run Y4GPU_PT_LANES=1
