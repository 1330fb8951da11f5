# Knob sweep of the path tracer headline (Cornell 1080p, 16 spp passes): lanes, stage register budget, wave size, traversal CTAs per SM.
run() { echo "== $*"; env "$@" timeout 300 python bench.py --no-raybatch --no-cpu-baseline --no-e2e --steps 6 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],3))"; }
run Y4GPU_PT_LANES=2
run Y4GPU_PT_LANES=3
run Y4GPU_PT_LANES=4
run Y4GPU_PT_LANES=2 Y4GPU_PT_MIN_BLOCKS=4
run Y4GPU_PT_LANES=2 Y4GPU_PT_TRACE_BLOCKS=3
run Y4GPU_PT_LANES=3 Y4GPU_PT_TRACE_BLOCKS=3
run Y4GPU_PT_LANES=2 Y4GPU_PT_TRACE_BLOCKS=1
run Y4GPU_PT_LANES=3 Y4GPU_PT_MIN_BLOCKS=4
echo "== raybench leaf_max 1 2 3 (SAH tree)"
for L in 1 2 3; do timeout 300 python tools/raybench.py 1000 4096 $L | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('n_nodes','build_ms','batch','mrays_s','nodes_per_ray','tris_per_ray')})"; done
echo "== treelet defer sweep (C3 build ms)"
for D in 0 32 64 128 256; do Y4GPU_TREELET_DEFER=$D timeout 300 python tools/raybench.py 1000 1024 | head -1; done
