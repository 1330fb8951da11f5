# Sweep of the SAH-treelet size (Y4GPU_TREELET = leaves per rebuilt subtree, 0 = plain LBVH) on C3 rays, a 2 M-triangle C4 city and the Cornell headline.
for K in ${KS:-0 512 2048 4096}; do
  echo "=== TREELET $K raybench" 
  Y4GPU_TREELET=$K timeout 300 python tools/raybench.py 1000 4096
  echo "=== TREELET $K c4 (2M tris 1080p)"
  Y4GPU_TREELET=$K timeout 300 python tools/c4bench.py 2000000 1920 1080 8 2 1
  echo "=== TREELET $K cornell ptbench"
  Y4GPU_TREELET=$K timeout 300 python bench.py --no-raybatch --no-cpu-baseline --no-e2e --steps 5 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])"
done
