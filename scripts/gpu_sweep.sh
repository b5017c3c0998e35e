mkdir -p gpurun_out
for rm in 4 8 16; do for nm in 4 10 16 24; do
  echo "refill=$rm node=$nm $(B200PT_REFILL_MIN=$rm B200PT_NODE_MIN=$nm timeout 300 python bench.py --steps 3 --warmup 2 --rays 8388608 --no-cpu-baseline 2>&1 | python -c 'import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["closest_mrays_per_s"]), round(d["any_mrays_per_s"]))')"
done; done | tee gpurun_out/sweep.txt
