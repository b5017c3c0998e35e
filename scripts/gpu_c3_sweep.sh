# usage: bash scripts/gpu_c3_sweep.sh <tag> - persist-kernel knobs on the config-3 traversal batch + one ncu --set full of it
TAG=${1:-x}
mkdir -p gpurun_out
for rm in 2 4 8; do for nm in 8 16 24 32; do
  echo "refill=$rm node=$nm $(B200PT_REFILL_MIN=$rm B200PT_NODE_MIN=$nm timeout 300 python bench.py --steps 3 --warmup 2 --rays 4194304 --no-render --no-config4 --no-config5 --no-cpu-baseline --config3-spp 2 2>&1 | python -c 'import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); t=d["config3"]["traversal"]; print(round(t["closest_mrays_per_s"]), round(t["any_mrays_per_s"]), round(d["closest_mrays_per_s"]), round(d["any_mrays_per_s"]))')"
done; done | tee gpurun_out/c3_sweep_$TAG.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:TraversePersist -s 12 -c 2 -o gpurun_out/prof_c3trav_$TAG python bench.py --steps 1 --warmup 1 --rays 4194304 --no-render --no-config4 --no-config5 --no-cpu-baseline --config3-spp 1 > gpurun_out/ncu_c3trav_$TAG.log 2>&1; echo "ncu rc=$?"
