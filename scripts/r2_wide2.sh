# usage: bash scripts/r2_wide2.sh <tag> : traversal gpu tests, wide vs binary bench (config 2 + config 3 batches), ncu of the wide kernel
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_traverse_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_trav_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_trav_$TAG.log
for mode in wide binary; do
B200PT_TRAVERSAL=$mode timeout 600 python bench.py --steps 5 --warmup 3 --no-render --no-config4 --no-config5 --no-cpu-baseline --config3-spp 2 > gpurun_out/bench_${mode}_$TAG.json 2> gpurun_out/bench_${mode}_$TAG.err; echo "bench $mode rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_${mode}_$TAG.json'))
print('$mode', 'value', round(d['value']), 'closest', round(d['closest_mrays_per_s']), 'any', round(d['any_mrays_per_s']), 'e2e', round(d['e2e']['value']))
t=d['config3']['traversal']; print(' c3 closest', round(t['closest_mrays_per_s']), 'any', round(t['any_mrays_per_s']))
PY
tail -2 gpurun_out/bench_${mode}_$TAG.err
done
python scripts/trav_one.py 22 > gpurun_out/trav_one_wide_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:TraversePersist -s 4 -c 1 -f -o gpurun_out/prof_wide_$TAG python scripts/trav_one.py 22 > gpurun_out/ncu_wide_$TAG.log 2>&1
echo "ncu wide rc=$?"
tail -n 2 gpurun_out/trav_one_wide_$TAG.log
