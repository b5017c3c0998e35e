# usage: bash scripts/r2_wide1.sh <tag> - round 2: traversal gpu tests + wide vs binary kernel on config 2 and config 3 batches
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_traverse_gpu.py -m gpu -x -q > gpurun_out/pytest_trav_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_trav_$TAG.log
for mode in wide binary; do
B200PT_TRAVERSAL=$mode timeout 600 python bench.py --steps 5 --warmup 3 --no-render --no-config4 --no-config5 --no-cpu-baseline --config3-spp 2 > gpurun_out/bench_${mode}_$TAG.json 2> gpurun_out/bench_${mode}_$TAG.err; echo "bench $mode rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_${mode}_$TAG.json'))
print('$mode', 'value', round(d['value']), 'closest', round(d['closest_mrays_per_s']), 'any', round(d['any_mrays_per_s']), 'e2e', round(d['e2e']['value']))
print(' roofline', json.dumps(d['roofline'])[:600])
t=d['config3']['traversal']; print(' c3', json.dumps(t)[:700])
PY
tail -2 gpurun_out/bench_${mode}_$TAG.err
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest all rc=$?"; tail -8 gpurun_out/pytest_gpu_$TAG.log
