# usage: bash scripts/gpu_c4.sh <tag>  - gpu tests, then the bench line (all sections)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_$TAG.log
timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; python -c "
import json,sys
d=json.load(open('gpurun_out/bench_$TAG.json'))
print(json.dumps(d.get('config4'))[:1500])
print('value',d['value'],'render',d['render']['value'],'c3',d['config3']['traversal']['value'],d['config3']['render']['value'])
"; tail -3 gpurun_out/bench_$TAG.err
