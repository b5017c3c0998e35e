# usage: bash scripts/gpu_c3.sh <tag> - bench line with only the config-3 section beside the headline
TAG=${1:-x}
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 5 --warmup 3 --no-render --no-config4 --no-config5 --no-cpu-baseline > gpurun_out/bench_c3_$TAG.json 2> gpurun_out/bench_c3_$TAG.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_c3_$TAG.json'))
print(json.dumps(d['config3']['traversal']))
print('value', d['value'], 'c3 render', d['config3']['render']['value'])
"; tail -3 gpurun_out/bench_c3_$TAG.err
