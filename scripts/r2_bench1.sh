# usage: bash scripts/r2_bench1.sh <tag> [extra bench args] : full bench line (wide walk) + the binary walk's traversal numbers for comparison
TAG=${1:-x}; shift
mkdir -p gpurun_out
( time timeout 2400 python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err ) 2> gpurun_out/bench_$TAG.time; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.time
tail -5 gpurun_out/bench_$TAG.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_$TAG.json'))
print('value', round(d['value']), 'closest', round(d['closest_mrays_per_s']), 'any', round(d['any_mrays_per_s']), 'e2e', round(d['e2e']['value']))
print('parity', d.get('parity'))
print('roofline', {k: (round(v, 4) if isinstance(v, float) else v) for k, v in d['roofline'].items()})
for k in ('render', 'config3', 'config4', 'config5'):
    c = d.get(k)
    if not c: continue
    r = c.get('render', c)
    print(k, 'value', round(r['value'], 1), 'e2e', round(r['e2e']['value'], 1), 'launches', r.get('launches_per_frame'), 'cpu', (r.get('cpu_baseline') or {}).get('value'))
    print('   parity', (r.get('cpu_baseline') or {}).get('parity'))
    if 'traversal' in c: print('   traversal', {kk: (round(v, 3) if isinstance(v, float) else v) for kk, v in c['traversal'].items()})
PY
B200PT_TRAVERSAL=binary timeout 900 python bench.py --steps 5 --warmup 3 --no-render --no-config4 --no-config5 --no-cpu-baseline --config3-spp 2 > gpurun_out/bench_binary_$TAG.json 2> gpurun_out/bench_binary_$TAG.err; echo "bench binary rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_binary_$TAG.json'))
print('binary value', round(d['value']), 'closest', round(d['closest_mrays_per_s']), 'any', round(d['any_mrays_per_s']))
print('   c3 traversal', {kk: (round(v, 3) if isinstance(v, float) else v) for kk, v in d['config3']['traversal'].items()})
PY
