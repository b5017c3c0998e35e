# usage: bash scripts/r2_render_only.sh <tag> : the bench's render sections only (configs 1, 4, 5; no config 3, no CPU legs) - quick A/B of the shading kernels
TAG=${1:-x}
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 5 --warmup 3 --no-config3 --no-cpu-baseline --rays 4194304 > gpurun_out/bench_ro_$TAG.json 2> gpurun_out/bench_ro_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_ro_$TAG.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_ro_$TAG.json'))
for k in ('render', 'config4', 'config5'):
    c = d.get(k)
    if not c: continue
    r = c.get('render', c)
    print(k, 'value', round(r['value'], 1), 'e2e', round(r['e2e']['value'], 1), 'launches', r.get('launches_per_frame'), 'parity', r.get('parity'))
PY
