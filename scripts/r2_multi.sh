# usage: bash scripts/r2_multi.sh <tag> <ngpus> : the bench line at N ranks the way the driver launches it
TAG=${1:-x}; N=${2:-2}
mkdir -p gpurun_out
( time timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err ) 2> gpurun_out/bench_n${N}_$TAG.time; echo "bench rc=$?"; tail -3 gpurun_out/bench_n${N}_$TAG.time
tail -5 gpurun_out/bench_n${N}_$TAG.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_n${N}_$TAG.json') if l.startswith('{')][-1])
print('N', d['n_gpus'], 'value', round(d['value']), 'closest', round(d['closest_mrays_per_s']), 'any', round(d['any_mrays_per_s']), 'e2e', round(d['e2e']['value']))
print('parity', d.get('parity', {}).get('mismatches'), 'roofline frac', round(d['roofline']['frac'], 3), 'frac_dram', d['roofline'].get('frac_dram'), 'issue_x_lanes', d['roofline'].get('issue_x_lanes'))
for k in ('render', 'config3', 'config4', 'config5'):
    c = d.get(k)
    if not c: continue
    r = c.get('render', c)
    print(k, 'value', round(r['value'], 1), 'e2e', round(r['e2e']['value'], 1), 'launches', r.get('launches_per_frame'), 'parity', ((r.get('cpu_baseline') or {}).get('parity') or {}).get('pass'))
    if 'traversal' in c: print('   traversal closest', round(c['traversal']['closest_mrays_per_s']), 'any', round(c['traversal']['any_mrays_per_s']), 'parity', c['traversal'].get('parity', {}).get('mismatches'))
PY
