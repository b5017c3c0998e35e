# usage: bash scripts/gpu_n2.sh <tag> <ngpus> - the bench line at N GPUs (torchrun), as the driver launches it
TAG=${1:-x}; N=${2:-2}
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_n${N}_$TAG.json') if l.startswith('{')][-1])
print('n_gpus',d['n_gpus'],'value',d['value'],'e2e',d['e2e']['value'])
for k in ('render','config4','config5'):
    if d.get(k): print(k, d[k]['value'], d[k]['ms_per_frame'])
if d.get('config3'): print('config3', d['config3']['traversal']['value'], d['config3']['render']['value'])
PY
tail -5 gpurun_out/bench_n${N}_$TAG.err
