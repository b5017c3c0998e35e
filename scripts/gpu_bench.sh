# usage: bash scripts/gpu_bench.sh <tag> [extra bench args]
TAG=${1:-x}; shift
mkdir -p gpurun_out
timeout 1500 python bench.py --steps 5 --warmup 3 "$@" > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json; tail -5 gpurun_out/bench_$TAG.err
