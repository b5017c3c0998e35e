# usage: bash scripts/gpu_cycle.sh <tag> [kernel-regex] [skip]   - tests, bench, launch list, one ncu --set full capture
TAG=${1:-x}; KRE=${2:-Traverse}; SKIP=${3:-6}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json; tail -3 gpurun_out/bench_$TAG.err
timeout 600 python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c 2 -o gpurun_out/prof_$TAG python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full_$TAG.log | cut -c1-200
