mkdir -p gpurun_out
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1_b.json 2> gpurun_out/bench_r1_b.err; echo "bench rc=$?"; cat gpurun_out/bench_r1_b.json; tail -3 gpurun_out/bench_r1_b.err
timeout 600 python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:TraverseBVH2Kernel -s 6 -c 2 -o gpurun_out/prof_r1_bvh2 python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out
