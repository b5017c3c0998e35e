mkdir -p gpurun_out
set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1_a.json 2> gpurun_out/bench_r1_a.err; echo "bench rc=$?"; cat gpurun_out/bench_r1_a.json; tail -3 gpurun_out/bench_r1_a.err
timeout 600 python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:Traverse --csv --log-file gpurun_out/launches_r1_a.csv python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:TraverseBVH2KernelILb0ELb0 -s 1 -c 1 -o gpurun_out/prof_r1_bvh2_closest python bench.py --steps 1 --warmup 1 --rays 4194304 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out
