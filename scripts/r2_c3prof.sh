# usage: bash scripts/r2_c3prof.sh <tag> : GPU render tests (tight noise gates), then ncu of the binary walk on the full-size config-3 batches
TAG=${1:-x}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_render_gpu.py -m gpu -x -q > gpurun_out/pytest_render_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_render_$TAG.log
B200PT_TRAVERSAL=binary COUNTERS=1 python scripts/trav_c3.py 22 > gpurun_out/trav_c3_binary_$TAG.log 2>&1 && \
B200PT_TRAVERSAL=binary ncu --set full --clock-control none --import-source on -k regex:TraverseBinary -s 2 -c 2 -f -o gpurun_out/prof_c3_binary_$TAG python scripts/trav_c3.py 22 > gpurun_out/ncu_c3_binary_$TAG.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/trav_c3_binary_$TAG.log
