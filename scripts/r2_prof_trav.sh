# round 2: full-size config-2 parity test, then ncu --set full of the wide and the binary closest-hit kernel
TAG=${1:-b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_traverse_gpu.py -m gpu -x -q -k "full_size" > gpurun_out/pytest_fullsize_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_fullsize_$TAG.log
python scripts/trav_one.py 22 > gpurun_out/trav_one_wide_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:TraversePersist -s 2 -c 2 -f -o gpurun_out/prof_wide_$TAG python scripts/trav_one.py 22 > gpurun_out/ncu_wide_$TAG.log 2>&1
echo "ncu wide rc=$?"
B200PT_TRAVERSAL=binary python scripts/trav_one.py 22 > gpurun_out/trav_one_binary_$TAG.log 2>&1 && \
B200PT_TRAVERSAL=binary ncu --set full --clock-control none --import-source on -k regex:TraversePersist -s 1 -c 2 -f -o gpurun_out/prof_binary_$TAG python scripts/trav_one.py 22 > gpurun_out/ncu_binary_$TAG.log 2>&1
echo "ncu binary rc=$?"
tail -3 gpurun_out/trav_one_wide_$TAG.log gpurun_out/trav_one_binary_$TAG.log
