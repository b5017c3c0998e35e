# usage: bash scripts/r2_wide5.sh <tag> : the wide walk in the loop shape: tests, then timing against the phase-per-round shape and the binary kernel
TAG=${1:-x}
mkdir -p gpurun_out
B200PT_WIDE_LOOP=1 timeout 900 python -m pytest tests/test_traverse_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_trav_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_trav_$TAG.log
B200PT_TRAVERSAL=binary timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
for nm in 8 12 16 20; do
echo -n "loop " | tee -a gpurun_out/sweep_$TAG.txt
B200PT_WIDE_LOOP=1 B200PT_NODE_MIN=$nm timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
done
B200PT_WIDE_LOOP=1 python scripts/trav_one.py 22 > gpurun_out/trav_one_wideloop_$TAG.log 2>&1 && \
B200PT_WIDE_LOOP=1 ncu --set full --clock-control none --import-source on -k regex:TraverseWide -s 2 -c 1 -f -o gpurun_out/prof_wideloop_$TAG python scripts/trav_one.py 22 > gpurun_out/ncu_wideloop_$TAG.log 2>&1
echo "ncu rc=$?"
