# usage: bash scripts/r2_trav6.sh <tag> : traversal tests with both walks, then config-2 and full-size config-3 timings of both
TAG=${1:-x}
mkdir -p gpurun_out
for mode in binary wide; do
B200PT_TRAVERSAL=$mode timeout 900 python -m pytest tests/test_traverse_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_trav_${mode}_$TAG.log 2>&1; echo "pytest $mode rc=$?"; tail -3 gpurun_out/pytest_trav_${mode}_$TAG.log
B200PT_TRAVERSAL=$mode timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
B200PT_TRAVERSAL=$mode timeout 600 python scripts/trav_c3.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
done
