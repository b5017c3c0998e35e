# usage: bash scripts/r2_wide3.sh <tag> : traversal gpu tests, then a scheduling-knob sweep of both kernels on the config-2 scene
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_traverse_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_trav_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_trav_$TAG.log
for mode in wide binary; do
for nm in 0 8 12 16 20 24; do
B200PT_TRAVERSAL=$mode B200PT_NODE_MIN=$nm timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
done; done
for rm in 1 8 16; do
B200PT_TRAVERSAL=wide B200PT_NODE_MIN=16 B200PT_REFILL_MIN=$rm timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
done
