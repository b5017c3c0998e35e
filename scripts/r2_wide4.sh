# usage: bash scripts/r2_wide4.sh <tag> : traversal tests; wide kernel at 6 and 5 resident blocks per SM x nodeMin; binary reference point
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_traverse_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_trav_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_trav_$TAG.log
B200PT_TRAVERSAL=binary timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
for blocks in 6 5; do
touch pbrt-v4_b200/csrc/traverse_persist.cu
B200PT_NVCC_EXTRA="-DB2_WIDE_BLOCKS=$blocks" python -c "import pbrt_v4_b200 as P; P.build()" || echo build failed
for nm in 10 14 16 18; do
echo -n "blocks=$blocks " | tee -a gpurun_out/sweep_$TAG.txt
B200PT_NODE_MIN=$nm timeout 300 python scripts/trav_one.py 23 2>&1 | tail -1 | tee -a gpurun_out/sweep_$TAG.txt
done; done
