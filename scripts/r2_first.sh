# round 2, first GPU call: the no-alloc entry point alone, then traversal tests + wide vs binary bench + all GPU tests
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 gpurun_out/pytest_noalloc_alone.log
bash scripts/r2_wide1.sh ${1:-a}
