TAG=${1:-x}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_dropin_gpu.py -m gpu -x -q > gpurun_out/pytest_dropin_$TAG.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_dropin_$TAG.log
