# usage: bash scripts/gpu_tests.sh <tag> [pytest -k expr]
TAG=${1:-x}; K=${2:-}
mkdir -p gpurun_out
if [ -n "$K" ]; then timeout 1500 python -m pytest tests -m gpu -x -q -k "$K" > gpurun_out/pytest_gpu_$TAG.log 2>&1; else timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1; fi
echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu_$TAG.log
