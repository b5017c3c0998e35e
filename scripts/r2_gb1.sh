# usage: bash scripts/r2_gb1.sh <tag> : render GPU tests (gbuffer film among them)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_render_gpu.py tests/test_dropin_gpu.py -m gpu -x -q > gpurun_out/pytest_render_$TAG.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_render_$TAG.log
