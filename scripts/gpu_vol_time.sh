# usage: bash scripts/gpu_vol_time.sh <tag> - volpath tests + config-5 timing
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "config5 or config4 or golden_image" > gpurun_out/pytest_vol_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_vol_$TAG.log
timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -5 gpurun_out/vol_time_$TAG.log
