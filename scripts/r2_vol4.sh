# usage: bash scripts/r2_vol4.sh <tag> : volume-path GPU tests, config-5 frame time
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_render_gpu.py tests/test_dropin_gpu.py -m gpu -x -q -k "config5 or volpath or tight" > gpurun_out/pytest_vol_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_vol_$TAG.log
timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/vol_time_$TAG.log
B200PT_PASS_PATHS=34000000 timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_big_$TAG.log 2>&1; echo "big-pass rc=$?"; tail -3 gpurun_out/vol_time_big_$TAG.log
