# usage: bash scripts/r2_gb2.sh <tag> : GBufferFilm tests (incl. albedo) + the render sections of the bench
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_render_gpu.py tests/test_dropin_gpu.py -m gpu -x -q -k "gbuffer or exr or refuses or filter" > gpurun_out/pytest_gb_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gb_$TAG.log
bash scripts/r2_render_only.sh gb_$TAG
