# first GPU run of the render path: gpu tests, then a timed config-1 render (512x512, 64 spp)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_render1.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_render1.log
timeout 600 python scripts/render_time.py > gpurun_out/render_time1.log 2>&1; echo "render_time rc=$?"; tail -20 gpurun_out/render_time1.log
