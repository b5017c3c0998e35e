# usage: bash scripts/gpu_times.sh <tag>  - frame times of config 1 and config 3 (8 spp), no profiler
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py > gpurun_out/c1_time_$TAG.log 2>&1; echo "c1 rc=$?"; tail -3 gpurun_out/c1_time_$TAG.log
timeout 600 python scripts/render_time.py config3_full 8 > gpurun_out/c3_time_$TAG.log 2>&1; echo "c3 rc=$?"; tail -3 gpurun_out/c3_time_$TAG.log
