# usage: bash scripts/gpu_render_prof.sh <tag>  - launch list of one config-1 render + ncu --set full of the shade kernel
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py > gpurun_out/render_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -5 gpurun_out/render_time_$TAG.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/render_launches_$TAG.csv python scripts/render_time.py > gpurun_out/ncu_render_$TAG.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ShadeKernel -s 2 -c 1 -o gpurun_out/prof_shade_$TAG python scripts/render_time.py > gpurun_out/ncu_shade_$TAG.log 2>&1; echo "ncu shade rc=$?"
