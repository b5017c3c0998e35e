# usage: bash scripts/gpu_shade_prof.sh <tag> - config-1 render: timing + ncu --set full of the diffuse shade kernel (2nd vertex)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py > gpurun_out/render_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -4 gpurun_out/render_time_$TAG.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ShadeKernel -s 1 -c 1 -o gpurun_out/prof_shade_$TAG python scripts/render_time.py > gpurun_out/ncu_shade_$TAG.log 2>&1; echo "ncu shade rc=$?"
