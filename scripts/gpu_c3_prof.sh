# usage: bash scripts/gpu_c3_prof.sh <tag>  - config 3 frame: plain timing, launch list of a 2-spp frame (one pass)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py config3_full 8 > gpurun_out/c3_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -5 gpurun_out/c3_time_$TAG.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/c3_launches_$TAG.csv python scripts/render_time.py config3_full 2 > gpurun_out/ncu_c3_$TAG.log 2>&1; echo "ncu list rc=$?"
if [ "$2" = "full" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ShadeKernel -s 3 -c 1 -o gpurun_out/prof_c3_coated_$TAG python scripts/render_time.py config3_full 2 > gpurun_out/ncu_c3_coated_$TAG.log 2>&1; echo "ncu coated rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:TraversePersistKernel -s 0 -c 1 -o gpurun_out/prof_c3_trav_$TAG python scripts/render_time.py config3_full 2 > gpurun_out/ncu_c3_trav_$TAG.log 2>&1; echo "ncu trav rc=$?"
fi
