# usage: bash scripts/r2_c3list.sh <tag> : config-3 frame (the committed coarse-tessellation tables: same materials, lights, light BVH)
# at 4 spp, then the ncu launch list of the same command
TAG=${1:-x}
mkdir -p gpurun_out
python scripts/render_time.py config3_full 4 > gpurun_out/c3_time_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/c3_launches_$TAG.csv python scripts/render_time.py config3_full 4 > gpurun_out/ncu_c3_$TAG.log 2>&1
python scripts/launch_summary.py gpurun_out/c3_launches_$TAG.csv | tee gpurun_out/c3_launches_summary_$TAG.txt | head -24
tail -3 gpurun_out/c3_time_$TAG.log
