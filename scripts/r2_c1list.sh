# usage: bash scripts/r2_c1list.sh <tag> : config-1 frame time, then (after it exited 0) the ncu launch list of the same command
TAG=${1:-x}
mkdir -p gpurun_out
python scripts/render_time.py cornell_r512_s64 > gpurun_out/c1_time_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/c1_launches_$TAG.csv python scripts/render_time.py cornell_r512_s64 > gpurun_out/ncu_c1_$TAG.log 2>&1
python scripts/launch_summary.py gpurun_out/c1_launches_$TAG.csv | tee gpurun_out/c1_launches_summary_$TAG.txt | head -20
tail -3 gpurun_out/c1_time_$TAG.log
