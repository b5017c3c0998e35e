# usage: bash scripts/r2_vol6.sh <tag> : config-5 frame time and the per-kernel launch list (ncu, after the plain run exited 0)
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -2 gpurun_out/vol_time_$TAG.log
timeout 600 python scripts/render_time.py config5_full 2 > gpurun_out/vol_time2_$TAG.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/vol_launches_$TAG.csv python scripts/render_time.py config5_full 2 > gpurun_out/ncu_vol_$TAG.log 2>&1; echo "ncu list rc=$?"
python scripts/launch_summary.py gpurun_out/vol_launches_$TAG.csv | tee gpurun_out/vol_launches_summary_$TAG.txt | head -20
