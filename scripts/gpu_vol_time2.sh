TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/vol_time_$TAG.log
B200PT_VOL_KERNEL=thread timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_thr_$TAG.log 2>&1; echo "thread rc=$?"; tail -2 gpurun_out/vol_time_thr_$TAG.log
