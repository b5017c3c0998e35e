# usage: bash scripts/gpu_vol_prof.sh <tag> [spp]  - config-5 render: timing, launch list, ncu --set full of the two volume kernels
TAG=${1:-x}; SPP=${2:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "config5 or config4 or golden_image" > gpurun_out/pytest_vol_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_vol_$TAG.log
timeout 600 python scripts/render_time.py config5_full 16 > gpurun_out/vol_time_$TAG.log 2>&1; echo "plain rc=$?"; tail -5 gpurun_out/vol_time_$TAG.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/vol_launches_$TAG.csv python scripts/render_time.py config5_full $SPP > gpurun_out/ncu_vol_$TAG.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:VolMediumKernel -s 3 -c 1 -o gpurun_out/prof_volmedium_$TAG python scripts/render_time.py config5_full $SPP > gpurun_out/ncu_volm_$TAG.log 2>&1; echo "ncu volmedium rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ShadowTrKernel -s 3 -c 1 -o gpurun_out/prof_shadowtr_$TAG python scripts/render_time.py config5_full $SPP > gpurun_out/ncu_shtr_$TAG.log 2>&1; echo "ncu shadowtr rc=$?"
