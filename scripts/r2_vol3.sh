# usage: bash scripts/r2_vol3.sh <tag> : ncu --set full of the two volume-walk kernels (config-5 frame, 2 spp), one launch each
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python scripts/render_time.py config5_full 2 > gpurun_out/vol_time2_$TAG.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none -k regex:ShadowTrKernel -s 1 -c 1 -f -o gpurun_out/prof_shadowtr_$TAG python scripts/render_time.py config5_full 2 > gpurun_out/ncu_vol_$TAG.log 2>&1; echo "ncu rc=$?"
timeout 1500 ncu --set full --clock-control none -k regex:VolMediumKernel -s 1 -c 1 -f -o gpurun_out/prof_volmedium_$TAG python scripts/render_time.py config5_full 2 > gpurun_out/ncu_vol2_$TAG.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
for f in shadowtr volmedium; do
python scripts/ncu_extract.py gpurun_out/prof_${f}_$TAG.ncu-rep > gpurun_out/prof_${f}_$TAG.json
python scripts/ncu_lines.py gpurun_out/prof_${f}_$TAG.ncu-rep 0 60 > gpurun_out/prof_${f}_${TAG}_lines.txt
ncu -i gpurun_out/prof_${f}_$TAG.ncu-rep --page details > gpurun_out/prof_${f}_${TAG}_details.txt 2>&1
rm -f gpurun_out/prof_${f}_$TAG.ncu-rep
done
