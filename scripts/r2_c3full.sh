# usage: bash scripts/r2_c3full.sh <tag> : config-3 frame (coarse tables, 4 spp; plain run first), then ncu --set full of the four
# first-vertex shade launches (one per material queue, light-BVH sampler)
TAG=${1:-x}
mkdir -p gpurun_out
python scripts/render_time.py config3_full 4 > gpurun_out/c3_time2_$TAG.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:ShadeKernel -s 0 -c 4 -f -o gpurun_out/prof_c3shade_$TAG python scripts/render_time.py config3_full 4 > gpurun_out/ncu_c3shade_$TAG.log 2>&1
python scripts/ncu_extract.py gpurun_out/prof_c3shade_$TAG.ncu-rep > gpurun_out/c3shade_ncu_$TAG.json
for i in 0 1 2 3; do python scripts/ncu_lines.py gpurun_out/prof_c3shade_$TAG.ncu-rep $i 40 > gpurun_out/c3shade_ncu_lines_${i}_$TAG.txt; done
rm -f gpurun_out/prof_c3shade_$TAG.ncu-rep
head -30 gpurun_out/c3shade_ncu_lines_0_$TAG.txt
