# usage: bash scripts/r2_c1full.sh <tag> : config-1 frame (plain run first), then ncu --set full of the diffuse shade kernel's first two launches
TAG=${1:-x}
mkdir -p gpurun_out
python scripts/render_time.py cornell_r512_s64 > gpurun_out/c1_time2_$TAG.log 2>&1 && \
ncu --set full --import-source on --clock-control none -k regex:ShadeKernel -s 0 -c 2 -f -o gpurun_out/prof_shade_$TAG python scripts/render_time.py cornell_r512_s64 > gpurun_out/ncu_shade_$TAG.log 2>&1
python scripts/ncu_extract.py gpurun_out/prof_shade_$TAG.ncu-rep > gpurun_out/shade_ncu_$TAG.json
python scripts/ncu_lines.py gpurun_out/prof_shade_$TAG.ncu-rep 0 60 > gpurun_out/shade_ncu_lines_$TAG.txt
ncu -i gpurun_out/prof_shade_$TAG.ncu-rep --page details > gpurun_out/shade_ncu_details_$TAG.txt 2>&1
rm -f gpurun_out/prof_shade_$TAG.ncu-rep
head -12 gpurun_out/shade_ncu_lines_$TAG.txt
