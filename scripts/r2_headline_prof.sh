# usage: bash scripts/r2_headline_prof.sh <tag> : ncu --set full of the headline closest-hit / any-hit kernels (config 2, 2^24 rays) ->
# gpurun_out/headline_ncu_<tag>.json (scripts/ncu_extract.py; copy to profiles/r02_headline_ncu.json: bench.py's roofline reads it),
# per-line summary, details page, and the launch list of a bench step
TAG=${1:-x}
mkdir -p gpurun_out
REPS=2 python scripts/trav_one.py 24 > gpurun_out/trav_one_$TAG.log 2>&1 && \
REPS=2 ncu --set full --clock-control none -k regex:TraverseBinaryKernel -s 8 -c 2 -f -o gpurun_out/prof_headline_$TAG python scripts/trav_one.py 24 > gpurun_out/ncu_headline_$TAG.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_extract.py gpurun_out/prof_headline_$TAG.ncu-rep 16777216 > gpurun_out/headline_ncu_$TAG.json
python scripts/ncu_lines.py gpurun_out/prof_headline_$TAG.ncu-rep 0 50 > gpurun_out/headline_ncu_lines_$TAG.txt
ncu -i gpurun_out/prof_headline_$TAG.ncu-rep --page details > gpurun_out/headline_ncu_details_$TAG.txt 2>&1
rm -f gpurun_out/prof_headline_$TAG.ncu-rep
tail -2 gpurun_out/trav_one_$TAG.log; grep -E "kernel|duration|dram_bytes_per_ray|issue_active|active_threads" gpurun_out/headline_ncu_$TAG.json
python bench.py --steps 2 --warmup 1 --no-render --no-config3 --no-config4 --no-config5 --no-cpu-baseline > gpurun_out/bench_short_$TAG.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/headline_launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-render --no-config3 --no-config4 --no-config5 --no-cpu-baseline > gpurun_out/ncu_bench_$TAG.log 2>&1
python scripts/launch_summary.py gpurun_out/headline_launches_$TAG.csv | tee gpurun_out/headline_launches_summary_$TAG.txt | head
