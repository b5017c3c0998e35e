# usage: bash scripts/gpu_traffic.sh <tag> - DRAM bytes of the headline traversal kernels (closest, any) at the bench's full batch
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum --clock-control none -k regex:TraversePersist -s 2 -c 4 --csv --log-file gpurun_out/traffic_$TAG.csv python bench.py --steps 1 --warmup 2 --no-render --no-config3 --no-config4 --no-config5 --no-cpu-baseline > gpurun_out/ncu_traffic_$TAG.log 2>&1; echo "ncu rc=$?"
grep -v "^==" gpurun_out/traffic_$TAG.csv | cut -d, -f5,13- | head -30
