# usage: bash scripts/gpu_retry.sh <timeout-seconds> '<command>' - gpurun, retried while the pod answers busy
T=$1; shift
mkdir -p gpurun_out
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > gpurun_out/.retry.log 2>&1
  if grep -q "status=transient" gpurun_out/.retry.log; then sleep 90; continue; fi
  break
done
cat gpurun_out/.retry.log
