# usage: bash scripts/gpurun_retry.sh <timeout-seconds> [--gpus N] '<command>' : retries while the pod answers busy (exit 3)
T=$1; shift
G=""
if [ "$1" = "--gpus" ]; then G="--gpus $2"; shift; shift; fi
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun --timeout $T $G -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
