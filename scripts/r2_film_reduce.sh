# usage: bash scripts/r2_film_reduce.sh <ngpus> : scripts/film_reduce_check.py under torchrun
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 scripts/film_reduce_check.py > gpurun_out/film_reduce_n$N.log 2>&1; echo "rc=$?"; grep -v "^$\|OMP_NUM\|\*\*\*\*" gpurun_out/film_reduce_n$N.log | tail -8
