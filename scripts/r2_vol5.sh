# usage: bash scripts/r2_vol5.sh <tag> : volume-walk kernels: spectral unroll x resident blocks sweep on the config-5 frame
TAG=${1:-x}
mkdir -p gpurun_out
for cfg in "8 3" "4 3" "2 3" "8 4" "8 5" "16 3" "4 4"; do
set -- $cfg
B200PT_BUILD_ONLY=render_4 B200PT_NVCC_EXTRA="-UB2_SPECTRAL_UNROLL -DB2_SPECTRAL_UNROLL=$1 -DB2_VOL_MIN_BLOCKS=$2" python -c "import pbrt_v4_b200 as P; P.build()" 2>&1 | tail -2
echo -n "unroll=$1 blocks=$2 " | tee -a gpurun_out/vol_sweep_$TAG.txt
timeout 600 python scripts/render_time.py config5_full 16 2>&1 | grep "iter 2" | tee -a gpurun_out/vol_sweep_$TAG.txt
done
