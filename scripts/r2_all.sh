# usage: bash scripts/r2_all.sh <tag> : what the driver runs at round end - the GPU test suite and smoke()
TAG=${1:-x}
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_all_$TAG.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_all_$TAG.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
