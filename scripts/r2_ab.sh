# usage: bash scripts/r2_ab.sh <tag> <variant>... : A/B of library builds kept under _ab/lib_<variant>.so - each is put in
# place of pbrt-v4_b200/libb200pt.so on the box and the render sections of the bench are run with it
TAG=${1:-x}; shift
mkdir -p gpurun_out
cp pbrt-v4_b200/libb200pt.so /tmp/lib_keep.so
for v in "$@"; do
  cp _ab/lib_$v.so pbrt-v4_b200/libb200pt.so
  timeout 900 python bench.py --steps 5 --warmup 3 ${AB_FLAGS---no-config3} --no-cpu-baseline --rays 4194304 ${AB_ARGS} > gpurun_out/bench_ab_${TAG}_$v.json 2> gpurun_out/bench_ab_${TAG}_$v.err; echo "$v bench rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_ab_${TAG}_$v.json'))
out=[]
for k in ("render", "config3", "config4", "config5"):
    c = d.get(k)
    if not c: continue
    r = c.get('render', c)
    out.append('%s %.1f' % (k, r['value']))
print('$v', ' | '.join(out))
PY
done
cp /tmp/lib_keep.so pbrt-v4_b200/libb200pt.so
