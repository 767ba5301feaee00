# A/B of prebuilt libraries: for every variants/<name>.so given (default: all), put it in place and print one bench summary
# line per workload in $WORKLOADS (default "me"). Parity is NOT checked here (run the suites on the variant you keep).
# usage (on the GPU box): bash profiles/ab_variants.sh [name ...]
cd "$(dirname "$0")/.."
LIB=svt-av1-1_b200/libsvtme_cuda.so
cp $LIB variants/_inplace.keep
names="$@"; [ -z "$names" ] && names=$(ls variants/*.so | xargs -n1 basename | sed 's/\.so$//')
for v in $names; do
  cp variants/$v.so $LIB
  for w in ${WORKLOADS:-me}; do
    timeout 300 python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline --no-strong $EXTRA 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v $w:', round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],3), {k:round(v,3) for k,v in (d.get('kernel_ms_per_step') or {}).items()}, round(d['roofline']['frac'],4), round((d.get('roofline_step') or {}).get('frac',0),4), d.get('result_checksum'))
for key in ('per_case', 'per_case_without_energy'):
    big = {k: round(v['ms'] * 1000, 1) for k, v in (d.get(key) or {}).items() if '64' in k}
    if big: print('   ', key, 'us:', big)"
  done
done
cp variants/_inplace.keep $LIB
