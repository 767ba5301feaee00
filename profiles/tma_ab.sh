# A/B of the raw-window staging of the full-pel kernels: SVTME_TMA=0 (16-byte cp.async granules) against the default (one 2-D
# TMA box per SB). usage (GPU box): bash profiles/tma_ab.sh [workload ...]    (default: me me209)
cd "$(dirname "$0")/.."
for w in ${@:-me me209}; do
  for t in 0 1 0 1; do
    SVTME_TMA=$t timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-strong 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('SVTME_TMA=$t $w:', round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()}, round(d['roofline']['frac'],4), d.get('result_checksum'))"
  done
done
