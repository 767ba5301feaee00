# Sweep behind the "chunked submission" paragraph of DESIGN.md section 4. SVTME_FP_SMEM_KB was an experiment-only knob of the
# full-pel kernel (inflated dynamic shared memory = one CTA per SM); it was removed with the experiment, so the "smem" column of
# this script only reproduces with that patch applied.
SVTME_CHUNKS=5 SVTME_FP_SMEM_KB=120 python -m pytest tests/test_me_parity_gpu.py -q -x 2>&1 | tail -1
for cfg in "1 0" "2 0" "4 0" "8 0" "4 120" "8 120" "15 120" "30 120" "8 150"; do set -- $cfg
  SVTME_CHUNKS=$1 SVTME_FP_SMEM_KB=$2 python bench.py --workload me --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunks $1 smem $2:', round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})"
done
