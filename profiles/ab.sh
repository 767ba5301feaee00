# A/B helper: parity of the two ME suites, then `bench.py --workload $w` for each workload given
timeout 900 python -m pytest tests/test_me_parity_gpu.py tests/test_me_ext209_gpu.py -q -x 2>&1 | tail -2
for w in "$@"; do timeout 300 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w:', round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['kernel_ms_per_step'].items()}, round(d['roofline']['frac'],4), round(d['roofline_step']['frac'],4))"; done
