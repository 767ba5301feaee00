# Does cutting the 120-picture batch into chunks small enough for their per-SB scratch (32 KB of prediction blocks per SB and
# list pair) to stay in the 126 MB L2 pay? chunks x streams, per-kernel event timing off (it forces one stream).
cd "$(dirname "$0")/.."
for cfg in "1 1" "30 2" "60 2" "15 2" "30 3" "120 2" "1 1"; do set -- $cfg
  SVTME_BENCH_NO_KERNEL_TIMES=1 SVTME_CHUNKS=$1 SVTME_CHUNK_STREAMS=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-strong 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunks $1 streams $2:', round(d['value'],1), round(d['ms_per_step'],3), d.get('result_checksum'))"
done
