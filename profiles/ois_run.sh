#!/bin/bash
# round 2: parity of svtme_ois for the kernel variants named in $1 (SVTME_OIS_KERNEL values), then the timing script plain
# and under the ncu launch-list pass for each of them
mkdir -p gpurun_out
for v in ${1:-0 1 2 3}; do
    SVTME_OIS_KERNEL=$v timeout 60 python -m pytest tests/test_zzz_ois_parity_gpu.py -q -x > gpurun_out/ois_tests_v$v.log 2>&1
    echo "pytest rc=$?" >> gpurun_out/ois_tests_v$v.log
    SVTME_OIS_KERNEL=$v timeout 30 python profiles/ois_time.py > gpurun_out/ois_time_v$v.log 2>&1 || continue
    SVTME_OIS_KERNEL=$v timeout 40 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ois --csv --log-file gpurun_out/ois_launches_v$v.csv \
        python profiles/ois_time.py > gpurun_out/ois_ncu_v$v.log 2>&1
done
exit 0
