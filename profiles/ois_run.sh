#!/bin/bash
# round 2: parity of svtme_ois for every kernel variant (SVTME_OIS_KERNEL: 0 = word-wise, 2 CTAs/SM; 1 = pixel-wise;
# 2 / 3 = word-wise at 3 / 4 CTAs/SM), then the timing script plain and under the ncu launch-list pass
mkdir -p gpurun_out
for v in 0 1 2 3; do
    SVTME_OIS_KERNEL=$v timeout 60 python -m pytest tests/test_zzz_ois_parity_gpu.py -q -x > gpurun_out/ois_tests_v$v.log 2>&1
    echo "pytest rc=$?" >> gpurun_out/ois_tests_v$v.log
done
for v in 0 1 2 3; do
    SVTME_OIS_KERNEL=$v timeout 30 python profiles/ois_time.py > gpurun_out/ois_time_v$v.log 2>&1
done
for v in 0 3; do
    SVTME_OIS_KERNEL=$v timeout 40 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ois --csv --log-file gpurun_out/ois_launches_v$v.csv \
        python profiles/ois_time.py > gpurun_out/ois_ncu_v$v.log 2>&1
done
exit 0
