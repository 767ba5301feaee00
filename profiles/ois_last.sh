#!/bin/bash
# round 2, last OIS call: the bench's ois key on a short sequence, parity of the three kernels, launch list of the shipped one
mkdir -p gpurun_out
timeout 60 python profiles/ois_bench_probe.py > gpurun_out/ois_bench_probe.log 2>&1; echo "rc=$?" >> gpurun_out/ois_bench_probe.log
for v in 0 1 2; do
    SVTME_OIS_KERNEL=$v timeout 60 python -m pytest tests/test_zzz_ois_parity_gpu.py -q -x > gpurun_out/ois_tests_v$v.log 2>&1
    echo "pytest rc=$?" >> gpurun_out/ois_tests_v$v.log
done
timeout 30 python profiles/ois_time.py > gpurun_out/ois_time_v0.log 2>&1 || exit 0
timeout 40 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ois --csv --log-file gpurun_out/ois_launches_v0.csv \
    python profiles/ois_time.py > gpurun_out/ois_ncu_v0.log 2>&1
exit 0
