#!/bin/bash
# round 2, last GPU call: parity of svtme_ois, then its timing and the ncu launch list of the timing command
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_zzz_ois_parity_gpu.py -q -x > gpurun_out/ois_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/ois_tests.log
timeout 40 python profiles/ois_time.py > gpurun_out/ois_time.log 2>&1 || exit 0
timeout 70 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:ois_kernel --csv --log-file gpurun_out/ois_launches.csv \
    python profiles/ois_time.py > gpurun_out/ois_ncu.log 2>&1
exit 0
