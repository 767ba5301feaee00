#!/bin/bash
# last GPU call of round 2: the bench's ois key on a short sequence, the three svtme_ois kernels against the oracle, and the
# tests that touch what changed in api.cu since the last full GPU suite (kernel kinds, context fields)
mkdir -p gpurun_out
timeout 60 python profiles/ois_bench_probe.py > gpurun_out/ois_bench_probe.log 2>&1; echo "rc=$?" >> gpurun_out/ois_bench_probe.log
for v in 0 1 2; do
    SVTME_OIS_KERNEL=$v timeout 60 python -m pytest tests/test_zzz_ois_parity_gpu.py -q -x > gpurun_out/ois_tests_v$v.log 2>&1
    echo "pytest rc=$?" >> gpurun_out/ois_tests_v$v.log
done
timeout 100 python -m pytest tests/test_zz_no_alloc.py tests/test_me_parity_gpu.py -q -x -m gpu -k "no_alloc or static or zz_sad or 64-64 or 200" > gpurun_out/sanity_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/sanity_tests.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 30 python profiles/ois_time.py > gpurun_out/ois_time_v0.log 2>&1 || exit 0
timeout 60 ncu --set full --clock-control none --import-source on -k regex:ois_lines --launch-skip 4 --launch-count 1 -o gpurun_out/ois_lines_full \
    python profiles/ois_time.py > gpurun_out/ois_ncu_full.log 2>&1
exit 0
