# The evidence run of a build (one GPU): full GPU suite, smoke(), the default bench line, the reference arm, the secondary
# workloads, then the ncu launch list of the bench command. Outputs under gpurun_out/final_*; copy what is kept into profiles/rNN/.
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/final_tests.log 2>&1; echo rc=$? >> gpurun_out/final_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo rc=$? >> gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench_me_n1.json 2> gpurun_out/final_bench_me_n1.err
timeout 600 python bench.py --impl reference > gpurun_out/final_ref_arm.json 2>/dev/null
for w in ${WORKLOADS:-me209 streams me4k}; do timeout 600 python bench.py --workload $w > gpurun_out/final_bench_$w.json 2>/dev/null; done
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-strong > /dev/null 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_me.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-strong > gpurun_out/final_ncu_launch.log 2>&1
