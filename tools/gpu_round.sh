#!/bin/bash
# One gpurun call: tests, default bench, per-step DRAM-byte pass, one full capture of the tcgen05 matmul.
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
python -m pytest tests -x -q -m gpu > gpurun_out/r1e_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r1e_tests.log
python bench.py > gpurun_out/r1e_bench_default.json 2> gpurun_out/r1e_bench_default.err; echo "bench rc=$?"
CMD="python bench.py --steps 6 --warmup 4 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/r1e_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 560 -c 224 --csv \
    --log-file gpurun_out/r1e_launches_tc_B256.csv $CMD > gpurun_out/r1e_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tf32x3 -s 40 -c 5 -o gpurun_out/r1e_gemm_tf32x3_B256 -f $CMD > gpurun_out/r1e_ncu2.log 2>&1
ls -la gpurun_out/
