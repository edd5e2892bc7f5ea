#!/bin/bash
# One gpurun call: GPU tests, smoke, default bench + a few configurations, launch list of the default bench.
cd "$GRAFT_REPO_ROOT" || exit 1
T=${1:-v}
python -m pytest tests -x -q -m gpu > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${T}_tests.log
python __graft_entry__.py --smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${T}_smoke.log
python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err; echo "bench rc=$?"; cat gpurun_out/${T}_bench_default.json
for cfg in "s110m f32 1" "s110m f32 8" "s110m f32 64" "7b q8 1" "7b q8 8" "7b q8 64" "7b f32 1"; do
  set -- $cfg
  SL=""; [ "$1" = 7b ] && SL="--seq-len 512"
  python bench.py --model $1 --quant $2 --batch $3 $SL --steps 48 --warmup 4 --no-cpu-baseline --no-e2e > gpurun_out/${T}_bench_$1_$2_$3.json 2>gpurun_out/${T}_bench_$1_$2_$3.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_bench_$1_$2_$3.json").read().strip().splitlines()[-1])
    print("$cfg", round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],3), "frac", round(d["roofline"]["frac"],3))
except Exception as e: print("$cfg", "FAILED", e)
PY
done
