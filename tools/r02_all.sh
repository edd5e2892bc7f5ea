#!/bin/bash
# the whole GPU suite + smoke + the default bench line, as the driver runs them at round end
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/all_tests.log 2>&1 ) 2> gpurun_out/all_tests.time; echo "rc=$?" >> gpurun_out/all_tests.log
tail -6 gpurun_out/all_tests.log | cut -c1-300; tail -3 gpurun_out/all_tests.time
timeout 600 python __graft_entry__.py --smoke > gpurun_out/all_smoke.log 2>&1; tail -3 gpurun_out/all_smoke.log
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/all_bench.json 2> gpurun_out/all_bench.err ) 2> gpurun_out/all_bench.time
tail -3 gpurun_out/all_bench.time; tail -3 gpurun_out/all_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/all_bench.json'))
print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'])
for k,v in d.get('extra',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a not in ('workload','note')} if isinstance(v,dict) else v)
PY
