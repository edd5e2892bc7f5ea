#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sampler.py -m gpu -q -s > gpurun_out/f_tests.log 2>&1; echo "rc=$?" >> gpurun_out/f_tests.log
grep -E "identical|passed|failed|rc=|Error" gpurun_out/f_tests.log | head -20
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err ) 2> gpurun_out/f_bench.time
tail -3 gpurun_out/f_bench.time; tail -5 gpurun_out/f_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/f_bench.json'))
print('value', d['value'], 'ms', d['ms_per_step'])
print('e2e', {k:(v['value'] if isinstance(v,dict) else v) for k,v in d['e2e'].items() if k in ('value','greedy_call','logits_call')})
print('shares', d['roofline']['live_shares_us'])
print('dominant', d['roofline']['dominant_kernel'])
print('parity', d.get('parity_sample'))
for k,v in d.get('extra',{}).items(): print(k, {a:b for a,b in v.items() if a!='workload'} if isinstance(v,dict) else v)
PY
