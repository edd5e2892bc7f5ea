#!/bin/bash
# one multi-GPU call: tools/gpu_scale.sh N  -> data-parallel default bench on N GPUs, TP-N parity check, TP-N 7B Q8_0 / fp32 decode
cd "$GRAFT_REPO_ROOT" || exit 1
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
show() { python -c "
import json,sys
try:
    d=json.loads(open('$1').read().strip().splitlines()[-1]); print('$2', round(d['value'],1), d['unit'], round(d['ms_per_step'],4), 'ms/step frac', round(d['roofline']['frac'],3), 'e2e', (round(d['e2e']['value'],1) if d.get('e2e') else None))
except Exception as e: print('$2 FAILED', e)"; }
timeout 400 $TR --master-port 29521 bench.py --gpus $N --no-cpu-baseline > gpurun_out/scale_dp${N}.json 2> gpurun_out/scale_dp${N}.err; show gpurun_out/scale_dp${N}.json "dp$N s110m f32 B256/gpu"
timeout 300 $TR --master-port 29522 tools/tp_check.py 2>&1 | grep -E "^tp|TP_CHECK|MISMATCH|timed out|disagree" | tail -8
for cfg in "q8 1" "q8 8" "q8 64" "f32 16"; do
  set -- $cfg
  timeout 400 $TR --master-port 29523 bench.py --gpus $N --tp $N --model 7b --quant $1 --batch $2 --seq-len 512 --steps 48 --warmup 4 --no-cpu-baseline > gpurun_out/scale_tp${N}_7b_$1_B$2.json 2> gpurun_out/scale_tp${N}_7b_$1_B$2.err; show gpurun_out/scale_tp${N}_7b_$1_B$2.json "tp$N 7b $1 B$2"
done
