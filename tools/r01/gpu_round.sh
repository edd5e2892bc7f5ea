#!/bin/bash
# Everything round 1 measured on ONE GPU, in one gpurun call (about 6 minutes of box time):
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh'
# tests + smoke + default bench + the configuration table of BASELINE.md section 4 (gpu_verify.sh), then the launch lists
# and the full capture of the dominant kernel (gpu_prof_r01k.sh).  Multi-GPU: tools/gpu_scale.sh N (gpurun --gpus N),
# tools/tp_check.py and tools/bench_prefill_tp.py under torchrun.  Sweeps that led to the current defaults:
# tools/attn_sweep.sh (attention segment width), tools/z_sweep.sh (split-K planner), tools/skip_sweep.sh (live share of
# each kernel class), tools/dual_sweep.sh (two concurrent half-batch branches: slower, off).
cd "$GRAFT_REPO_ROOT" || exit 1
bash tools/gpu_verify.sh round
bash tools/gpu_prof_r01k.sh
ls -la gpurun_out/ | tail -30
