#!/bin/bash
# round 2 profiles: launch list (time + DRAM bytes) of the TIMED positions of the default bench, full captures of the
# kernels that ship (persistent gemm_tf32x3 in-step, gemm_q8i in-step at 1 and 64 slots, the 7B prefill matmul and attention)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
K="gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu|tp_|chain_|sample_"
B="--no-cpu-baseline --no-e2e --no-extra --no-shares"
CMD="python bench.py --steps 20 --warmup 5 $B"
timeout 300 $CMD > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err || { tail -5 gpurun_out/r02_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:"$K" -s $((5 * 113)) -c $((20 * 113)) --csv --log-file gpurun_out/r02_launches_s110m_f32_B256_pos5-24.csv $CMD > gpurun_out/r02_ll.log 2>&1
tail -1 gpurun_out/r02_ll.log | cut -c1-200
S="python bench.py --steps 4 --warmup 3 $B"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tf32x3 -s 200 -c 4 -o gpurun_out/r02_gemm_tf32x3_B256 -f $S > gpurun_out/r02_full1.log 2>&1; tail -1 gpurun_out/r02_full1.log | cut -c1-160
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_q8i -s 200 -c 4 -o gpurun_out/r02_gemm_q8i_7b_B1 -f $S --model 7b --quant q8 --batch 1 --seq-len 512 > gpurun_out/r02_full2.log 2>&1; tail -1 gpurun_out/r02_full2.log | cut -c1-160
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_q8i -s 200 -c 4 -o gpurun_out/r02_gemm_q8i_7b_B64 -f $S --model 7b --quant q8 --batch 64 --seq-len 512 > gpurun_out/r02_full3.log 2>&1; tail -1 gpurun_out/r02_full3.log | cut -c1-160
P="python tools/bench_prefill.py --model 7b --len 2048 --reps 1"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_tf32x3|attn_prefill" -s 170 -c 6 -o gpurun_out/r02_prefill_7b_2048 -f $P > gpurun_out/r02_full4.log 2>&1; tail -1 gpurun_out/r02_full4.log | cut -c1-160
ls -la gpurun_out/*.ncu-rep
