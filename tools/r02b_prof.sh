#!/bin/bash
# round 2, second pass (after the read-out / attention / SwiGLU / prep changes): launch list of the timed positions of the default
# bench, full captures of the shipped kernels in-step, of the bf16 prompt pass (tcgen05 kind::f16 matmul + tensor-core attention)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
K="gemm_|tc_|attn_|argmax|advance|gemv|q8_prep|mega|rmsnorm|rope|swiglu|tp_|chain_|sample_|bf_"
B="--no-cpu-baseline --no-e2e --no-extra --no-shares"
CMD="python bench.py --steps 20 --warmup 5 $B"
timeout 300 $CMD > gpurun_out/r02b_plain.json 2> gpurun_out/r02b_plain.err || { tail -5 gpurun_out/r02b_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:"$K" -s $((5 * 101)) -c $((20 * 101)) --csv --log-file gpurun_out/r02b_launches_s110m_f32_B256_pos5-24.csv $CMD > gpurun_out/r02b_ll.log 2>&1
tail -1 gpurun_out/r02b_ll.log | cut -c1-200
S="python bench.py --steps 4 --warmup 3 $B"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_tf32x3|attn_fused" -s 200 -c 6 -o gpurun_out/r02b_step_s110m_f32_B256 -f $S > gpurun_out/r02b_full1.log 2>&1; tail -1 gpurun_out/r02b_full1.log | cut -c1-160
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_q8i -s 200 -c 4 -o gpurun_out/r02b_gemm_q8i_7b_B64 -f $S --model 7b --quant q8 --batch 64 --seq-len 512 > gpurun_out/r02b_full3.log 2>&1; tail -1 gpurun_out/r02b_full3.log | cut -c1-160
P="python tools/bench_prefill.py --model 7b --len 2048 --reps 1 --bf16 1"
timeout 300 $P > gpurun_out/r02b_prefill_bf16_plain.txt 2>&1; tail -1 gpurun_out/r02b_prefill_bf16_plain.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|attn_prefill_tc" -s 170 -c 6 -o gpurun_out/r02b_prefill_7b_bf16_2048 -f $P > gpurun_out/r02b_full4.log 2>&1; tail -1 gpurun_out/r02b_full4.log | cut -c1-160
ls -la gpurun_out/*.ncu-rep
