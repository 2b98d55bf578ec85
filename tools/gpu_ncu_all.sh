#!/bin/bash
# One gpurun call: `ncu --set full` captures of every kernel family of the forward pass, plus the
# launch list of a cfg2 job (B200_PROFILING.md recipe).  usage: bash tools/gpu_ncu_all.sh <tag>
TAG=${1:-r2}
O=gpurun_out; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on -f"
T=tools/ncu_target.py
cap() {  # name, kernel regex, skip, count, target args...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  timeout 600 $NCU -k regex:$rx -s $skip -c $cnt -o $O/full_${name}_$TAG python $T "$@" > $O/ncu_$name.log 2>&1
  echo "$name exit $?"
}
# plain runs first (a number or report taken from a failing program is worthless)
timeout 300 python $T > $O/plain_cfg2.log 2>&1 || { echo "plain cfg2 failed"; tail $O/plain_cfg2.log; exit 1; }
timeout 300 python $T gpt2 1 32 > $O/plain_cfg1.log 2>&1 || { echo "plain cfg1 failed"; tail $O/plain_cfg1.log; exit 1; }
timeout 300 python $T granite-moe-1b 128 128 > $O/plain_cfg3.log 2>&1 || { echo "plain cfg3 failed"; tail $O/plain_cfg3.log; exit 1; }
# launch list of the cfg2 job (prefill + 3 decode steps)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv \
    --log-file $O/launches_$TAG.csv python $T > $O/ncu_l.log 2>&1; echo "launch list exit $?"
# cfg2 (llama-3.2-1b, B=64): prefill kernels are the first launches, decode after ~150
cap gemm_prefill   gemm_tc5          0 4
cap attn_prefill   attn_prefill_tc5  1 1
cap rope_kv        rope_kv           1 1
cap embed          embed_kernel      0 2
cap gemm_decode    gemm_tc5        130 5
cap lmhead_decode  gemm_tc5        129 1
cap attn_decode    attn_decode_t    17 1
cap norm_decode    norm_kernel       4 2
cap sample         sample_kernel     1 1
cap advance        advance_kernel    0 1
# cfg1 (gpt2, B=1): GEMV / act / norm / MHA decode attention
cap gemv           gemv_kernel      10 4 gpt2 1 32
cap act            act_kernel       13 1 gpt2 1 32
cap norm_b1        norm_kernel      30 1 gpt2 1 32
cap attn_dec_mha   attn_decode      13 1 gpt2 1 32
# cfg3 (granite-moe, B=128): router-in-norm, sort+gather, grouped GEMMs
cap moe_norm       norm_kernel      52 2 granite-moe-1b 128 128
cap moe_sort       moe_sort         25 1 granite-moe-1b 128 128
cap moe_gemm       gemm_tc5        104 6 granite-moe-1b 128 128
# falcon-style MQA decode attention (71 heads -> attn_decode_mma) on a 2-layer falcon-7b width
NVLLM_NCU_LAYERS=2 cap attn_dec_mqa attn_decode_mma 1 1 falcon-7b 1 512
NVLLM_NCU_SAMPLING=1 cap sample_filtered sample_filtered 0 1 gpt2 1 32
ls -la $O | head -60
