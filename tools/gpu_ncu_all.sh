#!/bin/bash
# One gpurun call: `ncu --set full` captures of every kernel family of the forward pass (windows of
# consecutive launches of this library's kernels), plus the launch list of a cfg2 job
# (B200_PROFILING.md recipe).  The reports are summarised ON the box (profiles/summarize.py) and
# deleted: only the markdown / csv summaries travel back (gpurun_out/ is capped at 64 MiB).
# usage: bash tools/gpu_ncu_all.sh <tag>
TAG=${1:-r2}
O=gpurun_out; mkdir -p $O
T=tools/ncu_target.py
OURS='regex:^(embed_kernel|norm_kernel|gemm_tc5_kernel|gemv_kernel|rope_kv_kernel|attn_prefill_tc5_kernel|attn_decode_t_kernel|attn_decode_mma_kernel|attn_combine_kernel|act_kernel|sample_kernel|sample_filtered_kernel|advance_kernel|make_decode_state_kernel|reduce_partials_kernel|moe_sort_gather_kernel|moe_sort_kernel|moe_gather_kernel)'
cap() {  # name, skip, count, target args...
  local name=$1 skip=$2 cnt=$3; shift 3
  timeout 900 ncu --set full --clock-control none -f -k "$OURS" -s $skip -c $cnt -o $O/full_${name}_$TAG python $T "$@" > $O/ncu_$name.log 2>&1
  echo "$name exit $?"
  if [ -f $O/full_${name}_$TAG.ncu-rep ]; then
    python profiles/summarize.py full $O/full_${name}_$TAG.ncu-rep $O/${TAG}_full_$name.md
    rm -f $O/full_${name}_$TAG.ncu-rep
  fi
}
export NVLLM_NCU_HOST_WEIGHTS=1   # no torch CUDA kernels in the launch stream
timeout 300 python $T > $O/plain_cfg2.log 2>&1 || { echo "plain cfg2 failed"; tail $O/plain_cfg2.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" -c 1200 --csv \
    --log-file $O/launches_$TAG.csv python $T > $O/ncu_l.log 2>&1; echo "launch list exit $?"
python profiles/summarize.py launches $O/launches_$TAG.csv $O/${TAG}_launches.md; rm -f $O/launches_$TAG.csv
# cfg2 (llama-3.2-1b, B=64, 512-token prompts): first prefill layer, then one whole decode step
cap cfg2_prefill   0   9
cap cfg2_decode  205 104
# cfg1 (gpt2, B=1): decode = GEMV / norm / act / MHA decode attention
cap cfg1_decode  215  26 gpt2 1 32
# cfg3 (granite-moe, B=128): router-in-norm, sort+gather, grouped GEMMs, combine-in-norm
cap cfg3_decode  460  24 granite-moe-1b 128 128
# Falcon-7B width, 2 layers, B=1, 512-token prompt: MQA prefill + decode attention, parallel block
NVLLM_NCU_LAYERS=2 cap cfg4_falcon 0 40 falcon-7b 1 512
# cfg5 head dim 128: prefill attention at 4096-token contexts (2 layers of llama-3.2-3b width)
NVLLM_NCU_LAYERS=2 cap cfg5_prefill 0 8 llama-3.2-3b 4 4096 0
NVLLM_NCU_SAMPLING=1 timeout 600 ncu --set full --clock-control none -f -k regex:sample_filtered -c 1 -o $O/full_sampling_$TAG python $T gpt2 1 32 > $O/ncu_sampling.log 2>&1; echo "sampling exit $?"
[ -f $O/full_sampling_$TAG.ncu-rep ] && python profiles/summarize.py full $O/full_sampling_$TAG.ncu-rep $O/${TAG}_full_sampling.md && rm -f $O/full_sampling_$TAG.ncu-rep
rm -f $O/*.ncu-rep
du -sh $O; ls $O | head -40
