#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 -k "head_dim or full_width or full_depth or long_context or prefix or engine or logits_and_greedy or free_running" > $O/pytest_attn.log 2>&1; echo "pytest exit $?"; tail -15 $O/pytest_attn.log
CFG=2 bash tools/gpu_ab.sh tc5:X=1 mma:NVLLM_ATTN_PREFILL_TC5=0
CFG=5 bash tools/gpu_ab.sh c5tc5:X=1 c5mma:NVLLM_ATTN_PREFILL_TC5=0
