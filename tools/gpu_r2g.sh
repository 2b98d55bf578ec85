#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "chained or full_width_batch64 or full_depth or granite_moe_full or logits_and_greedy" > $O/pytest_chain.log 2>&1; echo "pytest exit $?"; tail -6 $O/pytest_chain.log
CFG=2 bash tools/gpu_ab.sh chain:X=1 nochain:NVLLM_CHAIN=0
