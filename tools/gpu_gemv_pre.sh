#!/bin/bash
# W2 GEMV with the activation built in: parity tests, then A/B on cfg1 (GPT-2 small, one sequence)
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "w2_gemv_builds or gpt2 or falcon or resident or matmul" > $O/gp_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $O/gp_pytest.log
CFG=1 bash tools/gpu_ab.sh c1_act_on:NVLLM_GEMV_PRE=1 c1_act_off:NVLLM_GEMV_PRE=0
