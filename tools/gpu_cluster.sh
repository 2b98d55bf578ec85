#!/bin/bash
# cluster / DSMEM fused-residual split-K epilogue (opt-in): parity tests
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -s -k "cluster_shared_memory or cfg2_llama or chained or w2_gemv" > $O/cl_pytest.log 2>&1; echo "pytest exit $?"; grep -a "nvllm:\|launches per\|passed\|failed\|Error\|error\|worst rel" $O/cl_pytest.log | sort | uniq -c | head -40
