#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -k "matmul or gpt2 or falcon or resident" > $O/pytest_gemv.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gemv.log
CFG=1 bash tools/gpu_ab.sh c1:X=1
