#!/bin/bash
# decode-attention variants on the cfg2 job: parity first, then decode us/step per variant
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest.log
NVLLM_FUSE_ROPE=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "head_dim or long_context or engine_batch or resident" > $O/pytest_unfused.log 2>&1; echo "pytest unfused exit $?"; tail -3 $O/pytest_unfused.log
for v in "NVLLM_ATTN_T=0" "NVLLM_ATTN_T_KPC=1" "NVLLM_ATTN_T_KPC=2" "NVLLM_ATTN_T_KPC=4" ${EXTRA_VARIANTS}; do
  echo "== $v"; env $v timeout 120 python tools/decode_time.py "$@" 2>&1 | tail -1
done
timeout 300 python tools/timeline.py > $O/timeline_attn_t.txt 2>&1; echo "timeline exit $?"; grep attn_dec $O/timeline_attn_t.txt | head -4
