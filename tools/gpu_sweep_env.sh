#!/bin/bash
# usage: bash tools/gpu_sweep_env.sh "VAR=a VAR2=b" "VAR=c" ... ; each argument is one env set
# for tools/decode_time.py (cfg2 job: prefill ms + decode us/step)
for v in "$@"; do
  echo "== $v"; env $v timeout 300 python tools/decode_time.py $SWEEP_ARGS 2>&1 | tail -1
done
