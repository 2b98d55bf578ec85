#!/bin/bash
# One gpurun call: device timeline of steady-state decode + ncu launch list + ncu --set full of
# the decode GEMMs, the decode attention and the prefill GEMMs (B200_PROFILING.md recipe).
# usage: bash tools/gpu_profile.sh <tag>
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
timeout 300 python tools/timeline.py > $O/timeline_$TAG.txt 2>&1; echo "timeline exit $?"
timeout 300 python tools/ncu_target.py > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail $O/plain_$TAG.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv \
    --log-file $O/launches_$TAG.csv python tools/ncu_target.py > $O/ncu_l.log 2>&1; echo "launch list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc5 -s 130 -c 4 \
    -f -o $O/full_gemm_decode_$TAG python tools/ncu_target.py > $O/ncu_a.log 2>&1; echo "full gemm decode exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 16 -c 1 \
    -f -o $O/full_attn_decode_$TAG python tools/ncu_target.py > $O/ncu_b.log 2>&1; echo "full attn decode exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc5 -s 0 -c 4 \
    -f -o $O/full_gemm_prefill_$TAG python tools/ncu_target.py > $O/ncu_c.log 2>&1; echo "full gemm prefill exit $?"
ls -la $O
