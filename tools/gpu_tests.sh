#!/bin/bash
# One gpurun call: the whole GPU test-suite with per-test durations, then smoke().
O=gpurun_out; mkdir -p $O
timeout ${TEST_TIMEOUT:-1500} python -m pytest tests -m gpu -x -q --durations=15 "$@" > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -30 $O/pytest.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $O/smoke.log
