#!/bin/bash
# the allocation-free entry point's test alone in a fresh process, then the whole GPU suite (it is collected last)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -5 $O/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_final.log 2>&1; echo "suite exit $?"; tail -8 $O/pytest_final.log
