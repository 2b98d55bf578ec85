#!/bin/bash
N=${1:-2}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_tp.py -m gpu -x -q -k "pipeline or nccl_path" > $O/pytest_tp_pipe$N.log 2>&1; echo "pytest exit $?"; tail -15 $O/pytest_tp_pipe$N.log | cut -c1-400
for pipe in 1 0; do
NVLLM_TP_PIPE=$pipe timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --no-profile --no-cpu-baseline > $O/bench_pipe${pipe}_n$N.json 2> $O/bench_pipe${pipe}_n$N.err; echo "bench pipe=$pipe exit $?"; tail -3 $O/bench_pipe${pipe}_n$N.err | cut -c1-300
done
python - $N <<'PY'
import json,sys
N=sys.argv[1]
for pipe in (1,0):
    try:
        d=json.loads(open(f"gpurun_out/bench_pipe{pipe}_n{N}.json").read().strip().splitlines()[-1])
        t=d["tp"]; print("pipe",pipe,{k:t.get(k) for k in ("tp","value","prefill_tok_s","decode_tok_s","prefill_ms","comm_share","error")}, t.get("parity"))
    except Exception as e: print("pipe",pipe,"ERR",e)
PY
