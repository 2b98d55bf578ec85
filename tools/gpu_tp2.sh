#!/bin/bash
# N-GPU box: TP parity tests, then the default bench under torchrun (cfg2 replicas + Falcon-7B at tp = N)
N=${1:-2}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_tp.py -m gpu -x -q > $O/pytest_tp$N.log 2>&1; echo "pytest tp exit $?"; tail -5 $O/pytest_tp$N.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n=$N exit $?"; tail -3 $O/bench_n$N.err | cut -c1-300
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_n{N}.json").read().strip().splitlines()[-1])
print("value",round(d["value"]),"e2e",round(d["e2e"]["value"]),"n_gpus",d["n_gpus"]); print(d["phases"]); print("tp",json.dumps(d["tp"],indent=1))
PY
