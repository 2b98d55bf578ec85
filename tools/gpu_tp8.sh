#!/bin/bash
# 8-GPU box: the default bench under torchrun (cfg2 x 8 replicas + Falcon-7B at tp = 8)
N=${1:-8}
O=gpurun_out; mkdir -p $O
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n=$N exit $?"; tail -3 $O/bench_n$N.err | cut -c1-300
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_n{N}.json").read().strip().splitlines()[-1])
print("value",round(d["value"]),"e2e",round(d["e2e"]["value"]),"n_gpus",d["n_gpus"]); print(d["phases"]); print("tp",json.dumps(d["tp"],indent=1))
PY
