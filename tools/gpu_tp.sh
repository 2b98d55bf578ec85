#!/bin/bash
# N-GPU box: TP parity tests, then cfg4 (Falcon-7B) at tp=N with the peer-memory all-reduce and with NCCL
N=${1:-2}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_tp.py -m gpu -x -q > $O/pytest_tp.log 2>&1; echo "pytest tp exit $?"; tail -5 $O/pytest_tp.log
for p2p in 1; do
NVLLM_TP_P2P=$p2p timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --config 4 --gpus $N --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_cfg4_tp${N}_p2p$p2p.json 2> $O/bench_cfg4_tp${N}_p2p$p2p.err; echo "cfg4 tp$N p2p=$p2p exit $?"; tail -3 $O/bench_cfg4_tp${N}_p2p$p2p.err | cut -c1-300
done
python - $N <<'PY'
import json,sys
N=sys.argv[1]
for f in (f"bench_cfg4_tp{N}_p2p1",f"bench_cfg4_tp{N}_p2p0"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "value",round(d["value"]),"e2e",round(d["e2e"]["value"]),d["phases"])
        for k,v in d["kernels"].items():
            if "comm" in k or "gemv" in k or "gemm_tc5" in k: print("  ",k,v)
    except Exception as e: print(f, "ERR", e)
PY
