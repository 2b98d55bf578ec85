"""One rank of the world_size-2 CPU (gloo) test of the multi-GPU host logic (tests/test_multiproc.py).
usage: mp_worker.py <mode: dp|tp>   (RANK / WORLD_SIZE / MASTER_* from the environment)"""
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import C  # noqa: E402
import nvllm_pkg  # noqa: E402

import torch.distributed as dist  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    mode = sys.argv[1]
    D = nvllm_pkg.sub("dist")
    rank, world, local_rank = D.env_ranks()
    dist.init_process_group("gloo")
    tp, replicas = D.plan(world, tensor_parallel=(mode == "tp"))
    cfg = C.tiny_llama()
    job = bench.Job(cfg, 4, 300, 8, seed0=D.request_seed(4321, rank, tp), shared_prefix=256)
    digest = hashlib.sha256(job.prompts.tobytes()).hexdigest()
    tid = None
    if tp > 1:
        tid = D.share_tp_id(dist, rank, lambda: bytes(range(128))).hex()
    # a fake per-rank device time: the slowest rank defines the job
    ms = D.reduce_max(dist, 10.0 * (rank + 1))
    value = D.aggregate_throughput(job.tokens_per_step, replicas, 2, ms)
    dist.barrier()
    print("MP " + json.dumps({"rank": rank, "world": world, "tp": tp, "replicas": replicas,
                              "digest": digest, "tid": tid, "ms": ms, "value": value,
                              "tokens": job.tokens_per_step, "cached": int(job.num_cached.sum()),
                              "tables": job.tables}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
