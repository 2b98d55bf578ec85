"""Host-side multi-GPU plumbing: one process per GPU under torchrun.

cfg1-3 / cfg5 are data-parallel replicas over disjoint request sets (no data-path collective);
cfg4 is one tensor-parallel group whose ranks are fed the SAME requests and joined inside the
library by NCCL (csrc/tp.cu).  torch.distributed is only the control plane here: rendezvous,
the broadcast of the group's NCCL id, barriers and the max-over-ranks of the device timings.
The same functions run on the gloo backend for the CPU tests (tests/test_multiproc.py).
"""
from __future__ import annotations

import os
from typing import Callable, Optional


def env_ranks():
    """(rank, world_size, local_rank) from the torchrun environment (1 process: 0, 1, 0)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def plan(world: int, tensor_parallel: bool):
    """(tp_size, replicas): a tensor-parallel config uses every rank for ONE model replica."""
    tp = world if (tensor_parallel and world > 1) else 1
    return tp, (1 if tp > 1 else world)


def request_seed(base: int, rank: int, tp: int) -> int:
    """Replicas draw disjoint request sets; the ranks of a tensor-parallel group share one."""
    return base + (0 if tp > 1 else 1000 * rank)


def share_tp_id(dist, rank: int, make_id: Callable[[], bytes]) -> bytes:
    """Rank 0 creates the group's 128-byte NCCL unique id; every rank returns the same bytes."""
    box = [make_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    tid = box[0]
    if not isinstance(tid, (bytes, bytearray)) or len(tid) != 128:
        raise RuntimeError("tensor-parallel id exchange failed")
    return bytes(tid)


def reduce_max(dist, value: float, device: Optional[str] = None) -> float:
    """max over ranks of a per-rank device time (the job is as slow as its slowest rank)."""
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_throughput(tokens_per_replica: int, replicas: int, steps: int, ms_max: float) -> float:
    """whole-job tokens/s: all replicas' tokens over the slowest rank's time."""
    return replicas * tokens_per_replica * steps / (ms_max / 1e3)
