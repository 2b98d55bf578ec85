"""world_size-2 gloo tests (CPU) of the N > 1 host logic bench.py runs under torchrun: replica
vs tensor-parallel planning, disjoint / shared request sets, the broadcast of the group's NCCL
id, max-over-ranks timing and the whole-job aggregate."""
import json
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(mode, world=2):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "mp_worker.py"), mode],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                      text=True))
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-2000:]
        line = [ln for ln in out.splitlines() if ln.startswith("MP ")][-1]
        outs.append(json.loads(line[3:]))
    return sorted(outs, key=lambda d: d["rank"])


def test_replicas_draw_disjoint_requests_and_aggregate():
    a, b = _run("dp")
    assert a["tp"] == b["tp"] == 1 and a["replicas"] == b["replicas"] == 2
    assert a["digest"] != b["digest"], "replicas must serve disjoint request sets"
    assert a["tables"] == b["tables"], "same shape of job => same block-table plan per replica"
    assert a["cached"] == b["cached"] == 3 * 256  # 3 of 4 prompts hit the shared 256-token block
    # slowest rank (20 ms) defines the job; both replicas' tokens count
    assert a["ms"] == b["ms"] == 20.0
    assert a["value"] == pytest.approx(2 * a["tokens"] * 2 / 0.020)


def test_tensor_parallel_ranks_share_requests_and_group_id():
    a, b = _run("tp")
    assert a["tp"] == b["tp"] == 2 and a["replicas"] == b["replicas"] == 1
    assert a["digest"] == b["digest"], "ranks of one TP group are fed the same requests"
    assert a["tid"] == b["tid"] == bytes(range(128)).hex()
    assert a["ms"] == b["ms"] == 20.0
    assert a["value"] == pytest.approx(a["tokens"] * 2 / 0.020)  # one replica's tokens only
