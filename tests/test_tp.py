"""Tensor-parallel parity: one process per GPU, each holding one rank's
handle, joined through NCCL inside the library.  Every rank must reproduce the single-device
CPU oracle within the precision's gate.  Needs >= 2 GPUs (skipped on a 1-GPU box):
  gpurun --gpus 2 -- python -m pytest tests/test_tp.py -m gpu -q
"""
import os
import subprocess
import sys
import tempfile

import pytest

from helpers import C

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _launch(tp, case, prec, timeout=600, p2p=True, two_shot=False, extra_env=None, expect=None):
    """p2p: the fused partial-sum + all-reduce kernel over NVLink peer memory (default);
    False forces the NCCL all-reduce path (NVLLM_TP_P2P=0)."""
    with tempfile.TemporaryDirectory() as td:
        id_file = os.path.join(td, "tp_id")
        procs = []
        env = dict(os.environ, NVLLM_TP_P2P="1" if p2p else "0", NVLLM_TP_DEBUG="1",
                   NVLLM_TP_TWOSHOT="1" if two_shot else "0", **(extra_env or {}))
        for r in range(tp):
            procs.append(subprocess.Popen(
                [sys.executable, os.path.join(HERE, "tp_worker.py"), str(r), str(tp), id_file, case,
                 str(prec)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
        outs = []
        try:
            for p in procs:
                out, _ = p.communicate(timeout=timeout)
                outs.append(out)
        finally:
            for p in procs:
                if p.poll() is None:
                    p.kill()
        for r, (p, out) in enumerate(zip(procs, outs)):
            assert p.returncode == 0 and "TP_OK" in out, f"rank {r} failed:\n{out[-3000:]}"
            if p2p:
                assert "peer-memory all-reduce enabled" in out, f"rank {r} fell back to NCCL:\n{out[-2000:]}"
            if expect:
                assert expect in out, f"rank {r}: '{expect}' not in the output:\n{out[-2000:]}"
    return outs


@pytest.mark.parametrize("case,prec", [
    ("falcon", C.PREC_FP32), ("falcon", C.PREC_BF16), ("falcon_g18", C.PREC_BF16),
    ("llama", C.PREC_FP32), ("llama", C.PREC_BF16), ("gpt2", C.PREC_FP32)])
def test_tp2_matches_oracle(case, prec):
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(2, case, prec)


@pytest.mark.parametrize("case,prec", [("falcon", C.PREC_FP32), ("llama", C.PREC_BF16)])
def test_tp2_nccl_path_matches_oracle(case, prec):
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(2, case, prec, p2p=False)


@pytest.mark.parametrize("case,prec", [("falcon", C.PREC_FP32), ("llama", C.PREC_BF16)])
def test_tp2_two_shot_peer_kernel_matches_oracle(case, prec):
    """The reduce-scatter + all-gather kernel over peer memory (opt-in: NVLLM_TP_TWOSHOT=1)."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(2, case, prec, two_shot=True)


@pytest.mark.parametrize("case,prec", [("falcon", C.PREC_BF16), ("llama", C.PREC_FP32)])
def test_tp4_matches_oracle(case, prec):
    if _n_gpus() < 4:
        pytest.skip("needs 4 GPUs")
    _launch(4, case, prec)


@pytest.mark.parametrize("tp", [2, 4])
@pytest.mark.parametrize("case,prec", [("falcon", C.PREC_FP32), ("falcon_g18", C.PREC_BF16), ("llama", C.PREC_BF16)])
def test_tp_two_chunk_prefill_pipeline_matches_oracle(tp, case, prec):
    """One long sequence: the prefill runs as two token chunks on two streams / communicators
    (forward.cu forward_step_auto); the second chunk's attention reads the first chunk's K / V
    through the per-layer events.  Parallel block (falcon), 71->72-style head padding (7 -> 8),
    sequential blocks with two reductions per layer (llama), SIMT and tcgen05 attention."""
    if _n_gpus() < tp:
        pytest.skip(f"needs {tp} GPUs")
    _launch(tp, case, prec, extra_env={"TP_SINGLE": "1", "NVLLM_TP_PIPE": "1", "NVLLM_TP_PIPE_MIN_T": "64"},
            expect="two-chunk prefill pipeline")
