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


def _launch(tp, case, prec, timeout=600):
    with tempfile.TemporaryDirectory() as td:
        id_file = os.path.join(td, "tp_id")
        procs = []
        for r in range(tp):
            procs.append(subprocess.Popen(
                [sys.executable, os.path.join(HERE, "tp_worker.py"), str(r), str(tp), id_file, case,
                 str(prec)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
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
    return outs


@pytest.mark.parametrize("case,prec", [
    ("falcon", C.PREC_FP32), ("falcon", C.PREC_BF16), ("falcon_g18", C.PREC_BF16),
    ("llama", C.PREC_FP32), ("llama", C.PREC_BF16), ("gpt2", C.PREC_FP32)])
def test_tp2_matches_oracle(case, prec):
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(2, case, prec)


@pytest.mark.parametrize("case,prec", [("falcon", C.PREC_BF16), ("llama", C.PREC_FP32)])
def test_tp4_matches_oracle(case, prec):
    if _n_gpus() < 4:
        pytest.skip("needs 4 GPUs")
    _launch(4, case, prec)
