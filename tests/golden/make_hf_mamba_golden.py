#!/usr/bin/env python
"""Golden vector for the Mamba2 block: one GraniteMoeHybridMambaLayer of
HF transformers (the implementation Granite-4 checkpoints are published for), pure-PyTorch path
(`torch_forward`, chunked SSD scan), fp32, on seeded weights and a seeded input.

    python tests/golden/make_hf_mamba_golden.py     ->  tests/golden/hf_mamba2.npz

A THIRD-PARTY cross-check of oracle/nvllm_oracle.c::mamba_forward (which restates the reference's
purego/tensor/mamba2.go:72-351), not a reference-held vector: the reference has no test for this
layer and cannot be built here (no Go toolchain).  It pins the math of ONE forward call from an
empty state — in_proj split, causal depthwise conv + SiLU, softplus(dt + bias), the selective
scan with grouped B / C, the D skip, the gated RMS norm over the whole inner width (eps 1e-5)
and out_proj.  What the reference does differently from HF is outside this vector on purpose:
it re-pads the convolution on every call (never reads its ConvCache), which only shows on decode
steps; tests/test_oracle.py covers that against the numpy restatement.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import nvllm_pkg  # noqa: E402

C = nvllm_pkg.sub("config")
W = nvllm_pkg.sub("weights")

# must match tests/test_oracle.py::MAMBA_GOLDEN_CFG
DIMS = dict(hidden=64, num_layers=1, mamba_num_heads=8, mamba_head_dim=16, mamba_state_size=16,
            mamba_n_groups=2, mamba_conv_kernel=4, hybrid_layers=(True,), vocab_size=64,
            num_heads=2, num_kv_heads=1, head_dim=32, ffn_dim=64)
SEED, T, CHUNK = 20261105, 21, 8   # 21 tokens over chunks of 8: two full chunks + a padded one


def main():
    from transformers.models.granitemoehybrid.configuration_granitemoehybrid import GraniteMoeHybridConfig
    from transformers.models.granitemoehybrid.modeling_granitemoehybrid import GraniteMoeHybridMambaLayer

    cfg = C.tiny_hybrid(**DIMS)
    E, conv, rows = cfg.mamba_dims()
    ws = {wid: np.asarray(a, dtype=np.float32) for wid, l, a in W.random_weights(cfg, SEED) if l == 0}
    hf_cfg = GraniteMoeHybridConfig(
        vocab_size=cfg.vocab_size, hidden_size=cfg.hidden, num_hidden_layers=1,
        num_attention_heads=cfg.num_heads, num_key_value_heads=cfg.num_kv_heads,
        intermediate_size=cfg.ffn_dim, shared_intermediate_size=cfg.ffn_dim,
        mamba_n_heads=cfg.mamba_num_heads, mamba_d_head=cfg.mamba_head_dim,
        mamba_d_state=cfg.mamba_state_size, mamba_n_groups=cfg.mamba_n_groups,
        mamba_d_conv=cfg.mamba_conv_kernel, mamba_expand=E // cfg.hidden, mamba_chunk_size=CHUNK,
        mamba_conv_bias=True, mamba_proj_bias=False, rms_norm_eps=1e-5, layer_types=["mamba"],
        hidden_act="silu")
    assert int(hf_cfg.mamba_expand * hf_cfg.hidden_size) == E
    torch.manual_seed(0)
    layer = GraniteMoeHybridMambaLayer(hf_cfg, 0).float().eval()
    with torch.no_grad():
        layer.in_proj.weight.copy_(torch.from_numpy(ws[C.W_MAMBA_IN_PROJ]))            # [rows, H]
        layer.conv1d.weight.copy_(torch.from_numpy(ws[C.W_MAMBA_CONV_W]))              # [conv, 1, K]
        layer.conv1d.bias.copy_(torch.from_numpy(ws[C.W_MAMBA_CONV_B]))
        layer.dt_bias.copy_(torch.from_numpy(ws[C.W_MAMBA_DT_BIAS]))
        layer.A_log.copy_(torch.from_numpy(ws[C.W_MAMBA_A_LOG]))
        layer.D.copy_(torch.from_numpy(ws[C.W_MAMBA_D]))
        layer.norm.weight.copy_(torch.from_numpy(ws[C.W_MAMBA_NORM]))
        layer.out_proj.weight.copy_(torch.from_numpy(ws[C.W_MAMBA_OUT_PROJ]))          # [H, E]
        assert tuple(layer.in_proj.weight.shape) == (rows, cfg.hidden)
        x = np.random.default_rng(SEED).standard_normal((T, cfg.hidden)).astype(np.float32)
        y = layer.torch_forward(torch.from_numpy(x)[None], None, None)[0].numpy()
    out = os.path.join(ROOT, "tests", "golden", "hf_mamba2.npz")
    np.savez_compressed(out, x=x, y=y.astype(np.float32), seed=np.int64(SEED),
                        **{f"w{wid}": ws[wid] for wid in (C.W_MAMBA_IN_PROJ, C.W_MAMBA_CONV_W,
                                                         C.W_MAMBA_CONV_B, C.W_MAMBA_DT_BIAS,
                                                         C.W_MAMBA_A_LOG, C.W_MAMBA_D,
                                                         C.W_MAMBA_NORM, C.W_MAMBA_OUT_PROJ)})
    print("wrote", out, "y abs max", float(np.abs(y).max()), "bytes", os.path.getsize(out))


if __name__ == "__main__":
    main()
