"""Random-init weights in the reference's POST-LOAD layouts (synthetic data).

Layouts are what ``TransformerModel`` holds after
/root/reference purego/tensor/generic_loader.go:184-265,353-604: every linear
is ``[in, out]`` row-major; Llama W1 is ``[gate | up]`` (``:531-536``); Falcon
KV is ``[H, 2*hd] = [K | V]`` (``:750-765``); MoE expert tensors keep the HF
layout ``[E, 2I, H]`` / ``[E, H, I]`` (``:578-586``); router is ``[H, E]``.

Distributions follow SURVEY.md §8d: linear / embedding ~ N(0, 0.02^2), output
projections ~ N(0, (0.02/sqrt(2L))^2), norm weight 1 + N(0, 0.02^2), biases
N(0, 0.02^2).  Each tensor has its own seed so tensors can be generated lazily,
one at a time, identically for the CUDA library and the CPU oracle.
"""
from __future__ import annotations

import math
from typing import Iterator, Tuple

import numpy as np
import torch

from . import config as C


def bf16_round(a: np.ndarray) -> np.ndarray:
    """fp32 -> nearest-even bf16 -> fp32 (what the reference loader sees for BF16 files,
    generic_loader.go:656-660,802-805)."""
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return t.to(torch.bfloat16).to(torch.float32).numpy()


def to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """fp32 array -> uint16 array of bf16 bit patterns (RNE)."""
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.bfloat16)
    return t.view(torch.int16).numpy().view(np.uint16)


def _normal(shape, std, seed, mean=0.0) -> np.ndarray:
    g = torch.Generator()
    g.manual_seed(seed)
    t = torch.empty(shape, dtype=torch.float32)
    t.normal_(mean, std, generator=g)
    return t.numpy()


def weight_specs(cfg: C.ModelConfig):
    """Yield (weight_id, layer, shape, std, mean) for every tensor the architecture has."""
    H, V, L = cfg.hidden, cfg.vocab_size, cfg.num_layers
    q = cfg.num_heads * cfg.head_dim
    kv = cfg.kv_heads() * cfg.head_dim
    F = cfg.ffn_dim
    s = 0.02
    so = 0.02 / math.sqrt(2 * L)
    layer_norm = cfg.norm_type == C.NORM_LAYER
    yield C.W_TOKEN_EMBEDDING, -1, (V, H), s, 0.0
    if cfg.position_type == C.POS_LEARNED:
        yield C.W_POS_EMBEDDING, -1, (cfg.max_seq_len, H), s, 0.0
    yield C.W_FINAL_NORM_W, -1, (H,), s, 1.0
    if layer_norm:
        yield C.W_FINAL_NORM_B, -1, (H,), s, 0.0
    if not cfg.tied_embedding:
        yield C.W_LM_HEAD, -1, (H, V), s, 0.0
    for l in range(L):
        yield C.W_ATTN_NORM_W, l, (H,), s, 1.0
        if layer_norm:
            yield C.W_ATTN_NORM_B, l, (H,), s, 0.0
        if cfg.block_style == C.BLOCK_SEQUENTIAL:
            yield C.W_FFN_NORM_W, l, (H,), s, 1.0
            if layer_norm:
                yield C.W_FFN_NORM_B, l, (H,), s, 0.0
        if cfg.is_mamba(l):
            # Mamba2 block (mamba2.go:9-28; tensors as the loader leaves them, un-transposed):
            # norms + the layer + the shared MLP, no attention tensors
            E, conv, rows = cfg.mamba_dims()
            nhm = cfg.mamba_num_heads
            yield C.W_MAMBA_IN_PROJ, l, (rows, H), s, 0.0
            yield C.W_MAMBA_CONV_W, l, (conv, 1, cfg.mamba_conv_kernel), 0.3, 0.0
            yield C.W_MAMBA_CONV_B, l, (conv,), s, 0.0
            yield C.W_MAMBA_DT_BIAS, l, (nhm,), 0.5, 0.0
            yield C.W_MAMBA_A_LOG, l, (nhm,), 0.5, 0.0
            yield C.W_MAMBA_D, l, (nhm,), s, 1.0
            yield C.W_MAMBA_NORM, l, (E,), s, 1.0
            yield C.W_MAMBA_OUT_PROJ, l, (H, E), so, 0.0
            w1_out = 2 * F if cfg.activation_type == C.ACT_SWIGLU else F
            yield C.W_FFN_W1, l, (H, w1_out), s, 0.0
            yield C.W_FFN_W2, l, (F, H), so, 0.0
            continue
        yield C.W_Q, l, (H, q), s, 0.0
        if cfg.attention_type == C.ATTN_MQA:
            yield C.W_KV, l, (H, 2 * cfg.head_dim), s, 0.0
        else:
            yield C.W_K, l, (H, kv), s, 0.0
            yield C.W_V, l, (H, kv), s, 0.0
        yield C.W_O, l, (q, H), so, 0.0
        if cfg.attention_type == C.ATTN_MHA:  # GPT-2 biases (generic_loader.go:417-447)
            yield C.W_Q_BIAS, l, (q,), s, 0.0
            yield C.W_K_BIAS, l, (kv,), s, 0.0
            yield C.W_V_BIAS, l, (kv,), s, 0.0
            yield C.W_O_BIAS, l, (H,), s, 0.0
        if cfg.use_moe:
            E = cfg.num_experts
            yield C.W_MOE_ROUTER, l, (H, E), s, 0.0
            yield C.W_MOE_INPUT, l, (E, 2 * F, H), s, 0.0
            yield C.W_MOE_OUTPUT, l, (E, H, F), so, 0.0
        else:
            w1_out = 2 * F if cfg.activation_type == C.ACT_SWIGLU else F
            yield C.W_FFN_W1, l, (H, w1_out), s, 0.0
            yield C.W_FFN_W2, l, (F, H), so, 0.0
            if cfg.has_ffn_bias and cfg.activation_type == C.ACT_GELU:
                yield C.W_FFN_B1, l, (F,), s, 0.0
                yield C.W_FFN_B2, l, (H,), s, 0.0


def random_weights(cfg: C.ModelConfig, seed: int = 1234,
                   round_bf16: bool = False) -> Iterator[Tuple[int, int, np.ndarray]]:
    """Yield (weight_id, layer, fp32 ndarray) one tensor at a time.

    round_bf16: round every tensor to bf16 (RNE) and return it as fp32 — the
    bf16-mode parity contract feeds these same values to the oracle.
    """
    for wid, layer, shape, std, mean in weight_specs(cfg):
        tseed = (seed * 1000003 + wid * 7919 + (layer + 1) * 104729) & 0x7FFFFFFF
        a = _normal(shape, std, tseed, mean)
        if round_bf16:
            a = bf16_round(a)
        yield wid, layer, a


def random_prompt(vocab: int, length: int, seed: int) -> np.ndarray:
    """Token ids uniform in [0, V) (seed 4321 + sequence index)."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, vocab, size=length, dtype=np.int32)
