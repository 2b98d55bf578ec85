"""Test infrastructure: write a HuggingFace-style safetensors checkpoint (single file or sharded
with model.safetensors.index.json) from weights held in the reference's POST-LOAD layouts — the
inverse of what purego/tensor/generic_loader.go:184-604 does on load: PyTorch families get their
linears transposed back to [out, in], GPT-2 gets its fused Conv1D c_attn [H, 3H] and bias [3H]
(splitGPT2QKV :674-702 inverted), Falcon its fused query_key_value [(n_h + 2) hd, H]
(splitFalconQKV :705-748 + combineMQAKV :751-765 inverted), Llama separate gate_proj / up_proj
(:531-536 inverted), Granite-MoE the router as [E, H] (:575 inverted).  The format itself:
8-byte little-endian header length, JSON header {name: {dtype, shape, data_offsets}}, raw data.
"""
import json
import os
import struct

import numpy as np

from helpers import C, W


def _bytes(a, dtype):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if dtype == "F32":
        return a.tobytes()
    if dtype == "F16":
        return a.astype(np.float16).tobytes()
    if dtype == "BF16":
        return W.to_bf16_bits(a).tobytes()
    raise ValueError(dtype)


def write_safetensors(path, tensors, dtype="F32", metadata=True):
    """tensors: {name: fp32 ndarray}; every tensor is stored as `dtype`."""
    header, blobs, off = {}, [], 0
    if metadata:
        header["__metadata__"] = {"format": "pt"}
    for name, a in tensors.items():
        b = _bytes(a, dtype)
        header[name] = {"dtype": dtype, "shape": list(a.shape), "data_offsets": [off, off + len(b)]}
        blobs.append(b)
        off += len(b)
    h = json.dumps(header, separators=(",", ":")).encode()
    h += b" " * ((8 - len(h) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(h)))
        f.write(h)
        for b in blobs:
            f.write(b)


def hf_tensors(cfg, weights, arch, with_unloaded_mlp_bias=False):
    """{HF tensor name: fp32 ndarray} for one of the reference's four mappings."""
    by = {(wid, layer): a for wid, layer, a in weights}
    H, F = cfg.hidden, cfg.ffn_dim
    out = {}
    T = lambda a: np.ascontiguousarray(a.T)
    if arch == C.ARCH_GPT2:
        out["wte.weight"] = by[(C.W_TOKEN_EMBEDDING, -1)]
        out["wpe.weight"] = by[(C.W_POS_EMBEDDING, -1)]
        out["ln_f.weight"] = by[(C.W_FINAL_NORM_W, -1)]
        out["ln_f.bias"] = by[(C.W_FINAL_NORM_B, -1)]
        for l in range(cfg.num_layers):
            p = f"h.{l}"
            out[p + ".ln_1.weight"] = by[(C.W_ATTN_NORM_W, l)]
            out[p + ".ln_1.bias"] = by[(C.W_ATTN_NORM_B, l)]
            out[p + ".ln_2.weight"] = by[(C.W_FFN_NORM_W, l)]
            out[p + ".ln_2.bias"] = by[(C.W_FFN_NORM_B, l)]
            out[p + ".attn.c_attn.weight"] = np.concatenate(
                [by[(C.W_Q, l)], by[(C.W_K, l)], by[(C.W_V, l)]], axis=1)
            out[p + ".attn.c_attn.bias"] = np.concatenate(
                [by[(C.W_Q_BIAS, l)], by[(C.W_K_BIAS, l)], by[(C.W_V_BIAS, l)]])
            out[p + ".attn.c_proj.weight"] = by[(C.W_O, l)]
            out[p + ".attn.c_proj.bias"] = by[(C.W_O_BIAS, l)]
            out[p + ".mlp.c_fc.weight"] = by[(C.W_FFN_W1, l)]
            out[p + ".mlp.c_proj.weight"] = by[(C.W_FFN_W2, l)]
            if with_unloaded_mlp_bias:
                # real GPT-2 checkpoints have these; the reference builds the key
                # "...c_fc.weight.bias" and never finds them (generic_loader.go:559-560)
                out[p + ".mlp.c_fc.bias"] = np.full(F, 0.5, np.float32)
                out[p + ".mlp.c_proj.bias"] = np.full(H, 0.5, np.float32)
        return out
    if arch == C.ARCH_FALCON:
        hd = cfg.head_dim
        out["transformer.word_embeddings.weight"] = by[(C.W_TOKEN_EMBEDDING, -1)]
        out["transformer.ln_f.weight"] = by[(C.W_FINAL_NORM_W, -1)]
        out["transformer.ln_f.bias"] = by[(C.W_FINAL_NORM_B, -1)]
        if not cfg.tied_embedding:
            out["lm_head.weight"] = T(by[(C.W_LM_HEAD, -1)])
        for l in range(cfg.num_layers):
            p = f"transformer.h.{l}"
            kv = by[(C.W_KV, l)]
            out[p + ".self_attention.query_key_value.weight"] = np.concatenate(
                [T(by[(C.W_Q, l)]), T(kv[:, :hd]), T(kv[:, hd:])], axis=0)
            out[p + ".self_attention.dense.weight"] = T(by[(C.W_O, l)])
            out[p + ".mlp.dense_h_to_4h.weight"] = T(by[(C.W_FFN_W1, l)])
            out[p + ".mlp.dense_4h_to_h.weight"] = T(by[(C.W_FFN_W2, l)])
            out[p + ".input_layernorm.weight"] = by[(C.W_ATTN_NORM_W, l)]
            out[p + ".input_layernorm.bias"] = by[(C.W_ATTN_NORM_B, l)]
        return out
    # Llama / Granite-MoE
    out["model.embed_tokens.weight"] = by[(C.W_TOKEN_EMBEDDING, -1)]
    out["model.norm.weight"] = by[(C.W_FINAL_NORM_W, -1)]
    if not cfg.tied_embedding and arch == C.ARCH_LLAMA:
        out["lm_head.weight"] = T(by[(C.W_LM_HEAD, -1)])
    for l in range(cfg.num_layers):
        p = f"model.layers.{l}"
        out[p + ".input_layernorm.weight"] = by[(C.W_ATTN_NORM_W, l)]
        out[p + ".post_attention_layernorm.weight"] = by[(C.W_FFN_NORM_W, l)]
        out[p + ".self_attn.q_proj.weight"] = T(by[(C.W_Q, l)])
        out[p + ".self_attn.k_proj.weight"] = T(by[(C.W_K, l)])
        out[p + ".self_attn.v_proj.weight"] = T(by[(C.W_V, l)])
        out[p + ".self_attn.o_proj.weight"] = T(by[(C.W_O, l)])
        if arch == C.ARCH_GRANITE_MOE:
            out[p + ".block_sparse_moe.router.layer.weight"] = T(by[(C.W_MOE_ROUTER, l)])
            out[p + ".block_sparse_moe.input_linear.weight"] = by[(C.W_MOE_INPUT, l)]
            out[p + ".block_sparse_moe.output_linear.weight"] = by[(C.W_MOE_OUTPUT, l)]
        else:
            w1 = by[(C.W_FFN_W1, l)]
            out[p + ".mlp.gate_proj.weight"] = T(w1[:, :F])
            out[p + ".mlp.up_proj.weight"] = T(w1[:, F:])
            out[p + ".mlp.down_proj.weight"] = T(by[(C.W_FFN_W2, l)])
    return out


def export_checkpoint(model_dir, cfg, weights, arch, dtype="F32", shards=1, **kw):
    """Write model.safetensors, or `shards` files + model.safetensors.index.json."""
    os.makedirs(model_dir, exist_ok=True)
    tensors = hf_tensors(cfg, list(weights), arch, **kw)
    if shards <= 1:
        write_safetensors(os.path.join(model_dir, "model.safetensors"), tensors, dtype)
        return tensors
    names = list(tensors)
    weight_map = {}
    for s in range(shards):
        fn = f"model-{s + 1:05d}-of-{shards:05d}.safetensors"
        part = {n: tensors[n] for n in names[s::shards]}
        write_safetensors(os.path.join(model_dir, fn), part, dtype)
        for n in part:
            weight_map[n] = fn
    with open(os.path.join(model_dir, "model.safetensors.index.json"), "w") as f:
        json.dump({"metadata": {"total_size": 0}, "weight_map": weight_map}, f, indent=1)
    return tensors
