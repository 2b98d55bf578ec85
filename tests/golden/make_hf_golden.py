"""Generate golden last-token logits from HuggingFace transformers (CPU, fp32).

The reference's own tests pin no forward-pass numerics, and the
reference (Go) cannot run in this image.  As an independent cross-check of the
oracle we load OUR deterministic random-init weights (nano-vllm-go_b200/weights.py,
reference post-load layouts) into the matching HF architecture by inverting the
reference loader's transforms (purego/tensor/generic_loader.go:353-604,674-765),
run HF's forward on a fixed prompt, and store the logits of every position.

Run here (needs transformers; not needed on the GPU box):
    python tests/golden/make_hf_golden.py
Writes tests/golden/hf_<arch>.npz  (config name, seed, prompt, logits[T,V]).

Known, documented difference: HF Falcon uses erf-GELU by default while the
reference uses the tanh approximation for every model (tensor.go:181-190); the
golden is generated with activation="gelu_pytorch_tanh" so that it checks the
rest of the pipeline (parity target is the reference, not HF).
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

SEED = 1234
PROMPT_LEN = 24


def wdict(cfg):
    return {(wid, l): torch.from_numpy(a.copy()) for wid, l, a in W.random_weights(cfg, SEED)}


def build_gpt2(cfg, w):
    from transformers import GPT2Config, GPT2LMHeadModel
    hc = GPT2Config(vocab_size=cfg.vocab_size, n_positions=cfg.max_seq_len, n_embd=cfg.hidden,
                    n_layer=cfg.num_layers, n_head=cfg.num_heads, n_inner=cfg.ffn_dim,
                    activation_function="gelu_new", layer_norm_epsilon=cfg.norm_eps,
                    resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    m = GPT2LMHeadModel(hc).eval()
    sd = {"transformer.wte.weight": w[(C.W_TOKEN_EMBEDDING, -1)],
          "transformer.wpe.weight": w[(C.W_POS_EMBEDDING, -1)],
          "transformer.ln_f.weight": w[(C.W_FINAL_NORM_W, -1)],
          "transformer.ln_f.bias": w[(C.W_FINAL_NORM_B, -1)],
          "lm_head.weight": w[(C.W_TOKEN_EMBEDDING, -1)]}
    for l in range(cfg.num_layers):
        p = f"transformer.h.{l}."
        sd[p + "ln_1.weight"] = w[(C.W_ATTN_NORM_W, l)]
        sd[p + "ln_1.bias"] = w[(C.W_ATTN_NORM_B, l)]
        sd[p + "ln_2.weight"] = w[(C.W_FFN_NORM_W, l)]
        sd[p + "ln_2.bias"] = w[(C.W_FFN_NORM_B, l)]
        # splitGPT2QKV inverse: c_attn.weight [H, 3H] = [Q | K | V] by columns
        sd[p + "attn.c_attn.weight"] = torch.cat([w[(C.W_Q, l)], w[(C.W_K, l)], w[(C.W_V, l)]], 1)
        sd[p + "attn.c_attn.bias"] = torch.cat(
            [w[(C.W_Q_BIAS, l)], w[(C.W_K_BIAS, l)], w[(C.W_V_BIAS, l)]])
        sd[p + "attn.c_proj.weight"] = w[(C.W_O, l)]
        sd[p + "attn.c_proj.bias"] = w[(C.W_O_BIAS, l)]
        sd[p + "mlp.c_fc.weight"] = w[(C.W_FFN_W1, l)]
        sd[p + "mlp.c_fc.bias"] = w[(C.W_FFN_B1, l)]
        sd[p + "mlp.c_proj.weight"] = w[(C.W_FFN_W2, l)]
        sd[p + "mlp.c_proj.bias"] = w[(C.W_FFN_B2, l)]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("attn.bias" in k or "masked_bias" in k for k in missing), missing
    return m


def _llama_like_sd(cfg, w, moe=False):
    F = cfg.ffn_dim
    sd = {"model.embed_tokens.weight": w[(C.W_TOKEN_EMBEDDING, -1)],
          "model.norm.weight": w[(C.W_FINAL_NORM_W, -1)],
          "lm_head.weight": w[(C.W_TOKEN_EMBEDDING, -1)]}
    for l in range(cfg.num_layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = w[(C.W_ATTN_NORM_W, l)]
        sd[p + "post_attention_layernorm.weight"] = w[(C.W_FFN_NORM_W, l)]
        sd[p + "self_attn.q_proj.weight"] = w[(C.W_Q, l)].t().contiguous()
        sd[p + "self_attn.k_proj.weight"] = w[(C.W_K, l)].t().contiguous()
        sd[p + "self_attn.v_proj.weight"] = w[(C.W_V, l)].t().contiguous()
        sd[p + "self_attn.o_proj.weight"] = w[(C.W_O, l)].t().contiguous()
        if moe:
            sd[p + "block_sparse_moe.router.layer.weight"] = w[(C.W_MOE_ROUTER, l)].t().contiguous()
            sd[p + "block_sparse_moe.input_linear.weight"] = w[(C.W_MOE_INPUT, l)]
            sd[p + "block_sparse_moe.output_linear.weight"] = w[(C.W_MOE_OUTPUT, l)]
        else:
            w1 = w[(C.W_FFN_W1, l)]
            sd[p + "mlp.gate_proj.weight"] = w1[:, :F].t().contiguous()
            sd[p + "mlp.up_proj.weight"] = w1[:, F:].t().contiguous()
            sd[p + "mlp.down_proj.weight"] = w[(C.W_FFN_W2, l)].t().contiguous()
    return sd


def build_llama(cfg, w):
    from transformers import LlamaConfig, LlamaForCausalLM
    hc = LlamaConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden,
                     intermediate_size=cfg.ffn_dim, num_hidden_layers=cfg.num_layers,
                     num_attention_heads=cfg.num_heads, num_key_value_heads=cfg.num_kv_heads,
                     head_dim=cfg.head_dim, max_position_embeddings=cfg.max_seq_len,
                     rms_norm_eps=cfg.norm_eps, rope_theta=cfg.rope_base, rope_scaling=None,
                     tie_word_embeddings=True, attention_bias=False, mlp_bias=False)
    m = LlamaForCausalLM(hc).eval()
    missing, unexpected = m.load_state_dict(_llama_like_sd(cfg, w), strict=False)
    assert not unexpected and not [k for k in missing if "rotary" not in k], (missing, unexpected)
    return m


def build_granite(cfg, w):
    from transformers import GraniteMoeConfig, GraniteMoeForCausalLM
    hc = GraniteMoeConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden,
                          intermediate_size=cfg.ffn_dim, num_hidden_layers=cfg.num_layers,
                          num_attention_heads=cfg.num_heads, num_key_value_heads=cfg.num_kv_heads,
                          max_position_embeddings=cfg.max_seq_len, rms_norm_eps=cfg.norm_eps,
                          rope_theta=cfg.rope_base, rope_scaling=None, tie_word_embeddings=True,
                          attention_bias=False, num_local_experts=cfg.num_experts,
                          num_experts_per_tok=cfg.num_experts_per_tok,
                          embedding_multiplier=cfg.embedding_multiplier,
                          attention_multiplier=cfg.attention_multiplier,
                          residual_multiplier=cfg.residual_multiplier,
                          logits_scaling=cfg.logits_scaling, attention_dropout=0.0)
    m = GraniteMoeForCausalLM(hc).eval()
    missing, unexpected = m.load_state_dict(_llama_like_sd(cfg, w, moe=True), strict=False)
    assert not unexpected and not [k for k in missing if "rotary" not in k], (missing, unexpected)
    return m


def build_falcon(cfg, w):
    from transformers import FalconConfig, FalconForCausalLM
    hc = FalconConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden,
                      num_hidden_layers=cfg.num_layers, num_attention_heads=cfg.num_heads,
                      layer_norm_epsilon=cfg.norm_eps, hidden_dropout=0.0, attention_dropout=0.0,
                      multi_query=True, parallel_attn=True, new_decoder_architecture=False,
                      bias=False, alibi=False, ffn_hidden_size=cfg.ffn_dim,
                      max_position_embeddings=cfg.max_seq_len, rope_theta=10000.0,
                      activation="gelu_pytorch_tanh", tie_word_embeddings=False)
    m = FalconForCausalLM(hc).eval()
    sd = {"transformer.word_embeddings.weight": w[(C.W_TOKEN_EMBEDDING, -1)],
          "transformer.ln_f.weight": w[(C.W_FINAL_NORM_W, -1)],
          "transformer.ln_f.bias": w[(C.W_FINAL_NORM_B, -1)],
          "lm_head.weight": w[(C.W_LM_HEAD, -1)].t().contiguous()}
    hd = cfg.head_dim
    for l in range(cfg.num_layers):
        p = f"transformer.h.{l}."
        sd[p + "input_layernorm.weight"] = w[(C.W_ATTN_NORM_W, l)]
        sd[p + "input_layernorm.bias"] = w[(C.W_ATTN_NORM_B, l)]
        kv = w[(C.W_KV, l)]
        # splitFalconQKV inverse: fused [H, (nh+2)*hd] = [Q heads | K | V], stored [out, in]
        fused = torch.cat([w[(C.W_Q, l)], kv[:, :hd], kv[:, hd:]], 1)
        sd[p + "self_attention.query_key_value.weight"] = fused.t().contiguous()
        sd[p + "self_attention.dense.weight"] = w[(C.W_O, l)].t().contiguous()
        sd[p + "mlp.dense_h_to_4h.weight"] = w[(C.W_FFN_W1, l)].t().contiguous()
        sd[p + "mlp.dense_4h_to_h.weight"] = w[(C.W_FFN_W2, l)].t().contiguous()
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and not [k for k in missing if "rotary" not in k], (missing, unexpected)
    return m


def build_hybrid(cfg, w):
    """Granite-4 hybrid: Mamba2 and attention blocks, shared SwiGLU MLP, no positional encoding,
    muP multipliers (reference: generic_model.go:162-202, mamba2.go)."""
    from transformers.models.granitemoehybrid.configuration_granitemoehybrid import GraniteMoeHybridConfig
    from transformers.models.granitemoehybrid.modeling_granitemoehybrid import GraniteMoeHybridForCausalLM
    E, conv, rows = cfg.mamba_dims()
    hc = GraniteMoeHybridConfig(
        vocab_size=cfg.vocab_size, hidden_size=cfg.hidden, intermediate_size=cfg.ffn_dim,
        shared_intermediate_size=cfg.ffn_dim, num_hidden_layers=cfg.num_layers,
        num_attention_heads=cfg.num_heads, num_key_value_heads=cfg.num_kv_heads,
        max_position_embeddings=cfg.max_seq_len, rms_norm_eps=cfg.norm_eps, tie_word_embeddings=True,
        attention_bias=False, attention_dropout=0.0, num_local_experts=0, num_experts_per_tok=0,
        embedding_multiplier=cfg.embedding_multiplier, attention_multiplier=cfg.attention_multiplier,
        residual_multiplier=cfg.residual_multiplier, logits_scaling=cfg.logits_scaling,
        position_embedding_type="nope", hidden_act="silu",
        layer_types=["mamba" if t else "attention" for t in cfg.hybrid_layers],
        mamba_n_heads=cfg.mamba_num_heads, mamba_d_head=cfg.mamba_head_dim,
        mamba_d_state=cfg.mamba_state_size, mamba_n_groups=cfg.mamba_n_groups,
        mamba_d_conv=cfg.mamba_conv_kernel, mamba_expand=E // cfg.hidden, mamba_chunk_size=8,
        mamba_conv_bias=True, mamba_proj_bias=False)
    m = GraniteMoeHybridForCausalLM(hc).float().eval()
    sd = {"model.embed_tokens.weight": w[(C.W_TOKEN_EMBEDDING, -1)],
          "model.norm.weight": w[(C.W_FINAL_NORM_W, -1)],
          "lm_head.weight": w[(C.W_TOKEN_EMBEDDING, -1)]}
    for l in range(cfg.num_layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = w[(C.W_ATTN_NORM_W, l)]
        sd[p + "post_attention_layernorm.weight"] = w[(C.W_FFN_NORM_W, l)]
        sd[p + "shared_mlp.input_linear.weight"] = w[(C.W_FFN_W1, l)].t().contiguous()   # [2F, H]: gate | up
        sd[p + "shared_mlp.output_linear.weight"] = w[(C.W_FFN_W2, l)].t().contiguous()
        if cfg.is_mamba(l):
            sd[p + "mamba.in_proj.weight"] = w[(C.W_MAMBA_IN_PROJ, l)]
            sd[p + "mamba.conv1d.weight"] = w[(C.W_MAMBA_CONV_W, l)]
            sd[p + "mamba.conv1d.bias"] = w[(C.W_MAMBA_CONV_B, l)]
            sd[p + "mamba.dt_bias"] = w[(C.W_MAMBA_DT_BIAS, l)]
            sd[p + "mamba.A_log"] = w[(C.W_MAMBA_A_LOG, l)]
            sd[p + "mamba.D"] = w[(C.W_MAMBA_D, l)]
            sd[p + "mamba.norm.weight"] = w[(C.W_MAMBA_NORM, l)]
            sd[p + "mamba.out_proj.weight"] = w[(C.W_MAMBA_OUT_PROJ, l)]
        else:
            sd[p + "self_attn.q_proj.weight"] = w[(C.W_Q, l)].t().contiguous()
            sd[p + "self_attn.k_proj.weight"] = w[(C.W_K, l)].t().contiguous()
            sd[p + "self_attn.v_proj.weight"] = w[(C.W_V, l)].t().contiguous()
            sd[p + "self_attn.o_proj.weight"] = w[(C.W_O, l)].t().contiguous()
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and not [k for k in missing if "rotary" not in k], (missing, unexpected)
    return m


CASES = {
    "gpt2": (lambda: C.tiny_gpt2(vocab_size=320, hidden=64, head_dim=16, ffn_dim=256,
                                 max_seq_len=64), build_gpt2),
    "llama": (lambda: C.tiny_llama(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                   head_dim=16, ffn_dim=128, max_seq_len=64), build_llama),
    "granite": (lambda: C.tiny_granite(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                       head_dim=16, ffn_dim=32, max_seq_len=64), build_granite),
    "falcon": (lambda: C.tiny_falcon(vocab_size=320, hidden=5 * 16, num_heads=5, head_dim=16,
                                     ffn_dim=4 * 5 * 16, max_seq_len=64), build_falcon),
    "hybrid": (lambda: C.tiny_hybrid(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                     head_dim=16, ffn_dim=96, max_seq_len=64, mamba_num_heads=8,
                                     mamba_head_dim=16, mamba_state_size=16, mamba_n_groups=2,
                                     mamba_conv_kernel=4, norm_eps=1e-5), build_hybrid),
}


def main():
    torch.manual_seed(0)
    out_dir = os.path.dirname(os.path.abspath(__file__))
    only = sys.argv[1:]  # optional: the cases to (re)generate
    for name, (mk, build) in CASES.items():
        if only and name not in only:
            continue
        cfg = mk()
        w = wdict(cfg)
        model = build(cfg, w)
        prompt = W.random_prompt(cfg.vocab_size, PROMPT_LEN, 4321)
        with torch.no_grad():
            logits = model(torch.from_numpy(prompt.astype(np.int64))[None]).logits[0].float().numpy()
        np.savez_compressed(os.path.join(out_dir, f"hf_{name}.npz"), prompt=prompt,
                            logits=logits.astype(np.float32), seed=np.int64(SEED))
        print(name, cfg.name, logits.shape, float(np.abs(logits).max()))


if __name__ == "__main__":
    main()
