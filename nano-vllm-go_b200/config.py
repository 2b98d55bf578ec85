"""Model configuration mirror of the reference's ``tensor.ModelConfig``.

Field meaning follows /root/reference purego/tensor/config.go:58-122; the
presets follow config.go:125-148 (GPT-2), :151-193 (Falcon), :330-376
(Granite-MoE, muP scalars from docs/ARCHITECTURE_GUIDE.md:1773-1776).  The
Llama-3.2 shapes are the HF config.json values the reference reads at load
(generic_loader.go:837-915); they are not in the reference tree.

``to_c()`` produces the ``nvllm_config`` struct of include/nvllm.h.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, replace

ABI_VERSION = 1
KV_BLOCK_SIZE = 256  # nanovllm/sequence.go:52

ATTN_MHA, ATTN_GQA, ATTN_MQA = 0, 1, 2
NORM_LAYER, NORM_RMS = 0, 1
POS_LEARNED, POS_ROPE, POS_NONE = 0, 1, 2
ACT_GELU, ACT_SWIGLU = 0, 1
BLOCK_SEQUENTIAL, BLOCK_PARALLEL = 0, 1
PREC_FP32, PREC_BF16 = 0, 1
# enum nvllm_arch (checkpoint families of generic_loader.go:60-182)
ARCH_GPT2, ARCH_FALCON, ARCH_LLAMA, ARCH_GRANITE_MOE = 0, 1, 2, 3
DT_F32, DT_BF16, DT_F16 = 0, 1, 2

# enum nvllm_weight_id
W_TOKEN_EMBEDDING, W_POS_EMBEDDING, W_FINAL_NORM_W, W_FINAL_NORM_B, W_LM_HEAD = 0, 1, 2, 3, 4
W_ATTN_NORM_W, W_ATTN_NORM_B, W_FFN_NORM_W, W_FFN_NORM_B = 16, 17, 18, 19
(W_MAMBA_IN_PROJ, W_MAMBA_CONV_W, W_MAMBA_CONV_B, W_MAMBA_DT_BIAS, W_MAMBA_A_LOG, W_MAMBA_D,
 W_MAMBA_NORM, W_MAMBA_OUT_PROJ) = 36, 37, 38, 39, 40, 41, 42, 43
W_Q, W_K, W_V, W_KV, W_O = 20, 21, 22, 23, 24
W_Q_BIAS, W_K_BIAS, W_V_BIAS, W_O_BIAS = 25, 26, 27, 28
W_FFN_W1, W_FFN_B1, W_FFN_W2, W_FFN_B2 = 29, 30, 31, 32
W_MOE_ROUTER, W_MOE_INPUT, W_MOE_OUTPUT = 33, 34, 35


class CConfig(ctypes.Structure):
    """``struct nvllm_config`` (include/nvllm.h)."""

    _fields_ = [
        ("abi_version", ctypes.c_int32),
        ("vocab_size", ctypes.c_int32),
        ("hidden", ctypes.c_int32),
        ("num_layers", ctypes.c_int32),
        ("num_heads", ctypes.c_int32),
        ("num_kv_heads", ctypes.c_int32),
        ("head_dim", ctypes.c_int32),
        ("ffn_dim", ctypes.c_int32),
        ("max_seq_len", ctypes.c_int32),
        ("attention_type", ctypes.c_int32),
        ("norm_type", ctypes.c_int32),
        ("position_type", ctypes.c_int32),
        ("activation_type", ctypes.c_int32),
        ("block_style", ctypes.c_int32),
        ("tied_embedding", ctypes.c_int32),
        ("use_moe", ctypes.c_int32),
        ("num_experts", ctypes.c_int32),
        ("num_experts_per_tok", ctypes.c_int32),
        ("norm_eps", ctypes.c_float),
        ("embedding_multiplier", ctypes.c_float),
        ("attention_multiplier", ctypes.c_float),
        ("residual_multiplier", ctypes.c_float),
        ("logits_scaling", ctypes.c_float),
        ("rope_base", ctypes.c_double),
        ("precision", ctypes.c_int32),
        ("num_kv_blocks", ctypes.c_int32),
        ("max_num_seqs", ctypes.c_int32),
        ("max_batched_tokens", ctypes.c_int32),
        ("tp_size", ctypes.c_int32),
        ("tp_rank", ctypes.c_int32),
        ("device", ctypes.c_int32),
        ("mamba_num_heads", ctypes.c_int32),
        ("mamba_head_dim", ctypes.c_int32),
        ("mamba_state_size", ctypes.c_int32),
        ("mamba_n_groups", ctypes.c_int32),
        ("mamba_conv_kernel", ctypes.c_int32),
    ]


@dataclass(frozen=True)
class ModelConfig:
    name: str
    vocab_size: int
    hidden: int
    num_layers: int
    num_heads: int
    num_kv_heads: int
    head_dim: int
    ffn_dim: int
    max_seq_len: int
    attention_type: int
    norm_type: int
    position_type: int
    activation_type: int
    block_style: int
    tied_embedding: bool
    norm_eps: float
    rope_base: float = 10000.0
    use_moe: bool = False
    num_experts: int = 0
    num_experts_per_tok: int = 0
    embedding_multiplier: float = 0.0
    attention_multiplier: float = 0.0
    residual_multiplier: float = 0.0
    logits_scaling: float = 0.0
    has_ffn_bias: bool = False  # GPT-2 c_fc/c_proj biases; optional at the boundary
    eos_token_id: int = -1
    # Granite-4 hybrid (purego/tensor/config.go:100-115): hybrid_layers[l] = True -> Mamba2 block
    mamba_num_heads: int = 0
    mamba_head_dim: int = 0
    mamba_state_size: int = 0
    mamba_n_groups: int = 0
    mamba_conv_kernel: int = 0
    hybrid_layers: tuple = ()

    def is_mamba(self, layer: int) -> bool:
        return bool(self.hybrid_layers) and bool(self.hybrid_layers[layer])

    def mamba_dims(self):
        """(E = expand * hidden, conv_dim, in_proj rows)."""
        E = self.mamba_num_heads * self.mamba_head_dim
        conv = E + 2 * self.mamba_n_groups * self.mamba_state_size
        return E, conv, E + conv + self.mamba_num_heads

    def kv_heads(self) -> int:
        if self.attention_type == ATTN_MQA:
            return 1
        if self.attention_type == ATTN_MHA:
            return self.num_heads
        return self.num_kv_heads

    def matmul_params(self) -> int:
        """Parameters of every matmul touched once per decode step."""
        H, F, L = self.hidden, self.ffn_dim, self.num_layers
        q = self.num_heads * self.head_dim
        kv = self.kv_heads() * self.head_dim
        attn = H * q + 2 * H * kv + q * H
        if self.use_moe:
            ffn = H * self.num_experts + self.num_experts * (2 * F * H + H * F)
        elif self.activation_type == ACT_SWIGLU:
            ffn = 3 * H * F
        else:
            ffn = 2 * H * F
        return L * (attn + ffn) + H * self.vocab_size

    def kv_bytes_per_token(self, bytes_per_el: int = 2) -> int:
        return 2 * self.num_layers * self.kv_heads() * self.head_dim * bytes_per_el

    def to_c(self, precision: int, num_kv_blocks: int = -1, max_num_seqs: int = 512,
             max_batched_tokens: int = 16384, device: int = 0, tp_size: int = 1,
             tp_rank: int = 0) -> CConfig:
        c = CConfig()
        c.abi_version = ABI_VERSION
        c.vocab_size = self.vocab_size
        c.hidden = self.hidden
        c.num_layers = self.num_layers
        c.num_heads = self.num_heads
        c.num_kv_heads = self.kv_heads()
        c.head_dim = self.head_dim
        c.ffn_dim = self.ffn_dim
        c.max_seq_len = self.max_seq_len
        c.attention_type = self.attention_type
        c.norm_type = self.norm_type
        c.position_type = self.position_type
        c.activation_type = self.activation_type
        c.block_style = self.block_style
        c.tied_embedding = int(self.tied_embedding)
        c.use_moe = int(self.use_moe)
        c.num_experts = self.num_experts
        c.num_experts_per_tok = self.num_experts_per_tok
        c.norm_eps = self.norm_eps
        c.embedding_multiplier = self.embedding_multiplier
        c.attention_multiplier = self.attention_multiplier
        c.residual_multiplier = self.residual_multiplier
        c.logits_scaling = self.logits_scaling
        c.rope_base = self.rope_base
        c.precision = precision
        c.num_kv_blocks = num_kv_blocks
        c.max_num_seqs = max_num_seqs
        c.max_batched_tokens = max_batched_tokens
        c.tp_size = tp_size
        c.tp_rank = tp_rank
        c.device = device
        c.mamba_num_heads = self.mamba_num_heads
        c.mamba_head_dim = self.mamba_head_dim
        c.mamba_state_size = self.mamba_state_size
        c.mamba_n_groups = self.mamba_n_groups
        c.mamba_conv_kernel = self.mamba_conv_kernel
        return c


def gpt2_small() -> ModelConfig:  # config.go:125-148
    return ModelConfig("gpt2", 50257, 768, 12, 12, 12, 64, 3072, 1024, ATTN_MHA, NORM_LAYER,
                       POS_LEARNED, ACT_GELU, BLOCK_SEQUENTIAL, True, 1e-5, has_ffn_bias=True,
                       eos_token_id=50256)


def llama32_1b() -> ModelConfig:  # HF meta-llama/Llama-3.2-1B; ref MaxSeqLen preset 4096 (config.go:222)
    return ModelConfig("llama-3.2-1b", 128256, 2048, 16, 32, 8, 64, 8192, 4096, ATTN_GQA, NORM_RMS,
                       POS_ROPE, ACT_SWIGLU, BLOCK_SEQUENTIAL, True, 1e-5, rope_base=500000.0,
                       eos_token_id=128001)


def llama32_3b() -> ModelConfig:  # HF meta-llama/Llama-3.2-3B
    return ModelConfig("llama-3.2-3b", 128256, 3072, 28, 24, 8, 128, 8192, 4096, ATTN_GQA, NORM_RMS,
                       POS_ROPE, ACT_SWIGLU, BLOCK_SEQUENTIAL, True, 1e-5, rope_base=500000.0,
                       eos_token_id=128001)


def granite_moe_1b() -> ModelConfig:  # config.go:330-376 + docs/ARCHITECTURE_GUIDE.md:1773-1776
    return ModelConfig("granite-3.0-1b-a400m", 49155, 1024, 24, 16, 8, 64, 512, 4096, ATTN_GQA,
                       NORM_RMS, POS_ROPE, ACT_SWIGLU, BLOCK_SEQUENTIAL, True, 1e-6,
                       rope_base=10000.0, use_moe=True, num_experts=32, num_experts_per_tok=8,
                       embedding_multiplier=12.0, attention_multiplier=0.015625,
                       residual_multiplier=0.22, logits_scaling=6.0, eos_token_id=0)


def falcon_7b() -> ModelConfig:  # config.go:151-177
    return ModelConfig("falcon-7b", 65024, 4544, 32, 71, 1, 64, 18176, 2048, ATTN_MQA, NORM_LAYER,
                       POS_ROPE, ACT_GELU, BLOCK_PARALLEL, False, 1e-5, rope_base=10000.0,
                       eos_token_id=11)


def tiny(base: ModelConfig, **kw) -> ModelConfig:
    """Scaled-down sibling of a preset for parity tests the CPU oracle finishes in seconds."""
    return replace(base, name=base.name + "-tiny", **kw)


def tiny_gpt2(**kw) -> ModelConfig:
    d = dict(vocab_size=1000, hidden=128, num_layers=2, num_heads=4, num_kv_heads=4, head_dim=32,
             ffn_dim=512, max_seq_len=640)
    d.update(kw)
    return tiny(gpt2_small(), **d)


def tiny_llama(**kw) -> ModelConfig:
    d = dict(vocab_size=1000, hidden=256, num_layers=2, num_heads=8, num_kv_heads=2, head_dim=32,
             ffn_dim=512, max_seq_len=1024)
    d.update(kw)
    return tiny(llama32_1b(), **d)


def tiny_granite(**kw) -> ModelConfig:
    d = dict(vocab_size=1000, hidden=128, num_layers=2, num_heads=4, num_kv_heads=2, head_dim=32,
             ffn_dim=64, max_seq_len=1024, num_experts=8, num_experts_per_tok=2)
    d.update(kw)
    return tiny(granite_moe_1b(), **d)


def tiny_hybrid(**kw) -> ModelConfig:
    """Granite-4-style hybrid: Mamba2 and GQA attention blocks (no positional encoding,
    config.go:241-300), shared SwiGLU MLP, muP multipliers."""
    d = dict(vocab_size=1000, hidden=128, num_layers=4, num_heads=4, num_kv_heads=2, head_dim=32,
             ffn_dim=256, max_seq_len=1024, position_type=POS_NONE, use_moe=False, num_experts=0,
             num_experts_per_tok=0, mamba_num_heads=8, mamba_head_dim=32, mamba_state_size=16,
             mamba_n_groups=2, mamba_conv_kernel=4, hybrid_layers=(True, False, True, True))
    d.update(kw)
    return replace(tiny(granite_moe_1b(), **d), name="granite-4-hybrid-tiny")


def tiny_falcon(**kw) -> ModelConfig:
    d = dict(vocab_size=1000, hidden=7 * 32, num_layers=2, num_heads=7, num_kv_heads=1, head_dim=32,
             ffn_dim=4 * 7 * 32, max_seq_len=1024)
    d.update(kw)
    return tiny(falcon_7b(), **d)


PRESETS = {
    "gpt2": gpt2_small,
    "llama-3.2-1b": llama32_1b,
    "llama-3.2-3b": llama32_3b,
    "granite-moe-1b": granite_moe_1b,
    "falcon-7b": falcon_7b,
}
