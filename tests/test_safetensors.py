"""nvllm_load_safetensors (the half that needs no Go): checkpoints written in
the HuggingFace layouts of the four families the reference maps (generic_loader.go:60-182) load
straight into the library and give logits BIT-IDENTICAL to the nvllm_set_weight path fed with
the reference's post-load layouts of the same weights (so every transpose / fused-QKV split /
gate-up concatenation of generic_loader.go:353-604,674-765 is reproduced by the device repack).
The CPU half checks the container parsing through nvllm_safetensors_find."""
import ctypes
import os
import struct

import numpy as np
import pytest

from helpers import C, W, TINY, ROOT
import nvllm_pkg
import st_writer

L = nvllm_pkg.sub("lib")

ARCH = {"gpt2": C.ARCH_GPT2, "falcon": C.ARCH_FALCON, "llama": C.ARCH_LLAMA, "granite": C.ARCH_GRANITE_MOE}


def find(model_dir, name):
    lib = L.load()
    dt, nd, nb, fb = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_uint64(), ctypes.c_uint64()
    shape = (ctypes.c_int64 * 8)()
    rc = lib.nvllm_safetensors_find(os.fsencode(model_dir), name.encode(), ctypes.byref(dt),
                                    ctypes.byref(nd), shape, ctypes.byref(nb), ctypes.byref(fb))
    if rc:
        return None
    return dt.value, list(shape[:nd.value]), nb.value, fb.value


def test_container_parsing_single_and_sharded(tmp_path):
    cfg = TINY["falcon"]()
    ws = list(W.random_weights(cfg, 77))
    for shards, dtype in ((1, "F32"), (3, "BF16"), (2, "F16")):
        d = str(tmp_path / f"falcon_{shards}_{dtype}")
        tensors = st_writer.export_checkpoint(d, cfg, ws, C.ARCH_FALCON, dtype=dtype, shards=shards)
        for name, a in tensors.items():
            got = find(d, name)
            assert got is not None, name
            dt, shape, nbytes, first = got
            assert dt == {"F32": C.DT_F32, "BF16": C.DT_BF16, "F16": C.DT_F16}[dtype]
            assert shape == list(a.shape)
            esz = 4 if dtype == "F32" else 2
            assert nbytes == a.size * esz
            want = st_writer._bytes(a.reshape(-1)[:8 // esz], dtype)
            assert struct.pack("<Q", first)[:len(want)] == want
        # generic_loader.go:621-629: a name is retried with the "transformer." prefix
        assert find(d, "word_embeddings.weight") is not None
        assert find(d, "ln_f.weight") is not None
        assert find(d, "no.such.tensor") is None
        assert b"not found" in L.load().nvllm_last_error(None)


def test_container_errors(tmp_path):
    d = tmp_path / "bad"
    d.mkdir()
    assert find(str(d), "x") is None  # no model.safetensors
    (d / "model.safetensors").write_bytes(struct.pack("<Q", 1 << 40) + b"{}")
    assert find(str(d), "x") is None
    assert b"header length" in L.load().nvllm_last_error(None)
    (d / "model.safetensors").write_bytes(struct.pack("<Q", 9) + b'{"a": [1}')
    assert find(str(d), "a") is None
    hdr = b'{"a":{"dtype":"F32","shape":[4],"data_offsets":[0,64]}}'
    (d / "model.safetensors").write_bytes(struct.pack("<Q", len(hdr)) + hdr + b"\0" * 16)
    assert find(str(d), "a") is None  # offsets outside the file
    assert b"outside" in L.load().nvllm_last_error(None)
    hdr = b'{"a":{"dtype":"I64","shape":[2],"data_offsets":[0,16]}}'
    (d / "model.safetensors").write_bytes(struct.pack("<Q", len(hdr)) + hdr + b"\0" * 16)
    assert find(str(d), "a")[0] == -1  # listed, dtype unsupported


def _logits(run, cfg, steps=3):
    prompt = [int(t) for t in W.random_prompt(cfg.vocab_size, 20, 99)]
    toks, outs = list(prompt), []
    for i in range(steps):
        out, lg = run.run_flat([toks], i == 0, [0], [[0]], [0.0], want_logits=True)
        outs.append(lg[0].copy())
        toks.append(int(out[0]))
    return np.stack(outs), toks


@pytest.mark.gpu
@pytest.mark.parametrize("family", ["gpt2", "llama", "granite", "falcon"])
@pytest.mark.parametrize("prec,dtype,shards", [(C.PREC_FP32, "F32", 1), (C.PREC_BF16, "BF16", 3),
                                               (C.PREC_BF16, "F16", 1)])
def test_loader_matches_set_weight_bit_for_bit(tmp_path, family, prec, dtype, shards):
    R = nvllm_pkg.sub("runner")
    cfg = TINY[family]()
    ws = list(W.random_weights(cfg, 4242, round_bf16=(dtype == "BF16")))
    if dtype == "F16":  # make the values exactly representable so both paths see the same bits
        ws = [(i, l, a.astype(np.float16).astype(np.float32)) for i, l, a in ws]
    d = str(tmp_path / "ckpt")
    st_writer.export_checkpoint(d, cfg, ws, ARCH[family], dtype=dtype, shards=shards,
                                **({"with_unloaded_mlp_bias": True} if family == "gpt2" else {}))
    # the set_weight side never hands GPT-2's MLP biases: the reference loader does not find them
    direct = [(i, l, a) for i, l, a in ws if i not in (C.W_FFN_B1, C.W_FFN_B2)]
    a = R.CudaModelRunner(cfg, direct, precision=prec, num_kv_blocks=4, sampling=None)
    la, ta = _logits(a, cfg)
    a.close()
    b = R.CudaModelRunner(cfg, None, precision=prec, num_kv_blocks=4, sampling=None, model_dir=d,
                          arch=ARCH[family])
    assert b.loaded_tensors > 0 and b.loaded_bytes > 0
    lb, tb = _logits(b, cfg)
    b.close()
    assert ta == tb
    assert np.array_equal(la, lb)


@pytest.mark.gpu
def test_loader_untied_without_lm_head_falls_back_to_embedding(tmp_path):
    """generic_loader.go:255-259: LMHead == nil -> Transpose(TokenEmbedding)."""
    from dataclasses import replace
    R = nvllm_pkg.sub("runner")
    cfg = replace(TINY["llama"](), tied_embedding=False)
    ws = [(i, l, a) for i, l, a in W.random_weights(cfg, 5) if i != C.W_LM_HEAD]
    emb = [a for i, l, a in ws if i == C.W_TOKEN_EMBEDDING][0]
    d = str(tmp_path / "ckpt")
    st_writer.export_checkpoint(d, replace(cfg, tied_embedding=True), ws, C.ARCH_LLAMA)
    a = R.CudaModelRunner(cfg, ws + [(C.W_LM_HEAD, -1, np.ascontiguousarray(emb.T))],
                          precision=C.PREC_FP32, num_kv_blocks=4, sampling=None)
    la, _ = _logits(a, cfg)
    a.close()
    b = R.CudaModelRunner(cfg, None, precision=C.PREC_FP32, num_kv_blocks=4, sampling=None,
                          model_dir=d, arch=C.ARCH_LLAMA)
    lb, _ = _logits(b, cfg)
    b.close()
    assert np.array_equal(la, lb)


@pytest.mark.gpu
def test_loader_reports_missing_tensor(tmp_path):
    R = nvllm_pkg.sub("runner")
    cfg = TINY["llama"]()
    ws = list(W.random_weights(cfg, 5))
    d = str(tmp_path / "ckpt")
    t = st_writer.hf_tensors(cfg, ws, C.ARCH_LLAMA)
    del t["model.layers.1.self_attn.v_proj.weight"]
    os.makedirs(d)
    st_writer.write_safetensors(os.path.join(d, "model.safetensors"), t)
    with pytest.raises(L.NvllmError, match="v_proj"):
        R.CudaModelRunner(cfg, None, precision=C.PREC_FP32, num_kv_blocks=4, sampling=None,
                          model_dir=d, arch=C.ARCH_LLAMA)
