"""CudaModelRunner — the ``ModelRunner`` of the reference, backed by libnvllm_b200.so.

Mirrors, method for method, what the Go engine sees:

  ModelRunner interface        /root/reference nanovllm/model_runner.go:9-16
      Run(seqs, isPrefill) ([]int, error)        -> run(seqs, is_prefill)
      Close() error                              -> close()
  TensorModelRunner extras     nanovllm/tensor_model_runner.go:35-52,100-111
      SetSamplingParams / ...WithRepetition      -> set_sampling_params(...)
      ClearCache / ClearAllCaches                -> clear_cache / clear_all_caches
  GenericModelRunner extras    purego/generic_runner.go:184-196
      GetVocabSize / GetModelConfig              -> get_vocab_size / get_model_config

The Go shim (go/cuda_model_runner.go) does the same flattening over cgo; this file
is that shim in Python because the image has no Go toolchain.  Differences from
TensorModelRunner that are deliberate and visible at the boundary:
  * sequences are batched into one ragged device step instead of a per-sequence loop
    (tensor_model_runner.go:58);
  * KV state lives in the library's paged pool and is addressed only through
    ``seq.block_table`` / ``seq.num_cached_tokens`` (which every reference runner
    ignores), so ClearCache has nothing to free;
  * errors come back as NvllmError (the Go shim wraps them with fmt.Errorf).
"""
from __future__ import annotations

import ctypes
import os
from typing import Iterable, List, Optional, Sequence as Seq, Tuple

import numpy as np

from . import config as C
from . import lib as L


# NewTensorModelRunner -> tensor.DefaultSamplingParams() (tensor_model_runner.go:27-31,
# sampling.go:17-24); the Go shim applies the same values in NewCudaModelRunner
DEFAULT_SAMPLING = dict(temperature=1.0, top_p=1.0, top_k=0, repetition_penalty=1.2)


class CudaModelRunner:
    def __init__(self, cfg: C.ModelConfig, weights: Iterable[Tuple[int, int, np.ndarray]],
                 precision: int = C.PREC_BF16, num_kv_blocks: int = -1, device: int = 0,
                 max_num_seqs: int = 512, max_batched_tokens: int = 16384,
                 weight_dtype: Optional[int] = None, tp_size: int = 1, tp_rank: int = 0,
                 tp_id: Optional[bytes] = None, sampling: Optional[dict] = DEFAULT_SAMPLING,
                 model_dir: Optional[str] = None, arch: Optional[int] = None):
        """Replaces NewTensorModelRunner(modelDir) (tensor_model_runner.go:21-32): instead of a
        directory, takes the tensors in the layouts LoadModel leaves them in
        (generic_loader.go:184-265), as (weight_id, layer, ndarray) triples.

        weight_dtype: dtype the host buffers are handed over in (DT_F32 default; DT_BF16 sends
        bf16 bit patterns, halving the upload).

        tp_size > 1: this runner is rank tp_rank of a tensor-parallel group (one runner per GPU,
        every rank handed the same FULL tensors and the same Run calls); tp_id is the 128-byte
        group id from tp_unique_id(), created on one rank and distributed by the host code.

        sampling: the runner-wide sampling parameters Run() draws with; the default is the
        reference runner's (temperature 1.0, repetition penalty 1.2: every row is drawn by the
        device-side SampleWithHistory).  None = greedy first-max rows, no filters (what the
        parity tests and the benchmark use: the reference's draw uses an unseeded global RNG).

        model_dir + arch (config.ARCH_*): what NewTensorModelRunner really takes — the library
        reads model.safetensors / the sharded index itself (nvllm_load_safetensors: mmap, file
        dtype kept, the loader's transposes and QKV splits done on the device); `weights` is
        then ignored and may be None."""
        self._lib = L.load()
        self.cfg = cfg
        self.precision = precision
        self._h = ctypes.c_void_p()
        cc = cfg.to_c(precision, num_kv_blocks=num_kv_blocks, max_num_seqs=max_num_seqs,
                      max_batched_tokens=max_batched_tokens, device=device, tp_size=tp_size,
                      tp_rank=tp_rank)
        idbuf = None
        if tp_size > 1:
            if tp_id is None or len(tp_id) != 128:
                raise ValueError("tp_size > 1 needs the 128-byte tp_id of the group")
            idbuf = (ctypes.c_uint8 * 128).from_buffer_copy(tp_id)
        L.check(self._lib.nvllm_create(ctypes.byref(cc), idbuf, ctypes.byref(self._h)))
        self.loaded_bytes, self.loaded_tensors = 0, 0
        try:
            if getattr(cfg, "hybrid_layers", ()):
                # ModelConfig.HybridLayers (config.go:113): Mamba2 blocks of the Granite-4 hybrid
                types = (ctypes.c_int32 * cfg.num_layers)(*[1 if t else 0 for t in cfg.hybrid_layers])
                L.check(self._lib.nvllm_set_layer_types(self._h, types, cfg.num_layers), self._h)
            if model_dir is not None:
                nb, nt = ctypes.c_uint64(), ctypes.c_int32()
                L.check(self._lib.nvllm_load_safetensors(self._h, os.fsencode(model_dir), int(arch),
                                                         ctypes.byref(nb), ctypes.byref(nt)), self._h)
                self.loaded_bytes, self.loaded_tensors = nb.value, nt.value
            else:
                for wid, layer, arr in weights:
                    self.set_weight(wid, layer, arr, weight_dtype)
            L.check(self._lib.nvllm_finalize(self._h), self._h)
        except Exception:
            self.close()
            raise
        # temperature <= 0 => greedy first-max (cmd/ask/main.go:389-402); deviation from the
        # reference runner, where 0 means "no temperature scaling" and the row is still drawn
        self.temperature = 0.0
        if sampling is not None:
            try:
                self.set_sampling_params(**sampling)
            except Exception:
                self.close()
                raise
        self.last_stats = L.RunStats()
        self.last_logits: Optional[np.ndarray] = None

    # ---- weights -----------------------------------------------------------
    def set_weight(self, wid: int, layer: int, arr: np.ndarray, weight_dtype: Optional[int] = None):
        if weight_dtype == C.DT_BF16:
            from .weights import to_bf16_bits
            a = to_bf16_bits(arr)
            dt = C.DT_BF16
        elif arr.dtype == np.uint16:
            a, dt = np.ascontiguousarray(arr), C.DT_BF16
        elif arr.dtype == np.float16:
            a, dt = np.ascontiguousarray(arr), C.DT_F16
        else:
            a, dt = np.ascontiguousarray(arr, dtype=np.float32), C.DT_F32
        L.check(self._lib.nvllm_set_weight(self._h, wid, layer, dt, a.ctypes.data_as(ctypes.c_void_p),
                                           a.size), self._h)

    # ---- ModelRunner -------------------------------------------------------
    def run(self, seqs: List, is_prefill: bool) -> List[int]:
        """ModelRunner.Run: next token for every scheduled sequence, same order."""
        if not seqs:
            return []
        toks, _ = self.run_flat(
            [s.token_ids for s in seqs], is_prefill,
            num_cached=[s.num_cached_tokens for s in seqs],
            block_tables=[s.block_table for s in seqs],
            temperature=[self.temperature] * len(seqs))
        return [int(t) for t in toks]

    def run_flat(self, token_lists: Seq[Seq[int]], is_prefill: bool,
                 num_cached: Optional[Seq[int]] = None,
                 block_tables: Seq[Seq[int]] = (), temperature: Optional[Seq[float]] = None,
                 uniforms: Optional[Seq[float]] = None, want_logits: bool = False):
        """The flattened call of include/nvllm.h::nvllm_run. Returns (tokens[n], logits[n,V]|None)."""
        n = len(token_lists)
        cu = np.zeros(n + 1, dtype=np.int32)
        np.cumsum([len(t) for t in token_lists], out=cu[1:])
        flat = np.concatenate([np.asarray(t, dtype=np.int32) for t in token_lists]) if n else \
            np.zeros(0, np.int32)
        cub = np.zeros(n + 1, dtype=np.int32)
        np.cumsum([len(b) for b in block_tables], out=cub[1:])
        bt = np.concatenate([np.asarray(b, dtype=np.int32) for b in block_tables]) \
            if n and cub[-1] else np.zeros(0, np.int32)
        nc = np.asarray(num_cached, dtype=np.int32) if num_cached is not None else None
        temp = np.asarray(temperature, dtype=np.float32) if temperature is not None else None
        uni = np.asarray(uniforms, dtype=np.float32) if uniforms is not None else None
        return self.run_arrays(n, is_prefill, flat, cu, nc, bt, cub, temp, uni, want_logits)

    def run_arrays(self, n, is_prefill, flat, cu, nc, bt, cub, temp, uni, want_logits=False):
        out = np.empty(n, dtype=np.int32)
        logits = np.empty((n, self.cfg.vocab_size), dtype=np.float32) if want_logits else None
        rc = self._lib.nvllm_run(self._h, n, int(is_prefill), L._ip(flat), L._ip(cu), L._ip(nc),
                                 L._ip(bt), L._ip(cub), L._fp(temp), L._fp(uni), L._ip(out),
                                 L._fp(logits), ctypes.byref(self.last_stats))
        L.check(rc, self._h)
        self.last_logits = logits
        return out, logits

    def run_resident(self) -> None:
        """One more decode step for the previous decode batch, fully device-resident."""
        L.check(self._lib.nvllm_run_resident(self._h, ctypes.byref(self.last_stats)), self._h)

    def rerun(self) -> None:
        """Re-execute the last prefill step from its device-resident inputs (no copies)."""
        L.check(self._lib.nvllm_rerun(self._h, ctypes.byref(self.last_stats)), self._h)

    def trace_begin(self, capacity: int = 1 << 16) -> None:
        L.check(self._lib.nvllm_trace_begin(self._h, capacity), self._h)

    def trace_end(self, max_records: int = 1 << 16) -> np.ndarray:
        """[n, 8] uint64: seq, kernel id, cta, sm, t_entry, t_after_wait, t_exit, extra (ns)."""
        out = np.zeros((max_records, 8), dtype=np.uint64)
        n = ctypes.c_int()
        L.check(self._lib.nvllm_trace_end(self._h, out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                          max_records, ctypes.byref(n)), self._h)
        return out[:n.value]

    def timer_start(self) -> None:
        L.check(self._lib.nvllm_timer_record(self._h, 0), self._h)

    def timer_stop(self) -> float:
        """Records the stop event, waits for the stream, returns device milliseconds."""
        L.check(self._lib.nvllm_timer_record(self._h, 1), self._h)
        ms = ctypes.c_float()
        L.check(self._lib.nvllm_timer_elapsed(self._h, ctypes.byref(ms)), self._h)
        return ms.value

    def set_profile(self, mode) -> None:
        """0 / False off, 1 / True CUDA events around every launch, 2 device timeline."""
        L.check(self._lib.nvllm_set_profile(self._h, int(mode)), self._h)

    def get_profile(self) -> dict:
        """{class: {ms, bytes, flops, launches}} accumulated since set_profile(True)."""
        arr = (L.ProfileEntry * len(L.KERNEL_CLASSES))()
        L.check(self._lib.nvllm_get_profile(self._h, arr), self._h)
        return {name: {"ms": e.ms, "bytes": e.bytes, "flops": e.flops, "launches": e.launches,
                       "impl_bytes": e.impl_bytes}
                for name, e in zip(L.KERNEL_CLASSES, arr) if e.launches}

    def sync(self, n: int = 0):
        out = np.empty(max(n, 1), dtype=np.int32)
        ms = ctypes.c_float()
        L.check(self._lib.nvllm_sync(self._h, L._ip(out), n, ctypes.byref(ms)), self._h)
        return out[:n], ms.value

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.nvllm_destroy(self._h)
            self._h = ctypes.c_void_p()

    # ---- TensorModelRunner / GenericModelRunner extras ---------------------
    def set_sampling_params(self, temperature: float, top_p: float = 1.0, top_k: int = 0,
                            repetition_penalty: float = 1.0):
        self.temperature = float(temperature)
        L.check(self._lib.nvllm_set_sampling(self._h, top_p, top_k, repetition_penalty), self._h)

    def set_seed(self, seed: int):
        L.check(self._lib.nvllm_set_seed(self._h, seed), self._h)

    def set_graphs(self, enabled: bool):
        L.check(self._lib.nvllm_set_graphs(self._h, int(enabled)), self._h)

    def set_speculation(self, enabled: bool):
        L.check(self._lib.nvllm_set_speculation(self._h, int(enabled)), self._h)

    def clear_cache(self, seq_id: int) -> None:  # state is addressed by block tables only
        return None

    def clear_all_caches(self) -> None:
        return None

    def get_vocab_size(self) -> int:
        return self.cfg.vocab_size

    def get_model_config(self) -> C.ModelConfig:
        return self.cfg

    def memory_info(self) -> dict:
        w, k, s = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        L.check(self._lib.nvllm_memory_info(self._h, ctypes.byref(w), ctypes.byref(k),
                                            ctypes.byref(s)), self._h)
        return {"weights": w.value, "kv": k.value, "workspace": s.value}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
