"""ctypes wrapper of oracle/libnvllm_oracle.so — TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
See nvllm_oracle.c for what it restates (reference file:line per function) and
for the "parity unpinned" statement.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libnvllm_oracle.so")
_lib = None

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "nvllm_oracle.c")
    hdr = os.path.join(_HERE, "..", "include", "nvllm.h")
    if force or not os.path.exists(_SO) or any(
            os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_SO) for p in (src, hdr)):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libnvllm_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        L.oracle_create.restype = ctypes.c_void_p
        L.oracle_create.argtypes = [ctypes.c_void_p]
        L.oracle_destroy.argtypes = [ctypes.c_void_p]
        L.oracle_set_weight.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _f32p]
        L.oracle_set_layer_types.argtypes = [ctypes.c_void_p, _i32p, ctypes.c_int]
        L.oracle_mamba_layer.argtypes = [ctypes.c_void_p, ctypes.c_int, _f32p, _f32p, ctypes.c_int, _f32p]
        L.oracle_kv_create.restype = ctypes.c_void_p
        L.oracle_kv_create.argtypes = [ctypes.c_void_p]
        L.oracle_kv_destroy.argtypes = [ctypes.c_void_p]
        L.oracle_kv_reset.argtypes = [ctypes.c_void_p]
        L.oracle_kv_len.argtypes = [ctypes.c_void_p]
        L.oracle_kv_read.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        L.oracle_forward.argtypes = [ctypes.c_void_p, _i32p, ctypes.c_int, ctypes.c_void_p,
                                     ctypes.c_int, _f32p, ctypes.c_int, _f32p]
        L.oracle_argmax.argtypes = [_f32p, ctypes.c_int]
        L.oracle_sample.argtypes = [_f32p, ctypes.c_int, _i32p, ctypes.c_int, ctypes.c_float,
                                    ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_float]
        L.oracle_split_falcon_qkv.argtypes = [_f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              _f32p, _f32p]
        L.oracle_split_gpt2_qkv.argtypes = [_f32p, ctypes.c_int, _f32p, _f32p, _f32p]
        L.oracle_matmul.argtypes = [_f32p, _f32p, _f32p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        L.oracle_layer_norm.argtypes = [_f32p, _f32p, _f32p, ctypes.c_float, _f32p, ctypes.c_int,
                                        ctypes.c_int]
        L.oracle_gelu.argtypes = [_f32p, ctypes.c_size_t]
        L.oracle_silu.argtypes = [_f32p, ctypes.c_size_t]
        _lib = L
    return _lib


def _fp(a: np.ndarray):
    return a.ctypes.data_as(_f32p)


def _ip(a: np.ndarray):
    return a.ctypes.data_as(_i32p)


class OracleKV:
    """purego/tensor/kv_cache.go KVCache for one sequence."""

    def __init__(self, model: "OracleModel"):
        self._m = model
        self._p = lib().oracle_kv_create(model._p)

    def __len__(self):
        return lib().oracle_kv_len(self._p)

    def reset(self):
        lib().oracle_kv_reset(self._p)

    def read(self, layer: int, head: int, which: int) -> np.ndarray:
        out = np.empty((len(self), self._m.cfg.head_dim), dtype=np.float32)
        lib().oracle_kv_read(self._p, layer, head, which, _fp(out))
        return out

    def __del__(self):
        try:
            lib().oracle_kv_destroy(self._p)
        except Exception:
            pass


class OracleModel:
    """TransformerModel restated on the CPU (generic_model.go:4-19)."""

    def __init__(self, cfg, weights):
        """cfg: package ModelConfig; weights: iterable of (weight_id, layer, fp32 ndarray)."""
        self.cfg = cfg
        self._c = cfg.to_c(precision=0)
        self._p = lib().oracle_create(ctypes.byref(self._c))
        if not self._p:
            raise RuntimeError("oracle_create failed")
        self._keep = []
        if getattr(cfg, "hybrid_layers", ()):
            types = np.asarray([1 if t else 0 for t in cfg.hybrid_layers], dtype=np.int32)
            if lib().oracle_set_layer_types(self._p, _ip(types), len(types)) != 0:
                raise ValueError("hybrid_layers must name every layer")
        for wid, layer, arr in weights:
            a = np.ascontiguousarray(arr, dtype=np.float32)
            self._keep.append(a)
            if lib().oracle_set_weight(self._p, wid, layer, _fp(a)) != 0:
                raise ValueError(f"oracle_set_weight({wid},{layer}) failed")

    def mamba_layer(self, layer: int, x: np.ndarray, state: np.ndarray) -> np.ndarray:
        """Mamba2Layer.Forward of one layer: x [T, H] -> [T, H]; `state` [heads, hd, N] is updated."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty_like(x)
        lib().oracle_mamba_layer(self._p, layer, _fp(x), _fp(out), x.shape[0], _fp(state))
        return out

    def new_cache(self) -> OracleKV:
        return OracleKV(self)

    def forward(self, tokens, cache: OracleKV, pos_offset: int | None = None, all_rows=False,
                return_hidden=False):
        """ForwardWithCache + GetLogitsForLastToken. Returns logits [V] (or [T,V])."""
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        T = int(t.shape[0])
        if pos_offset is None:
            pos_offset = len(cache)
        V = self.cfg.vocab_size
        out = np.empty((T if all_rows else 1, V), dtype=np.float32)
        hid = np.empty((self.cfg.hidden,), dtype=np.float32)
        rc = lib().oracle_forward(self._p, _ip(t), T, cache._p, pos_offset, _fp(out),
                                  int(all_rows), _fp(hid))
        if rc != 0:
            raise RuntimeError("oracle_forward failed (bad token id or position >= max_seq_len)")
        res = out if all_rows else out[0]
        return (res, hid) if return_hidden else res

    def greedy(self, prompt, steps: int):
        """cmd/ask/main.go:287-360 generateResponse with argmax (:389-402).
        Returns (tokens[steps], logits[steps, V])."""
        cache = self.new_cache()
        toks, logs = [], []
        logits = self.forward(prompt, cache)
        for _ in range(steps):
            nt = argmax(logits)
            toks.append(nt)
            logs.append(logits.copy())
            if len(toks) == steps:
                break
            logits = self.forward([nt], cache)
        return np.array(toks, dtype=np.int32), np.stack(logs)

    def greedy_batch(self, prompts, steps: int, threads: int | None = None):
        """greedy() for independent sequences, one host thread each (ctypes releases the GIL;
        oracle_forward only reads the model).  The first sequence runs alone: a tied LM head is
        materialised lazily by the first forward call (generic_loader.go:255-259 transposes the
        embedding once), which must not race.  Returns ([tokens], [logits])."""
        from concurrent.futures import ThreadPoolExecutor
        if not len(prompts):
            return [], []
        res = [self.greedy(prompts[0], steps)]
        if len(prompts) > 1:
            nt = max(1, min(len(prompts) - 1, threads or os.cpu_count() or 1))
            with ThreadPoolExecutor(max_workers=nt) as ex:
                res += list(ex.map(lambda p: self.greedy(p, steps), prompts[1:]))
        return [r[0] for r in res], [r[1] for r in res]

    def __del__(self):
        try:
            lib().oracle_destroy(self._p)
        except Exception:
            pass


def argmax(logits: np.ndarray) -> int:
    a = np.ascontiguousarray(logits, dtype=np.float32)
    return int(lib().oracle_argmax(_fp(a), a.shape[0]))


def sample(logits, history, temperature=1.0, top_p=1.0, top_k=0, rep_penalty=1.0, u=0.5) -> int:
    a = np.ascontiguousarray(logits, dtype=np.float32)
    h = np.ascontiguousarray(history if history is not None else [], dtype=np.int32)
    return int(lib().oracle_sample(_fp(a), a.shape[0], _ip(h), h.shape[0], temperature, top_p,
                                   top_k, rep_penalty, u))


def matmul(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    m, k = a.shape
    n = b.shape[1]
    c = np.empty((m, n), dtype=np.float32)
    lib().oracle_matmul(_fp(a), _fp(b), _fp(c), m, k, n)
    return c


def layer_norm(x, w, b, eps):
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    y = np.empty_like(x)
    bp = _fp(np.ascontiguousarray(b, dtype=np.float32)) if b is not None else None
    lib().oracle_layer_norm(_fp(x), _fp(w), bp, eps, _fp(y), x.shape[0], x.shape[1])
    return y


def gelu(x):
    y = np.array(x, dtype=np.float32, copy=True)
    lib().oracle_gelu(_fp(y), y.size)
    return y


def silu(x):
    y = np.array(x, dtype=np.float32, copy=True)
    lib().oracle_silu(_fp(y), y.size)
    return y


def split_falcon_qkv(qkv, hidden, nh, hd):
    qkv = np.ascontiguousarray(qkv, dtype=np.float32)
    Q = np.empty((hidden, nh * hd), dtype=np.float32)
    KV = np.empty((hidden, 2 * hd), dtype=np.float32)
    lib().oracle_split_falcon_qkv(_fp(qkv), hidden, nh, hd, _fp(Q), _fp(KV))
    return Q, KV


def split_gpt2_qkv(qkv, hidden):
    qkv = np.ascontiguousarray(qkv, dtype=np.float32)
    Q = np.empty((hidden, hidden), dtype=np.float32)
    K = np.empty_like(Q)
    V = np.empty_like(Q)
    lib().oracle_split_gpt2_qkv(_fp(qkv), hidden, _fp(Q), _fp(K), _fp(V))
    return Q, K, V
