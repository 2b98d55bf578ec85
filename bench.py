#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark (driver contract: see DESIGN.md "Measurement").

Workload at N=1 (BASELINE.json configs[1]): Llama-3.2-1B (GQA, bf16 weights, fp32 accumulate),
random-init weights, 64 sequences continuously batched: one 64 x 512-token prefill step, then
256 greedy decode steps.  One bench "step" = that whole job (49152 tokens).  N>1: one replica
per GPU over disjoint request batches (weak scaling, no data-path collective).

  value  whole-job tokens/s with every input already resident in HBM (nvllm_rerun +
         nvllm_run_resident: no host<->device copies, no host sync inside the timed region)
  e2e    the same job through the C-ABI call nvllm_run with HOST buffers every step
         (token ids / block tables in, sampled tokens out; copies inside the timed region)
  roofline / kernels  per-kernel-class device time from a separate profiled pass of the
         same job (events around every launch), algorithmic bytes / flops per DESIGN.md
  cpu_baseline  the CPU oracle (restated reference, 1 thread) on a bounded sample

`--impl reference` times the reference's CPU algorithm (oracle port; the Go reference cannot be
built in this image) on the same config, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import nvllm_pkg  # noqa: E402

C = nvllm_pkg.sub("config")
W = nvllm_pkg.sub("weights")
from harness import go_engine as E  # noqa: E402  (restated Go scheduler / block manager: test + bench harness)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.device), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "power_w_max": float(max(power)) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# BASELINE.json configs[0..4] (SURVEY.md §8d "Prompts")
CONFIGS = {
    1: dict(model="gpt2", batch=1, prompt_len=32, decode_steps=64, shared_prefix=0, tp=False,
            label="configs[0]: 1 sequence, 32-token prompt + 64 greedy decode steps"),
    2: dict(model="llama-3.2-1b", batch=64, prompt_len=512, decode_steps=256, shared_prefix=0, tp=False,
            label="configs[1]: batch 64 continuous batching, 512-token prefill + 256 greedy decode steps"),
    3: dict(model="granite-moe-1b", batch=128, prompt_len=128, decode_steps=256, shared_prefix=0,
            tp=False, label="configs[2]: batch 128, 128-token prefill + 256 decode steps (context 128->384), "
                            "top-8 of 32 experts"),
    4: dict(model="falcon-7b", batch=1, prompt_len=2048, decode_steps=32, shared_prefix=0, tp=True,
            label="configs[3]: tensor parallel over all ranks, 1 sequence, 2048-token prefill + 32 greedy "
                  "decode steps (RoPE table extended to 2112 positions for the decode tail)"),
    5: dict(model="llama-3.2-3b", batch=16, prompt_len=4096, decode_steps=0, shared_prefix=3840,
            tp=False, label="configs[4]: per replica 16 prompts x 4096 tokens sharing a 3840-token "
                            "prefix (15 KV blocks): request 0 cold, 15 prefix-cache hits "
                            "(NumCachedTokens=3840); prefill only (positions stop at MaxSeqLen 4096)"),
}


class Job:
    """One bench "step" on one replica (or one tensor-parallel group): the prefill step(s) the Go
    scheduler would issue for the request set, then D decode steps.  Block tables and
    NumCachedTokens come from the restated Go block manager (harness/go_engine.py)."""

    def __init__(self, cfg, batch, prompt_len, decode_steps, seed0, shared_prefix=0):
        self.cfg, self.n, self.P, self.D = cfg, batch, prompt_len, decode_steps
        self.max_len = prompt_len + decode_steps + (1 if decode_steps else 0)
        V = cfg.vocab_size
        if shared_prefix:
            common = W.random_prompt(V, shared_prefix, seed0 + 99991)
            self.prompts = np.stack([np.concatenate(
                [common, W.random_prompt(V, prompt_len - shared_prefix, seed0 + i)]) for i in range(batch)])
        else:
            self.prompts = np.stack([W.random_prompt(V, prompt_len, seed0 + i) for i in range(batch)])
        nb = (self.max_len + 255) // 256
        self.nb = nb
        bm = E.BlockManager(batch * nb, 256)
        seqs = []
        for i in range(batch):
            s = E.Sequence([int(t) for t in self.prompts[i]])
            bm.allocate(s)
            seqs.append(s)
        # decode tail: take the blocks the sequence will grow into up front (the ABI allows a
        # block table longer than needed), from the same free list the Go engine would use
        for s in seqs:
            while len(s.block_table) < nb:
                bid = bm.free_block_ids[0]
                bm._allocate_block(bid)
                s.block_table.append(bid)
        self.tables = [s.block_table for s in seqs]
        self.num_cached = np.asarray([s.num_cached_tokens for s in seqs], dtype=np.int32)
        self.num_blocks = batch * nb
        self.bt = np.asarray(self.tables, dtype=np.int32).reshape(-1)
        self.cub = np.arange(0, (batch + 1) * nb, nb, dtype=np.int32)
        self.temp = np.zeros(batch, dtype=np.float32)
        # one prefill step for the whole request set, as scheduler.go:45-71 issues it: cache hits
        # ride in the SAME step as the request whose blocks they hit (both are allocated before
        # Run); the library appends every new K/V of a layer before any attention read of it
        self.prefill_steps = [list(range(batch))]
        self.tokens_per_step = batch * (prompt_len + decode_steps)
        self.computed_tokens_per_step = int(batch * (prompt_len + decode_steps) - self.num_cached.sum())

    def _prefill_args(self, idx):
        flat = self.prompts[idx].reshape(-1).astype(np.int32)
        cu = np.arange(0, (len(idx) + 1) * self.P, self.P, dtype=np.int32)
        bt = np.asarray([self.tables[i] for i in idx], dtype=np.int32).reshape(-1)
        cub = np.arange(0, (len(idx) + 1) * self.nb, self.nb, dtype=np.int32)
        return len(idx), True, flat, cu, self.num_cached[idx].copy(), bt, cub, self.temp[:len(idx)], None

    @property
    def simple(self):
        return len(self.prefill_steps) == 1

    def stage(self, run, step=0):
        """Upload one prefill step's inputs (outside any timed region; runs the step once)."""
        run.run_arrays(*self._prefill_args(self.prefill_steps[step]))

    # ---- device-resident pass: nothing crosses PCIe inside a timed segment
    def run_resident(self, run):
        """Simple jobs (one prefill step, staged beforehand): rerun + D resident decode steps."""
        launches = 0
        run.rerun()
        launches += run.last_stats.kernel_launches
        for _ in range(self.D):
            run.run_resident()
            launches += run.last_stats.kernel_launches
        return launches

    def run_resident_timed(self, run):
        """Jobs with several prefill steps: stage each (untimed), time its device-resident
        re-execution; returns (device ms summed over the segments, launches)."""
        ms, launches = 0.0, 0
        for k in range(len(self.prefill_steps)):
            self.stage(run, k)
            run.sync()
            run.timer_start()
            run.rerun()
            if k == len(self.prefill_steps) - 1:
                for _ in range(self.D):
                    run.run_resident()
            ms += run.timer_stop()
            launches += run.last_stats.kernel_launches
        return ms, launches

    # ---- end-to-end pass through nvllm_run with host buffers every step
    def run_e2e(self, run):
        n, P = self.n, self.P
        h2d = d2h = 0
        first = np.zeros(n, dtype=np.int32)
        for idx in self.prefill_steps:
            out, _ = run.run_arrays(*self._prefill_args(idx))
            h2d += run.last_stats.h2d_bytes
            d2h += run.last_stats.d2h_bytes
            first[idx] = out
        if not self.D:
            return h2d, d2h, first
        tok = np.zeros((n, self.max_len), dtype=np.int32)
        tok[:, :P] = self.prompts
        tok[:, P] = first
        lens = np.full(n, P + 1, dtype=np.int64)
        cols = np.arange(self.max_len)[None, :]
        cu = np.zeros(n + 1, dtype=np.int32)
        for _ in range(self.D):
            np.cumsum(lens, out=cu[1:])
            # every sequence of the job has the same length: the ragged flatten is a plain copy
            flat = np.ascontiguousarray(tok[:, :int(lens[0])]).reshape(-1)
            out, _ = run.run_arrays(n, False, flat, cu, None, self.bt, self.cub, self.temp, None)
            h2d += run.last_stats.h2d_bytes
            d2h += run.last_stats.d2h_bytes
            tok[np.arange(n), lens] = out
            lens += 1
        return h2d, d2h, tok


def device_random_weights(cfg, device, seed=1234):
    """Random-init weights of the same distributions as W.random_weights, drawn on the GPU and
    handed over as bf16 bit patterns (uint16): a 7 B model takes seconds instead of minutes of
    host RNG.  Every rank of a tensor-parallel group draws the SAME full tensors (same Philox
    seed per tensor).  Not bit-equal to the host generator: used where no CPU oracle is run on
    the same weights (the full-depth Falcon-7B timing)."""
    import torch
    g = torch.Generator(device=f"cuda:{device}")
    for wid, layer, shape, std, mean in W.weight_specs(cfg):
        g.manual_seed((seed * 1000003 + wid * 7919 + (layer + 1) * 104729) & 0x7FFFFFFF)
        t = torch.empty(shape, dtype=torch.float32, device=f"cuda:{device}")
        t.normal_(mean, std, generator=g)
        yield wid, layer, t.to(torch.bfloat16).view(torch.int16).cpu().numpy().view(np.uint16)
        del t


def make_runner(cfg, job, device, keep_host=None, tp_size=1, tp_rank=0, tp_id=None, fast_weights=False):
    R = nvllm_pkg.sub("runner")

    def gen():
        if fast_weights:
            yield from device_random_weights(cfg, device)
            return
        for wid, layer, a in W.random_weights(cfg, 1234, round_bf16=True):
            if keep_host is not None:
                keep_host.append((wid, layer, a))
            yield wid, layer, a

    max_T = max(len(g) for g in job.prefill_steps) * job.P
    return R.CudaModelRunner(cfg, gen(), precision=C.PREC_BF16, num_kv_blocks=job.num_blocks,
                             device=device, max_num_seqs=job.n, max_batched_tokens=max_T,
                             weight_dtype=None if fast_weights else C.DT_BF16, tp_size=tp_size,
                             tp_rank=tp_rank, tp_id=tp_id,
                             sampling=None)  # greedy rows: the job is BASELINE's greedy decode


def cpu_oracle_sample(cfg, weights, budget_s, prompt_tokens=None):
    """Time the restated reference (1 thread): prefill `p` tokens + decode `d` tokens of one
    sequence, sized to ~budget_s after a 2-token calibration.  The last-token logits of every
    step and the greedy tokens are kept so that the GPU can be checked on the same prompt."""
    from oracle import oracle as O
    m = O.OracleModel(cfg, weights)
    prompt = W.random_prompt(cfg.vocab_size, 64, 4321)
    t0 = time.perf_counter()
    m.forward(prompt[:2], m.new_cache())
    t_tok = (time.perf_counter() - t0) / 2
    total = int(max(4, min(48, budget_s / max(t_tok, 1e-3))))
    # a FIXED 12-token prompt whenever the budget allows (the parity check below then runs on the
    # same prompt on every box); slow models fall back to 2/3 of what fits
    p = prompt_tokens if prompt_tokens else (12 if total >= 18 else max(2, (2 * total) // 3))
    d = max(1, total - p)
    cache = m.new_cache()
    t0 = time.perf_counter()
    logits = m.forward(prompt[:p], cache)
    t_prefill = time.perf_counter() - t0
    step_logits, step_tokens = [], []
    t0 = time.perf_counter()
    for _ in range(d):
        nt = O.argmax(logits)
        step_logits.append(np.array(logits, dtype=np.float32, copy=True))
        step_tokens.append(int(nt))
        logits = m.forward([nt], cache)
    t_decode = time.perf_counter() - t0
    return {"tokens": p + d, "seconds": t_prefill + t_decode, "prefill_tokens": p,
            "decode_tokens": d, "prefill_tok_s": p / t_prefill, "decode_tok_s": d / t_decode,
            "prompt": [int(t) for t in prompt[:p]], "logits": step_logits,
            "greedy": step_tokens}, m


def gpu_parity(run, ref):
    """The same prompt through nvllm_run (bf16 weights, the library's production kernels at the
    full model size), last-token logits of every step against the oracle's.  Teacher-forced with
    the ORACLE's greedy tokens, so a near-tie flip (random-init logits are nearly flat) does not
    make the later steps incomparable; every flip is reported with the oracle's own margin
    between the two tokens.  Gate (BASELINE.json north_star, bf16 mode): max|dlogit| /
    max|logit| <= 2e-2 per step and identical greedy tokens."""
    toks = list(ref["prompt"])
    steps = len(ref["greedy"])
    errs, flips, margins = [], [], []
    for i in range(steps):
        out, logits = run.run_flat([toks], i == 0, [0], [[0]], [0.0], want_logits=True)
        r = ref["logits"][i]
        errs.append(float(np.abs(logits[0] - r).max() / np.abs(r).max()))
        if int(out[0]) != ref["greedy"][i]:
            flips.append(i)
            margins.append(float((r[ref["greedy"][i]] - r[int(out[0])]) / np.abs(r).max()))
        toks.append(ref["greedy"][i])
    worst = max(errs)
    return {"max_rel_err": worst, "tolerance": 2e-2, "tokens_equal": not flips, "steps": steps,
            "prompt_tokens": len(ref["prompt"]), "within_tolerance": bool(worst <= 2e-2),
            "flipped_steps": flips, "flip_margins_rel": margins, "teacher_forced": True,
            "per_step_rel_err": [round(e, 5) for e in errs],
            "against": "oracle/nvllm_oracle.c on the same bf16-rounded weights and prompt "
                       "(full-size model)"}


def tp_leg(dist, rank, world, device, steps, warmup, barrier):
    """BASELINE.json configs[3] on ALL ranks of this launch as ONE tensor-parallel group (tp =
    world; tp = 1 is the anchor of the strong-scaling curve): Falcon-7B bf16, full depth, one
    2048-token prompt + 32 greedy decode steps, random-init weights drawn on the device.  Then
    the parity check the single-GPU test-suite cannot run at tp > 1: a 2-layer full-width
    Falcon-7B (host-generated, bf16-rounded weights) against the CPU oracle on rank 0."""
    from dataclasses import replace
    D = nvllm_pkg.sub("dist")
    spec = CONFIGS[4]
    cfg = C.PRESETS[spec["model"]]()
    need = spec["prompt_len"] + spec["decode_steps"] + 1
    if cfg.max_seq_len < need:
        cfg = replace(cfg, max_seq_len=need + 31)
    tp = world
    tp_id = D.share_tp_id(dist, rank, nvllm_pkg.sub("lib").tp_unique_id) if tp > 1 else None
    job = Job(cfg, spec["batch"], spec["prompt_len"], spec["decode_steps"], seed0=4321)
    t0 = time.perf_counter()
    run = make_runner(cfg, job, device, None, tp_size=tp, tp_rank=rank if tp > 1 else 0, tp_id=tp_id,
                      fast_weights=True)
    load_s = time.perf_counter() - t0
    run.set_speculation(False)
    job.stage(run)
    for _ in range(warmup):
        job.run_resident(run)
    run.sync()
    barrier()
    run.timer_start()
    for _ in range(steps):
        job.run_resident(run)
    ms = D.reduce_max(dist, run.timer_stop(), f"cuda:{device}")
    barrier()
    # phase times (un-instrumented) and the communication share (events around every launch:
    # the only instrument that sees the NCCL calls; serialised, so shares not absolutes)
    job.stage(run)
    run.sync()
    barrier()
    run.timer_start()
    run.rerun()
    t_prefill = D.reduce_max(dist, run.timer_stop(), f"cuda:{device}")
    run.timer_start()
    for _ in range(job.D):
        run.run_resident()
    t_decode = D.reduce_max(dist, run.timer_stop(), f"cuda:{device}")
    job.stage(run)
    run.set_profile(1)
    run.rerun()
    pp = run.get_profile()
    run.set_profile(1)
    for _ in range(job.D):
        run.run_resident()
    pd = run.get_profile()
    run.set_profile(0)
    run.close()
    out = {"config": f"{cfg.name} bf16, {spec['label']}", "tp": tp, "scaling": "strong",
           "value": job.tokens_per_step * steps / (ms / 1e3), "unit": "tokens/s",
           "ms_per_step": ms / steps, "steps": steps,
           "prefill_tok_s": job.P / (t_prefill / 1e3), "decode_tok_s": job.D / (t_decode / 1e3),
           "prefill_ms": t_prefill, "decode_ms": t_decode, "weights_load_s": load_s,
           "prefill_gemm_TFLOPs_per_rank": (pp["gemm_tc5"]["flops"] / (pp["gemm_tc5"]["ms"] * 1e9))
           if "gemm_tc5" in pp else None}
    if tp > 1 and "comm" in pp:
        tot = sum(e["ms"] for e in pp.values())
        c = pp["comm"]
        algbw = c["bytes"] / (c["ms"] * 1e6)
        out.update({"comm_ms": c["ms"] * t_prefill / tot, "comm_share": c["ms"] / tot,
                    "comm_calls": c["launches"], "wire_MB_per_call": c["bytes"] / c["launches"] / 1e6,
                    "busbw_GBs": algbw * 2 * (tp - 1) / tp,
                    "decode_comm_share": pd["comm"]["ms"] / sum(e["ms"] for e in pd.values())
                    if "comm" in pd else None,
                    "comm_note": "shares from the per-launch event pass (serialised); comm_ms = "
                                 "share x the un-instrumented prefill time"})
    # ---- parity at tp = world: 2 layers at full width against the oracle
    cfg2 = replace(cfg, num_layers=2)
    ws = list(W.random_weights(cfg2, 1234, round_bf16=True))
    R = nvllm_pkg.sub("runner")
    tp_id2 = D.share_tp_id(dist, rank, nvllm_pkg.sub("lib").tp_unique_id) if tp > 1 else None
    r2 = R.CudaModelRunner(cfg2, ws, precision=C.PREC_BF16, num_kv_blocks=4, device=device,
                           weight_dtype=C.DT_BF16, tp_size=tp, tp_rank=rank if tp > 1 else 0,
                           tp_id=tp_id2, sampling=None)
    # prompt seed 4330: the oracle's top-2 margin stays >= 0.044 of max|logit| over these steps
    # (seed 4321 has a 0.0008 near-tie at step 0, inside the 2e-2 gate: an argmax flip there says
    # nothing; found by scanning seeds 4321..4334 with the CPU oracle)
    prompt = [int(t) for t in W.random_prompt(cfg2.vocab_size, 16, 4330)]
    n_steps = 4
    # rank 0 runs the oracle first; every rank then feeds the ORACLE's tokens (teacher forcing:
    # random-init logits are nearly flat, a near-tie flip would make the later steps incomparable)
    box = [None]
    ref_logits = None
    if rank == 0:
        from oracle import oracle as O
        m = O.OracleModel(cfg2, ws)
        ref_toks, ref_logits = m.greedy(np.asarray(prompt, dtype=np.int32), n_steps)
        box[0] = [int(t) for t in ref_toks]
        del m
    if dist is not None:
        dist.broadcast_object_list(box, src=0)
    ref_toks = box[0]
    toks, got = list(prompt), []
    for i in range(n_steps):
        o, lg = r2.run_flat([toks], i == 0, [0], [[0]], [0.0], want_logits=True)
        got.append((int(o[0]), lg[0].copy()))
        toks.append(ref_toks[i])
    r2.close()
    if rank == 0:
        errs = [float(np.abs(got[i][1] - ref_logits[i]).max() / np.abs(ref_logits[i]).max())
                for i in range(n_steps)]
        flips = [i for i in range(n_steps) if got[i][0] != ref_toks[i]]
        # a flip is only forgiven when the oracle's own margin to the GPU's token is inside the gate
        margins = [float((ref_logits[i][ref_toks[i]] - ref_logits[i][got[i][0]]) / np.abs(ref_logits[i]).max())
                   for i in flips]
        out["parity"] = {"max_rel_err": max(errs), "tolerance": 2e-2,
                         "within_tolerance": bool(max(errs) <= 2e-2),
                         "tokens_equal": not flips, "flipped_steps": flips,
                         "flip_margins_rel": margins, "steps": n_steps, "prompt_tokens": len(prompt),
                         "teacher_forced": True,
                         "against": "oracle/nvllm_oracle.c, Falcon-7B width, 2 of 32 layers, same "
                                    "bf16-rounded weights"}
    barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS),
                    help="BASELINE.json configs[N-1]; the default (2) is the config the metric is quoted on")
    ap.add_argument("--model", default=None, choices=sorted(C.PRESETS))
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--prompt-len", type=int, default=None)
    ap.add_argument("--decode-steps", type=int, default=None)
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-tp", action="store_true",
                    help="skip the Falcon-7B tensor-parallel leg (configs[3] at tp = --gpus)")
    args = ap.parse_args()

    D = nvllm_pkg.sub("dist")
    rank, world, local_rank = D.env_ranks()
    spec = dict(CONFIGS[args.config])
    for k, v in (("model", args.model), ("batch", args.batch), ("prompt_len", args.prompt_len),
                 ("decode_steps", args.decode_steps)):
        if v is not None:
            spec[k] = v
            spec["label"] = "custom"
    cfg = C.PRESETS[spec["model"]]()
    need_pos = spec["prompt_len"] + spec["decode_steps"] + 1
    if args.config == 4 and cfg.max_seq_len < need_pos:
        from dataclasses import replace
        cfg = replace(cfg, max_seq_len=need_pos + 31)
    tp, replicas = D.plan(world, spec["tp"])
    config = {"workload": f"{cfg.name} bf16 weights/fp32 accumulate, {spec['label']} "
                          f"(BASELINE.json configs[{args.config - 1}])",
              "global_batch": spec["batch"] * replicas, "prompt_len": spec["prompt_len"],
              "decode_steps": spec["decode_steps"],
              "parallelism": f"tp{tp}" if tp > 1 else f"dp{world} replicas",
              "l2": "inputs larger than L2 (weights + KV streamed per step >> 126 MB)"
                    if cfg.matmul_params() * 2 > 2 * 126e6 else
                    "weights + KV of this model fit the 126 MB L2 only partly; no flush between steps"}
    scaling = "strong" if tp > 1 else "weak"
    metric = "prefill+decode tokens/s (whole job)"

    if args.impl == "reference":
        if rank != 0:
            return
        keep = list(W.random_weights(cfg, 1234, round_bf16=True))
        per_step = max(2.0, min(args.cpu_budget_s, 150.0 / max(1, args.steps + args.warmup)))
        vals = []
        for i in range(args.warmup + args.steps):
            r, _m = cpu_oracle_sample(cfg, keep, per_step)
            del _m
            if i >= args.warmup:
                vals.append(r)
        tok = sum(v["tokens"] for v in vals)
        sec = sum(v["seconds"] for v in vals)
        sample = (f"1 sequence, {vals[-1]['prefill_tokens']}-token prefill + "
                  f"{vals[-1]['decode_tokens']} decode steps per step, full-size {cfg.name} fp32")
        line = {"impl": "reference", "metric": metric, "value": tok / sec, "unit": "tokens/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * sec / len(vals), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": tok / sec, "unit": "tokens/s", "cores": 1, "kind": "port",
                                 "sample": sample,
                                 "note": "restated reference (oracle/nvllm_oracle.c), 1 thread: the Go "
                                         "reference has no goroutines on this path; no Go toolchain here"},
                "e2e": {"value": tok / sec, "unit": "tokens/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": 0, "host_cores": os.cpu_count()}
        print(json.dumps(line))
        return

    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    device = local_rank

    # replicas draw disjoint request sets; the ranks of a tensor-parallel group share one
    job = Job(cfg, spec["batch"], spec["prompt_len"], spec["decode_steps"],
              seed0=D.request_seed(4321, rank, tp), shared_prefix=spec["shared_prefix"])
    # cpu_baseline (and the oracle parity it enables) on rank 0 at N = 1 only
    keep = [] if (rank == 0 and world == 1 and not args.no_cpu_baseline) else None
    tp_id = None
    if tp > 1:
        tp_id = D.share_tp_id(dist, rank, nvllm_pkg.sub("lib").tp_unique_id)
    run = make_runner(cfg, job, device, keep, tp_size=tp, tp_rank=rank if tp > 1 else 0, tp_id=tp_id)
    run.set_speculation(False)  # the timed legs below choose it explicitly
    mem = run.memory_info()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- value leg: device-resident
    job.stage(run)
    for _ in range(args.warmup):
        job.run_resident(run)
    run.sync()
    sampler = ClockSampler(device)
    barrier()
    sampler.start()
    run.timer_start()
    launches = 0
    for _ in range(args.steps):
        launches += job.run_resident(run)
    ms = run.timer_stop()
    clocks = sampler.stop()
    barrier()
    ms = D.reduce_max(dist, ms, f"cuda:{device}")
    value = D.aggregate_throughput(job.tokens_per_step, replicas, args.steps, ms)

    # ---- e2e leg: host buffers through nvllm_run every step.  The library's speculative next
    # step is OFF here so that every step really copies its inputs host -> device inside the
    # timed region; the same leg with it on (the library default) is reported beside it.
    run.set_speculation(False)
    job.run_e2e(run)  # warm (staging buffers sized)
    barrier()
    t0 = time.perf_counter()
    h2d = d2h = 0
    for _ in range(args.steps):
        a, b, _tok = job.run_e2e(run)
        h2d += a
        d2h += b
    torch.cuda.synchronize(device)
    e2e_s = time.perf_counter() - t0
    e2e_s = D.reduce_max(dist, e2e_s, f"cuda:{device}")
    e2e = D.aggregate_throughput(job.tokens_per_step, replicas, args.steps, e2e_s * 1e3)
    e2e_spec, h2d_spec = None, 0
    if tp == 1:
        run.set_speculation(True)
        job.run_e2e(run)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            a, _b, _tok = job.run_e2e(run)
            h2d_spec += a
        torch.cuda.synchronize(device)
        spec_s = D.reduce_max(dist, time.perf_counter() - t0, f"cuda:{device}")
        e2e_spec = D.aggregate_throughput(job.tokens_per_step, replicas, args.steps, spec_s * 1e3)
        run.set_speculation(False)

    # ---- separate profiled pass: per-kernel-class time, phase split, roofline
    peaks = load_peaks()
    kernels, roofline, phases = {}, None, {}
    if not args.no_profile:
        # tp == 1: device timeline (mode 2) — launches stay back to back with their PDL edges and
        # each is charged the time from its predecessor's last exit to its own, so the classes
        # add up to the phase's device time.  tp > 1: CUDA events around every launch (mode 1),
        # the only instrument that sees the NCCL calls; serialised, an upper bound.
        prof_mode = 2 if tp == 1 else 1
        job.stage(run)
        run.set_profile(prof_mode)
        run.rerun()
        prof_prefill = run.get_profile()
        run.set_profile(prof_mode)
        for _ in range(job.D):
            run.run_resident()
        prof_decode = run.get_profile()
        run.set_profile(0)
        # un-instrumented phase times
        job.stage(run)
        run.sync()
        run.timer_start()
        run.rerun()
        t_prefill = run.timer_stop()
        run.timer_start()
        for _ in range(job.D):
            run.run_resident()
        t_decode = run.timer_stop()
        new_tok = int(job.n * job.P - job.num_cached.sum())
        phases = {"prefill_ms": t_prefill, "decode_ms": t_decode,
                  "prefill_tok_s": job.n * job.P / (t_prefill / 1e3),
                  "prefill_computed_tok_s": new_tok / (t_prefill / 1e3),
                  "prefix_cached_tokens": int(job.num_cached.sum()),
                  "decode_tok_s": job.n * job.D / (t_decode / 1e3) if job.D else None,
                  "decode_us_per_step": 1e3 * t_decode / job.D if job.D else None,
                  # whole-phase roofline fractions on ALGORITHMIC work: prefill
                  # flops against the sustained bf16 tensor peak, decode bytes (weights + K/V +
                  # logits + one pass over each activation) against the measured HBM peak
                  "prefill_frac_of_tensor_peak":
                      sum(e["flops"] for e in prof_prefill.values()) / (t_prefill * 1e9) /
                      peaks["bf16_tflops_sustained"],
                  "decode_frac_of_hbm_peak":
                      (sum(e["bytes"] for e in prof_decode.values()) / (t_decode * 1e6) /
                       peaks["hbm_gbs"]) if job.D else None,
                  "profile_instrument": "device timeline (globaltimer, PDL intact)"
                                        if prof_mode == 2 else "CUDA events per launch (serialised)",
                  # the profiled pass's own total next to the un-instrumented phase time
                  "profiled_prefill_ms": sum(e["ms"] for e in prof_prefill.values()),
                  "profiled_decode_ms": sum(e["ms"] for e in prof_decode.values())}
        best = None
        for phase, prof in (("prefill", prof_prefill), ("decode", prof_decode)):
            for name, e in prof.items():
                gbs = e["bytes"] / (e["ms"] * 1e6) if e["ms"] > 0 else 0.0
                tfs = e["flops"] / (e["ms"] * 1e9) if e["ms"] > 0 else 0.0
                kernels[f"{phase}.{name}"] = {"ms": round(e["ms"], 3), "launches": e["launches"],
                                              "GB/s": round(gbs, 1), "TFLOP/s": round(tfs, 2),
                                              "alg_MB_per_launch": round(e["bytes"] / e["launches"] / 1e6, 3),
                                              "impl_MB_per_launch": round(e["impl_bytes"] / e["launches"] / 1e6, 3)}
                if best is None or e["ms"] > best[2]["ms"]:
                    best = (phase, name, e)
        phase, name, e = best
        tensor_bound = phase == "prefill" and name in ("gemm_tc5", "attn_prefill")
        per = e["ms"] / e["launches"]
        if tensor_bound:
            ach = e["flops"] / e["launches"] / (per * 1e9)
            peak = peaks["bf16_tflops_sustained"]
            roofline = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s"}
        else:
            ach = e["bytes"] / e["launches"] / (per * 1e6)
            peak = peaks["hbm_gbs"]
            roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s"}
        traffic = None
        try:  # per-launch DRAM bytes of the dominant kernel from the committed ncu --set full capture
            tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            ent = tr.get(cfg.name.lower(), {}).get(f"{phase}.{name}")
            if ent:
                traffic = ent["dram_bytes_per_launch"]
        except (OSError, ValueError, KeyError):
            traffic = None
        roofline.update({"frac": ach / peak, "traffic": traffic, "kernel": f"{phase}.{name}",
                         "avg_launch_us": per * 1e3, "launches": e["launches"],
                         "algorithmic_bytes_per_launch": e["bytes"] / e["launches"],
                         "implementation_bytes_per_launch": e["impl_bytes"] / e["launches"],
                         "instrument": phases["profile_instrument"],
                         "peak_source": peaks["source"],
                         "share_of_step": e["ms"] / sum(x["ms"] for p in (prof_prefill, prof_decode)
                                                        for x in p.values())})

    cpu, parity = None, None
    if rank == 0 and keep is not None:
        r, _m = cpu_oracle_sample(cfg, keep, args.cpu_budget_s)
        del _m
        if tp == 1:
            parity = gpu_parity(run, r)
        cpu = {"value": r["tokens"] / r["seconds"], "unit": "tokens/s", "cores": 1, "kind": "port",
               "sample": f"1 sequence, {r['prefill_tokens']}-token prefill + {r['decode_tokens']} decode "
                         f"steps, full-size {cfg.name} (bf16-rounded weights as fp32)",
               "prefill_tok_s": r["prefill_tok_s"], "decode_tok_s": r["decode_tok_s"],
               "host_cores": os.cpu_count()}
    run.close()
    del keep
    # the one config whose path shards with a collective: Falcon-7B at tp = N, every N incl. 1
    tp_line = None
    if args.config == 2 and not args.no_tp and args.model is None:
        try:
            tp_line = tp_leg(dist, rank, world, device, max(2, min(args.steps, 5)), 3, barrier)
        except Exception as ex:  # the headline line must survive a failure of the extra leg
            tp_line = {"error": f"{type(ex).__name__}: {ex}"[:400]}
    if rank == 0:
        line = {"metric": metric, "value": value, "unit": "tokens/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": h2d // args.steps,
                        "d2h_bytes_per_step": d2h // args.steps,
                        "speculative_next_step": {"value": e2e_spec, "unit": "tokens/s",
                                                  "h2d_bytes_per_step": h2d_spec // args.steps}},
                "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
                "parity": parity, "tp": tp_line,
                "phases": phases, "kernels": kernels, "memory": mem}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
