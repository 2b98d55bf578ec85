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
E = nvllm_pkg.sub("engine")


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


class Job:
    """The cfg2 job on one replica: block tables from the restated Go block manager."""

    def __init__(self, cfg, batch, prompt_len, decode_steps, seed0):
        self.cfg, self.n, self.P, self.D = cfg, batch, prompt_len, decode_steps
        self.max_len = prompt_len + decode_steps + 1
        self.prompts = np.stack([W.random_prompt(cfg.vocab_size, prompt_len, seed0 + i)
                                 for i in range(batch)])
        nb = (self.max_len + 255) // 256
        bm = E.BlockManager(batch * nb, 256)
        self.tables = []
        for i in range(batch):  # allocate the final extent up front (ABI allows nb >= needed)
            s = E.Sequence(list(range(i * 7, i * 7 + self.max_len)))
            bm.allocate(s)
            self.tables.append(s.block_table)
        self.num_blocks = batch * nb
        self.bt = np.asarray(self.tables, dtype=np.int32).reshape(-1)
        self.cub = np.arange(0, (batch + 1) * nb, nb, dtype=np.int32)
        self.temp = np.zeros(batch, dtype=np.float32)
        self.tokens_per_step = batch * (prompt_len + decode_steps)

    # ---- device-resident pass: nothing crosses PCIe inside
    def run_resident(self, run):
        launches = 0
        run.rerun()
        launches += run.last_stats.kernel_launches
        for _ in range(self.D):
            run.run_resident()
            launches += run.last_stats.kernel_launches
        return launches

    def stage(self, run):
        """Upload the prefill step's inputs once (outside any timed region)."""
        flat = self.prompts.reshape(-1).astype(np.int32)
        cu = np.arange(0, (self.n + 1) * self.P, self.P, dtype=np.int32)
        run.run_arrays(self.n, True, flat, cu, np.zeros(self.n, np.int32), self.bt, self.cub,
                       self.temp, None)

    # ---- end-to-end pass through nvllm_run with host buffers every step
    def run_e2e(self, run):
        n, P = self.n, self.P
        tok = np.zeros((n, self.max_len), dtype=np.int32)
        tok[:, :P] = self.prompts
        lens = np.full(n, P, dtype=np.int64)
        cols = np.arange(self.max_len)[None, :]
        h2d = d2h = 0
        cu = np.zeros(n + 1, dtype=np.int32)
        nc = np.zeros(n, np.int32)
        for step in range(self.D + 1):
            np.cumsum(lens, out=cu[1:])
            flat = tok[cols < lens[:, None]]
            out, _ = run.run_arrays(n, step == 0, flat, cu, nc if step == 0 else None, self.bt,
                                    self.cub, self.temp, None)
            h2d += run.last_stats.h2d_bytes
            d2h += run.last_stats.d2h_bytes
            tok[np.arange(n), lens] = out
            lens += 1
        return h2d, d2h, tok


def make_runner(cfg, job, device, keep_host=None):
    R = nvllm_pkg.sub("runner")

    def gen():
        for wid, layer, a in W.random_weights(cfg, 1234, round_bf16=True):
            if keep_host is not None:
                keep_host.append((wid, layer, a))
            yield wid, layer, a

    return R.CudaModelRunner(cfg, gen(), precision=C.PREC_BF16, num_kv_blocks=job.num_blocks,
                             device=device, max_num_seqs=job.n,
                             max_batched_tokens=job.n * job.P, weight_dtype=C.DT_BF16)


def cpu_oracle_sample(cfg, weights, budget_s, prompt_tokens=None):
    """Time the restated reference (1 thread): prefill `p` tokens + decode `d` tokens of one
    sequence, sized to ~budget_s after a 2-token calibration."""
    from oracle import oracle as O
    m = O.OracleModel(cfg, weights)
    prompt = W.random_prompt(cfg.vocab_size, 64, 4321)
    t0 = time.perf_counter()
    m.forward(prompt[:2], m.new_cache())
    t_tok = (time.perf_counter() - t0) / 2
    total = int(max(4, min(48, budget_s / max(t_tok, 1e-3))))
    p = prompt_tokens if prompt_tokens else max(2, (2 * total) // 3)
    d = max(1, total - p)
    cache = m.new_cache()
    t0 = time.perf_counter()
    logits = m.forward(prompt[:p], cache)
    t_prefill = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in range(d):
        nt = O.argmax(logits)
        logits = m.forward([nt], cache)
    t_decode = time.perf_counter() - t0
    return {"tokens": p + d, "seconds": t_prefill + t_decode, "prefill_tokens": p,
            "decode_tokens": d, "prefill_tok_s": p / t_prefill, "decode_tok_s": d / t_decode}, m


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama-3.2-1b", choices=sorted(C.PRESETS))
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prompt-len", type=int, default=512)
    ap.add_argument("--decode-steps", type=int, default=256)
    ap.add_argument("--cpu-budget-s", type=float, default=20.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = C.PRESETS[args.model]()
    config = {"workload": f"{cfg.name} bf16 weights/fp32 accumulate, batch {args.batch} continuous "
                          f"batching, {args.prompt_len}-token prefill + {args.decode_steps} greedy "
                          f"decode steps (BASELINE.json configs[1])",
              "global_batch": args.batch * world, "prompt_len": args.prompt_len,
              "decode_steps": args.decode_steps, "parallelism": f"dp{world} replicas",
              "l2": "inputs larger than L2 (2.5 GB of weights + KV per step >> 126 MB)"}
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

    job = Job(cfg, args.batch, args.prompt_len, args.decode_steps, seed0=4321 + 1000 * rank)
    keep = [] if (rank == 0 and not args.no_cpu_baseline) else None
    run = make_runner(cfg, job, device, keep)
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
    if dist is not None:
        t = torch.tensor([ms], device=f"cuda:{device}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * job.tokens_per_step * args.steps / (ms / 1e3)

    # ---- e2e leg: host buffers through nvllm_run every step
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
    if dist is not None:
        t = torch.tensor([e2e_s], device=f"cuda:{device}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = world * job.tokens_per_step * args.steps / e2e_s

    # ---- separate profiled pass: per-kernel-class time, phase split, roofline
    peaks = load_peaks()
    kernels, roofline, phases = {}, None, {}
    if not args.no_profile:
        job.stage(run)
        run.set_profile(True)
        run.rerun()
        prof_prefill = run.get_profile()
        run.set_profile(True)
        for _ in range(job.D):
            run.run_resident()
        prof_decode = run.get_profile()
        run.set_profile(False)
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
        phases = {"prefill_ms": t_prefill, "decode_ms": t_decode,
                  "prefill_tok_s": job.n * job.P / (t_prefill / 1e3),
                  "decode_tok_s": job.n * job.D / (t_decode / 1e3)}
        best = None
        for phase, prof in (("prefill", prof_prefill), ("decode", prof_decode)):
            for name, e in prof.items():
                gbs = e["bytes"] / (e["ms"] * 1e6) if e["ms"] > 0 else 0.0
                tfs = e["flops"] / (e["ms"] * 1e9) if e["ms"] > 0 else 0.0
                kernels[f"{phase}.{name}"] = {"ms": round(e["ms"], 3), "launches": e["launches"],
                                              "GB/s": round(gbs, 1), "TFLOP/s": round(tfs, 2)}
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
        roofline.update({"frac": ach / peak, "traffic": None, "kernel": f"{phase}.{name}",
                         "avg_launch_us": per * 1e3, "launches": e["launches"],
                         "peak_source": peaks["source"],
                         "share_of_step": e["ms"] / sum(x["ms"] for p in (prof_prefill, prof_decode)
                                                        for x in p.values())})

    cpu = None
    if rank == 0 and keep is not None:
        run.close()
        r, _m = cpu_oracle_sample(cfg, keep, args.cpu_budget_s)
        cpu = {"value": r["tokens"] / r["seconds"], "unit": "tokens/s", "cores": 1, "kind": "port",
               "sample": f"1 sequence, {r['prefill_tokens']}-token prefill + {r['decode_tokens']} decode "
                         f"steps, full-size {cfg.name} (bf16-rounded weights as fp32)",
               "prefill_tok_s": r["prefill_tok_s"], "decode_tok_s": r["decode_tok_s"],
               "host_cores": os.cpu_count()}
    if rank == 0:
        line = {"metric": metric, "value": value, "unit": "tokens/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e, "unit": "tokens/s", "h2d_bytes_per_step": h2d // args.steps,
                        "d2h_bytes_per_step": d2h // args.steps},
                "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
                "phases": phases, "kernels": kernels, "memory": mem}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
