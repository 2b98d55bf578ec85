#!/usr/bin/env python
"""profiles/ncu_traffic.json from the committed `ncu --set full` summaries (profiles/r2_full_*.md):
per workload and kernel class, the measured DRAM bytes per launch (dram__bytes_read.sum +
dram__bytes_write.sum), averaged over the launches of that class in the captured window, next to
the launch count and the mean duration under ncu.  bench.py reports the entry of its dominant
kernel as `roofline.traffic`.

    python tools/ncu_traffic.py > profiles/ncu_traffic.json
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")
# summary file -> (workload name as bench.py's cfg.name, phase)
FILES = {"r2_full_cfg2_prefill.md": ("llama-3.2-1b", "prefill", "batch 64 x 512-token prompts"),
         "r2_full_cfg2_decode.md": ("llama-3.2-1b", "decode", "batch 64, context ~514"),
         "r2_full_cfg1_decode.md": ("gpt2", "decode", "1 sequence, context ~34"),
         "r2_full_cfg3_decode.md": ("granite-moe-1b", "decode", "batch 128, context ~130"),
         "r2_full_cfg5_prefill.md": ("llama-3.2-3b", "prefill", "4 x 4096-token prompts, 2 layers"),
         "r2_full_cfg4_falcon.md": ("falcon-7b", "mixed", "1 x 512-token prompt + decode steps, 2 layers")}
CLASS = [("gemm_tc5_kernel", "gemm_tc5"), ("gemv_kernel", "gemv"), ("norm_kernel", "norm"),
         ("attn_prefill", "attn_prefill"), ("attn_decode", "attn_decode"), ("attn_combine", "attn_decode"),
         ("rope_kv", "rope_kv"), ("embed", "embed"), ("act_kernel", "act"), ("sample", "sample"),
         ("moe_sort", "moe_route"), ("moe_gather", "moe_route"), ("advance", "embed"),
         ("make_decode_state", "embed"), ("reduce_partials", "misc")]


def cls(kernel):
    for pat, c in CLASS:
        if kernel.startswith(pat):
            return c
    return "misc"


def main():
    out = {}
    for fn, (wl, phase, what) in FILES.items():
        path = os.path.join(P, fn)
        if not os.path.exists(path):
            continue
        lines = [l for l in open(path) if l.startswith("|")]
        head = [h.strip() for h in lines[0].strip().strip("|").split("|")]
        col = {re.sub(r" \[.*$", "", h): i for i, h in enumerate(head)}
        unit = {re.sub(r" \[.*$", "", h): (re.search(r"\[(.*)\]", h).group(1) if "[" in h else "") for h in head}
        agg = {}
        for l in lines[2:]:
            f = [x.strip() for x in l.strip().strip("|").split("|")]
            k = f[col["kernel"]].strip("`")
            def val(name):
                v = float(f[col[name]].replace(",", ""))
                u = unit[name].lower()
                return v * {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "byte": 1.0, "us": 1.0,
                            "ms": 1e3, "ns": 1e-3}.get(u, 1.0)
            b = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
            t = val("gpu__time_duration.sum")
            a = agg.setdefault(cls(k), {"launches": 0, "bytes": 0.0, "us": 0.0, "kernels": set()})
            a["launches"] += 1
            a["bytes"] += b
            a["us"] += t
            a["kernels"].add(re.sub(r"<.*$", "", k))
        for c, a in agg.items():
            key = f"{phase}.{c}" if phase != "mixed" else c
            out.setdefault(wl, {})[key] = {
                "dram_bytes_per_launch": round(a["bytes"] / a["launches"]),
                "launches_in_window": a["launches"],
                "avg_us_under_ncu": round(a["us"] / a["launches"], 2),
                "kernels": sorted(a["kernels"]), "window": what, "source": f"profiles/{fn}"}
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
