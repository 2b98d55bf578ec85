#!/usr/bin/env python
"""Extracts the Go declarations the cgo shim (go/cuda_model_runner.go) binds against from the
reference tree and writes them to tests/golden/go_symbols.json.

    python tests/golden/make_go_symbols.py [/root/reference]

No Go toolchain is needed: struct fields, exported package-level identifiers (types, consts,
vars, funcs) and method sets are read with a small line parser.  tests/test_go_shim.py checks
the shim's selectors against the committed JSON, and regenerates it for comparison whenever the
reference tree is present.
"""
import json
import os
import re
import sys

PKGS = {"tensor": "purego/tensor", "nanovllm": "nanovllm"}


def parse_pkg(root, rel):
    structs, exports, methods, interfaces = {}, set(), {}, {}
    d = os.path.join(root, rel)
    for fn in sorted(os.listdir(d)):
        if not fn.endswith(".go") or fn.endswith("_test.go"):
            continue
        lines = open(os.path.join(d, fn), encoding="utf-8").read().split("\n")
        i = 0
        while i < len(lines):
            ln = lines[i]
            m = re.match(r"^type\s+(\w+)\s+struct\s*\{", ln)
            if m:
                name, fields = m.group(1), {}
                i += 1
                while i < len(lines) and not lines[i].startswith("}"):
                    body = lines[i].split("//")[0].strip()
                    fm = re.match(r"^((?:\w+\s*,\s*)*\w+)\s+(\S.*)$", body)
                    if fm:
                        for f in fm.group(1).split(","):
                            fields[f.strip()] = fm.group(2).strip()
                    i += 1
                structs[name] = fields
                exports.add(name)
            else:
                m = re.match(r"^type\s+(\w+)\s+interface\s*\{", ln)
                if m:
                    name, ms = m.group(1), []
                    i += 1
                    while i < len(lines) and not lines[i].startswith("}"):
                        mm = re.match(r"^\s*(\w+)\(", lines[i])
                        if mm:
                            ms.append(mm.group(1))
                        i += 1
                    interfaces[name] = ms
                    exports.add(name)
                else:
                    m = re.match(r"^type\s+(\w+)\s", ln)
                    if m:
                        exports.add(m.group(1))
                    m = re.match(r"^func\s+(\w+)\(", ln)
                    if m:
                        exports.add(m.group(1))
                    m = re.match(r"^func\s+\(\w+\s+\*?(\w+)\)\s+(\w+)\(", ln)
                    if m:
                        methods.setdefault(m.group(1), []).append(m.group(2))
                    if re.match(r"^(const|var)\s*\($", ln):
                        i += 1
                        while i < len(lines) and not lines[i].startswith(")"):
                            mm = re.match(r"^\s*(\w+)\b", lines[i].split("//")[0])
                            if mm:
                                exports.add(mm.group(1))
                            i += 1
                    m = re.match(r"^(const|var)\s+(\w+)\b", ln)
                    if m:
                        exports.add(m.group(2))
            i += 1
    return {"structs": structs,
            "exports": sorted(e for e in exports if e[:1].isupper()),
            "methods": {k: sorted(v) for k, v in sorted(methods.items())},
            "interfaces": interfaces}


def extract(root):
    return {name: parse_pkg(root, rel) for name, rel in PKGS.items()}


if __name__ == "__main__":
    root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "go_symbols.json")
    json.dump(extract(root), open(out, "w"), indent=1, sort_keys=True)
    print("wrote", out)
