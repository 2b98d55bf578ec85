"""Static check of the cgo shim go/cuda_model_runner.go (no Go toolchain in this image).

Every selector the shim uses on a reference type — model.X, mc.X, engineCfg.X, b.X, b.X.Y,
a.X inside the attention type switch, s.X, t.X, tensor.X — must exist in the reference's Go
declarations (tests/golden/go_symbols.json, extracted by tests/golden/make_go_symbols.py), every
C.X must be declared by include/nvllm.h, and the shim's method set must cover the reference's
ModelRunner interface plus the TensorModelRunner extras.  When /root/reference is present the
committed JSON is also checked against a fresh extraction.
"""
import importlib.util
import json
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "go", "cuda_model_runner.go")
SYMS = os.path.join(ROOT, "tests", "golden", "go_symbols.json")
HDR = os.path.join(ROOT, "include", "nvllm.h")
REF = "/root/reference"

# variable -> (package, struct) as declared in the shim
VAR_TYPES = {
    "model": ("tensor", "TransformerModel"),
    "mc": ("tensor", "ModelConfig"),
    "engineCfg": ("nanovllm", "Config"),
    "b": ("tensor", "GenericBlock"),
    "s": ("nanovllm", "Sequence"),
    "t": ("tensor", "Tensor"),
}
CGO_BUILTINS = {"int", "float", "double", "int32_t", "uint64_t", "GoString", "uint8_t", "CString", "free"}


def _code():
    src = open(SHIM, encoding="utf-8").read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)          # cgo preamble / block comments
    return "\n".join(ln.split("//")[0] for ln in src.split("\n"))


def _syms():
    return json.load(open(SYMS))


def _field_struct(syms, pkg, struct, field):
    """struct type name a field points at (same package), or None for scalars / slices."""
    ty = syms[pkg]["structs"][struct][field].lstrip("*[]")
    return ty if ty in syms[pkg]["structs"] else None


def test_fixture_matches_reference_tree():
    if not os.path.isdir(REF):
        pytest.skip("reference tree not present on this box")
    spec = importlib.util.spec_from_file_location(
        "make_go_symbols", os.path.join(ROOT, "tests", "golden", "make_go_symbols.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    fresh = json.loads(json.dumps(mod.extract(REF), sort_keys=True))
    assert fresh == _syms(), "tests/golden/go_symbols.json is stale: rerun make_go_symbols.py"


def test_struct_selectors_exist():
    code, syms = _code(), _syms()
    checked = 0
    for var, (pkg, struct) in VAR_TYPES.items():
        for m in re.finditer(rf"(?<![\w.]){var}\.([A-Z]\w*)(?:\.([A-Z]\w*))?", code):
            f1, f2 = m.group(1), m.group(2)
            fields = syms[pkg]["structs"][struct]
            assert f1 in fields, f"{var}.{f1}: {pkg}.{struct} has no field {f1}"
            checked += 1
            if f2:
                sub = _field_struct(syms, pkg, struct, f1)
                assert sub is not None, f"{var}.{f1}.{f2}: {f1} is not a struct"
                assert f2 in syms[pkg]["structs"][sub], f"{var}.{f1}.{f2}: {pkg}.{sub} has no field {f2}"
                checked += 1
    assert checked > 60  # the shim touches every weight of every block type


def test_attention_type_switch_selectors():
    """`switch a := b.Attention.(type)`: a.X is checked against the struct named by its case."""
    code, syms = _code(), _syms()
    cases = list(re.finditer(r"case \*tensor\.(\w+):", code))
    assert {c.group(1) for c in cases} == {"MultiHeadAttention", "GroupedQueryAttention",
                                           "MultiQueryAttention", "Mamba2Layer"}
    for i, c in enumerate(cases):
        end = cases[i + 1].start() if i + 1 < len(cases) else code.index("if b.MoE != nil")
        body = code[c.end():end]
        fields = syms["tensor"]["structs"][c.group(1)]
        used = re.findall(r"(?<![\w.])a\.([A-Z]\w*)", body)
        assert used, c.group(1)
        for f in used:
            assert f in fields, f"a.{f}: tensor.{c.group(1)} has no field {f}"
    # every weight-bearing field of the three attention structs is sent
    # Mamba2Layer: XProj / DtProj are never read by Forward (mamba2.go:72-207), ConvCache / SSMState
    # are run-time state
    skip = {"Mamba2Layer": {"XProj", "DtProj", "ConvCache", "SSMState"}}
    for name in ("MultiHeadAttention", "GroupedQueryAttention", "MultiQueryAttention", "Mamba2Layer"):
        tensors = [f for f, ty in syms["tensor"]["structs"][name].items()
                   if ty == "*Tensor" and f not in skip.get(name, ())]
        body = code[code.index(f"case *tensor.{name}:"):]
        for f in tensors:
            assert re.search(rf"a\.{f}\b", body), f"{name}.{f} is never handed to the library"


def test_package_identifiers_exist():
    code, syms = _code(), _syms()
    for ident in set(re.findall(r"(?<![\w.])tensor\.([A-Z]\w*)", code)):
        assert ident in syms["tensor"]["exports"], f"tensor.{ident} is not exported by the reference"
    # unqualified reference types of package nanovllm used in signatures
    for ident in ("Sequence", "Config"):
        assert ident in syms["nanovllm"]["exports"]
    assert syms["nanovllm"]["interfaces"]["ModelRunner"] == ["Run", "Close"]


def test_c_identifiers_are_declared_by_the_header():
    code = _code()
    hdr = open(HDR).read()
    for ident in set(re.findall(r"(?<![\w.])C\.(\w+)", code)):
        if ident in CGO_BUILTINS:
            continue
        assert re.search(rf"\b{ident}\b", hdr), f"C.{ident} is not declared in include/nvllm.h"


def test_method_set_covers_the_reference_runner():
    code, syms = _code(), _syms()
    have = set(re.findall(r"func \(r \*CudaModelRunner\) (\w+)\(", code))
    want = set(syms["nanovllm"]["interfaces"]["ModelRunner"]) | \
        set(syms["nanovllm"]["methods"]["TensorModelRunner"]) | {"GetVocabSize", "GetModelConfig"}
    assert want <= have, f"missing methods: {sorted(want - have)}"
    # Run's signature is the interface's (model_runner.go:11)
    assert re.search(r"func \(r \*CudaModelRunner\) Run\(seqs \[\]\*Sequence, isPrefill bool\) "
                     r"\(\[\]int, error\)", code)
    assert re.search(r"func \(r \*CudaModelRunner\) Close\(\) error", code)


def test_defaults_match_new_tensor_model_runner():
    """NewTensorModelRunner -> DefaultSamplingParams: temperature 1.0, top-p 1.0, top-k 0,
    repetition penalty 1.2 (tensor_model_runner.go:27-31, sampling.go:17-24)."""
    code = _code()
    assert re.search(r"temperature:\s*1\.0", code)
    assert re.search(r"nvllm_set_sampling\(h, C\.float\(1\.0\), C\.int\(0\), C\.float\(1\.2\)\)", code)
    # the python mirror keeps the same defaults
    import nvllm_pkg
    R = nvllm_pkg.sub("runner")
    assert R.DEFAULT_SAMPLING == dict(temperature=1.0, top_p=1.0, top_k=0, repetition_penalty=1.2)


# ---- go/engine_fixes.go: same static discipline
FIXES = os.path.join(ROOT, "go", "engine_fixes.go")
FIX_VARS = {  # variable -> struct of package nanovllm
    "seq": "Sequence", "last": "Sequence", "v": "Sequence", "s": "Scheduler", "config": "Config",
}


def _fix_code():
    src = open(FIXES, encoding="utf-8").read()
    return "\n".join(ln.split("//")[0] for ln in src.split("\n"))


def test_engine_fixes_selectors_exist_in_the_reference():
    code, syms = _fix_code(), _syms()["nanovllm"]
    for var, struct in FIX_VARS.items():
        members = set(syms["structs"][struct]) | set(syms["methods"].get(struct, []))
        for sel in set(re.findall(rf"(?<![\w.]){var}\.([A-Za-z]\w*)", code)):
            assert sel in members, f"{var}.{sel}: {struct} has no such field or method"
    # chained selectors on the scheduler's block manager and queues
    for sel in set(re.findall(r"s\.blockManager\.(\w+)", code)):
        assert sel in syms["methods"]["BlockManager"], f"BlockManager.{sel} does not exist"
    for ident in ("NewScheduler", "StatusRunning", "StatusFinished", "Sequence", "Scheduler", "Config"):
        assert ident in syms["exports"]
    # the runner methods the fixes call exist in the shim
    shim = _code()
    for meth in set(re.findall(r"(?<![\w.])r\.(\w+)\(", code)):
        assert re.search(rf"func \(r \*CudaModelRunner\) {meth}\(", shim) or \
            re.search(rf"func \(r \*CudaModelRunner\) {meth}\(", code), f"CudaModelRunner.{meth} is not defined"


def test_engine_fixes_match_the_python_mirror():
    """Both files must implement the same four fixes (names kept in step by hand)."""
    code = _fix_code()
    mirror = open(os.path.join(ROOT, "harness", "engine_fixes.py")).read()
    for go_name, py_name in (("ScheduleMixed", "schedule_mixed"), ("PostprocessMixed", "postprocess_mixed"),
                             ("GenerateByRequest", "def generate"), ("CheckBlockSize", "check_block_size"),
                             ("NumKVBlocksFromFreeHBM", "num_kv_blocks_from_free_hbm"),
                             ("blockReady", "block_ready"), ("allocLen", "alloc_len")):
        assert go_name in code and py_name in mirror
