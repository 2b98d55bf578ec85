"""CPU checks of the drop-in boundary: libnvllm_b200.so loads and exports every symbol that
include/nvllm.h declares (no compute calls: there is no GPU here), the ctypes mirror of
``struct nvllm_config`` matches the C layout, and the product package never imports the oracle."""
import ctypes
import os
import re
import subprocess

import pytest

from helpers import C, ROOT
import nvllm_pkg

HEADER = os.path.join(ROOT, "include", "nvllm.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nvllm_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ["nvllm_create", "nvllm_set_weight", "nvllm_finalize", "nvllm_run", "nvllm_destroy",
              "nvllm_last_error", "nvllm_run_resident", "nvllm_op_matmul"]:
        assert s in syms


def test_library_exports_every_declared_symbol():
    L = nvllm_pkg.sub("lib")
    lib = L.load()
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/nvllm.h but not exported"
    # and the python binding types every one of them
    assert set(declared_symbols()) == set(L.SIGNATURES)
    assert lib.nvllm_abi_version() == C.ABI_VERSION


def test_config_struct_layout_matches_c(tmp_path):
    """Compile a tiny C program against the header and compare sizeof/offsetof."""
    fields = [f for f, _ in C.CConfig._fields_ if f != "reserved"]
    prog = '#include <stdio.h>\n#include <stddef.h>\n#include "nvllm.h"\nint main(){\n'
    prog += 'printf("%zu\\n", sizeof(nvllm_config));\n'
    for f in fields:
        prog += f'printf("%zu\\n", offsetof(nvllm_config, {f}));\n'
    prog += 'printf("%zu\\n", sizeof(nvllm_run_stats));return 0;}\n'
    src = tmp_path / "t.c"
    src.write_text(prog)
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out[0] == ctypes.sizeof(C.CConfig)
    for f, off in zip(fields, out[1:-1]):
        assert getattr(C.CConfig, f).offset == off, f
    L = nvllm_pkg.sub("lib")
    assert out[-1] == ctypes.sizeof(L.RunStats)


def test_no_gpu_means_loud_failure():
    """There is no CPU fallback: without an sm_100 device nvllm_create must fail with ERR_CUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = nvllm_pkg.sub("lib")
    lib = L.load()
    cc = C.tiny_gpt2().to_c(C.PREC_FP32)
    h = ctypes.c_void_p()
    rc = lib.nvllm_create(ctypes.byref(cc), None, ctypes.byref(h))
    assert rc == L.ERR_CUDA and not h
    assert L.last_error()


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "nano-vllm-go_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("the oracle", "").replace("CPU oracle", ""), \
                    f"{f} references the oracle"


def test_noalloc_entry_point_source_allocates_nothing():
    """nvllm_op_gemv_noalloc (csrc/noalloc.cu): static check of what its header comment promises —
    scratch in a file-scope static __device__ array of the size the header states, its address
    taken with cudaGetSymbolAddress, and no allocating call anywhere in the file."""
    path = os.path.join(os.path.dirname(HEADER), "..", "nano-vllm-go_b200", "csrc", "noalloc.cu")
    src = open(path).read()
    code = re.sub(r"//[^\n]*", "", src)
    code = re.sub(r"/\*.*?\*/", "", code, flags=re.S)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "malloc(", "new ",
                   "std::vector", "cudaStreamCreate", "cudaEventCreate"):
        assert banned not in code, banned
    assert re.search(r"static\s+__device__\s+__align__\(256\)\s+float\s+\w+\[NOALLOC_FLOATS\]", code)
    assert "cudaGetSymbolAddress" in code
    hdr = open(HEADER).read()
    m = re.search(r"#define\s+NVLLM_NOALLOC_MAX_SCRATCH_FLOATS\s+\(1\s*<<\s*(\d+)\)", hdr)
    assert m and int(m.group(1)) == 20  # 4 MiB: the limit DESIGN.md and INTEGRATION.md state
