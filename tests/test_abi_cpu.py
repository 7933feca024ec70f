"""CPU-only: the C-ABI library loads without a GPU and exports every symbol include/peft_b200.h declares, the ctypes
binding table covers exactly those symbols, and compute entry points fail loudly (status code + message) without a
device instead of falling back."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "peft_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(peft_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from peft_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib


def test_header_symbols_exported(lib):
    handle = lib.load()
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in include/peft_b200.h but not exported by libpeftb200.so"


def test_binding_table_matches_header(lib):
    assert sorted(lib.SIGNATURES) == declared_symbols()


def test_version_and_error_string(lib):
    handle = lib.load()
    assert handle.peft_version() >= 100
    assert isinstance(lib.last_error(), str)
    assert lib.launch_count() >= 0


def test_invalid_arguments_are_rejected_without_gpu(lib):
    handle = lib.load()
    # shape validation happens before any CUDA call: status < 0 and a message, never a silent success
    rc = handle.peft_gemm_bf16(None, 0, None, 0, 0, 0, 0, None, None, 0, None, 0, None, 0, 0, None)
    assert rc < 0
    assert "peft_gemm_bf16" in lib.last_error()
    rc = handle.peft_set_gemm_algo(7)
    assert rc < 0


def test_python_ops_refuse_cpu_tensors(lib):
    import torch
    from peft_b200 import ops
    a = torch.zeros(16, 16, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        ops.gemm(a, a, a.clone())
