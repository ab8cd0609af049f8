"""CPU tests: the C-ABI library loads without a GPU and exports every symbol include/*.h declares; the ctypes
prototype tables cover the same set; argument errors come back as codes + messages (no compute calls here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(nvllm_[a-z0-9_]+)\s*\(", text))


def test_library_exports_every_declared_symbol():
    from nano_vllm_rs_b200 import _ffi
    from nano_vllm_rs_b200 import engine as E
    lib = E.lib()
    kernels, engine = declared_symbols("nvllm.h"), declared_symbols("nvllm_engine.h")
    assert len(kernels) >= 25 and len(engine) >= 30
    for sym in sorted(kernels | engine):
        assert hasattr(lib, sym), f"{sym} declared in include/ but not exported by libnvllm.so"
    assert kernels == set(_ffi.PROTOTYPES), kernels ^ set(_ffi.PROTOTYPES)
    assert engine == set(E.ENGINE_PROTOTYPES), engine ^ set(E.ENGINE_PROTOTYPES)


def test_version_and_error_convention_without_gpu():
    from nano_vllm_rs_b200 import _ffi
    lib = _ffi.lib()
    assert lib.nvllm_version() == 100
    assert lib.nvllm_device_ok() in (0, 1)
    lay = _ffi.KvLayout.reference(256, 8, 100)   # head_dim 100 is not supported
    rc = lib.nvllm_store_kvcache(None, None, None, None, None, 4, 1024, lay, 0, None)
    assert rc == -1 and b"head_dim" in lib.nvllm_last_error()
    rc = lib.nvllm_add_rmsnorm(None, None, None, None, 4, 1001, 1e-6, 0, None)
    assert rc == -1 and b"hidden" in lib.nvllm_last_error()
    rc = lib.nvllm_gemm_bf16(None, None, None, 4, 8, 100, 100, 100, 8, 0, None, 0, None)
    assert rc == -1 and b"k % 64" in lib.nvllm_last_error()
    rc = lib.nvllm_silu_mul(None, None, 2, 8, 1, None)   # dtype f32 unsupported
    assert rc == -3
    with pytest.raises(_ffi.NvllmError):
        _ffi.check(rc)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nano-vllm-rs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src.replace("oracle/weights.py", "").replace("oracle/host.py", "").replace("oracle/qwen3.py", "").replace("oracle/layers.py", ""), f
