import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from nano_vllm_rs_b200 import _ffi
    lib = _ffi.lib()  # raises if the extension is missing: GPU tests must never pass on a fallback
    assert lib.nvllm_device_ok() == 1, "libnvllm.so loaded but no sm_100 device visible"
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def _nvllm_knobs_follow_the_environment(monkeypatch):
    """libnvllm.so parses its NVLLM_* knobs once (nvllm::knobs()); tests that force a kernel variant with
    monkeypatch.setenv / delenv must have the library re-read them — and again once the environment is restored."""
    touched = []

    def reload():
        from nano_vllm_rs_b200 import _ffi
        _ffi.lib().nvllm_reload_env()

    set_orig, del_orig = monkeypatch.setenv, monkeypatch.delenv

    def setenv(name, value, *a, **k):
        set_orig(name, value, *a, **k)
        if name.startswith("NVLLM_"):
            touched.append(name)
            reload()

    def delenv(name, *a, **k):
        del_orig(name, *a, **k)
        if name.startswith("NVLLM_"):
            touched.append(name)
            reload()

    monkeypatch.setenv, monkeypatch.delenv = setenv, delenv
    yield
    if touched:
        monkeypatch.undo()
        reload()
