"""Import shim: the package directory is named after the reference crate (`nano-vllm-rs_b200/`,
with a hyphen), which Python cannot import by name.  This module turns itself into that package
(`__path__` points at the directory), so `import nano_vllm_rs_b200.ops` etc. work from the repo root."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "nano-vllm-rs_b200")]
__file__ = _os.path.join(__path__[0], "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
