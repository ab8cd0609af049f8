"""nano-vllm-rs_b200 — B200-native (sm_100a) paged-attention decode hot path of ssvgopal/nano-vllm-rs.

Layout: csrc/ (hand-written CUDA kernels + C++ host runtime behind include/nvllm.h, built into
lib/libnvllm.so), _ffi.py (ctypes loader — fails loudly if the library is missing), ops.py (one
function per kernel entry point, taking torch CUDA tensors), engine.py (Python mirror of the
reference API: Config, SamplingParams, LLMEngine, LLMEngineBuilder, ModelRunner ... over the C++ host).
There is no CPU fallback anywhere in this package."""

__version__ = "0.1.0"
