"""ctypes loader for lib/libnvllm.so (the C ABI declared in include/nvllm.h).

The product path must fail loudly when the CUDA extension is missing: `lib()` raises, nothing here
falls back to PyTorch or to the CPU oracle."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NVLLM_LIB_PATH: debug builds only (e.g. lib/libnvllm_timeline.so for tools/gemm_timeline.py); a missing file still raises
LIB_PATH = os.environ.get("NVLLM_LIB_PATH") or os.path.join(_HERE, "lib", "libnvllm.so")
_lib = None

i32, i64, u64, f32, f64, vp, sz = C.c_int32, C.c_int64, C.c_uint64, C.c_float, C.c_double, C.c_void_p, C.c_size_t


class KvLayout(C.Structure):
    """nvllm_kv_layout (include/nvllm.h)."""
    _fields_ = [("block_size", i32), ("num_kv_heads", i32), ("head_dim", i32), ("reserved", i32),
                ("block_stride", i64), ("token_stride", i64), ("head_stride", i64)]

    @staticmethod
    def reference(block_size, num_kv_heads, head_dim):
        """[num_blocks, block_size, Hkv, D] — the reference's allocation (model_runner.rs:379-389)."""
        return KvLayout(block_size, num_kv_heads, head_dim, 0, block_size * num_kv_heads * head_dim,
                        num_kv_heads * head_dim, head_dim)

    @staticmethod
    def head_major(block_size, num_kv_heads, head_dim):
        """[num_blocks, Hkv, block_size, D] — the engine's layout (64 KiB contiguous per block/head)."""
        return KvLayout(block_size, num_kv_heads, head_dim, 0, block_size * num_kv_heads * head_dim,
                        head_dim, block_size * head_dim)


LP = C.POINTER(KvLayout)

# name -> (restype, argtypes); every symbol include/nvllm.h declares must appear here
# (tests/test_abi.py checks the header against this table and against the built library).
PROTOTYPES = {
    "nvllm_version": (i32, []),
    "nvllm_last_error": (C.c_char_p, []),
    "nvllm_device_ok": (i32, []),
    "nvllm_reload_env": (i32, []),
    "nvllm_store_kvcache": (i32, [vp, vp, vp, vp, vp, i32, i64, LP, i32, vp]),
    "nvllm_paged_decode_workspace_bytes": (sz, [i32, i32, i32, i32]),
    "nvllm_paged_decode_choose_splits": (i32, [i32, i32, i32, i32]),
    "nvllm_paged_decode_attention": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i64, LP, f32, vp, sz, i32, i32, vp]),
    "nvllm_paged_decode_static_scratch_bytes": (sz, []),
    "nvllm_paged_decode_attention_static": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i64, i64, LP, f32, i32, i32, vp]),
    "nvllm_paged_decode_attention_fused": (i32, [vp, vp, vp, vp, vp, vp, f32, i32, vp, vp, vp, vp, vp, i32, i32, i32, i64, i64,
                                                 LP, f32, vp, sz, i32, i32, vp]),
    "nvllm_varlen_prefill_attention": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i64, i64, i64,
                                             vp, vp, vp, i32, LP, f32, i32, vp]),
    "nvllm_gemm_workspace_bytes": (sz, [i32, i32, i32]),
    "nvllm_gemm_bf16": (i32, [vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, vp, sz, vp]),
    "nvllm_gemm_silu_mul_bf16": (i32, [vp, vp, vp, i32, i32, i32, i64, i64, i64, vp, sz, vp]),
    "nvllm_gemm_qkv_rope_store": (i32, [vp, vp, vp, i32, i32, i64, i64, vp, vp, vp, vp, f32, i32, i32, i32, i32, vp, vp, vp, LP, vp]),
    "nvllm_gemm_bf16_simt": (i32, [vp, vp, vp, i32, i32, i32, i64, i64, i64, i32, vp]),
    "nvllm_add_rmsnorm": (i32, [vp, vp, vp, vp, i32, i32, f32, i32, vp]),
    "nvllm_qknorm_rope_store": (i32, [vp, vp, vp, vp, vp, f32, i32, i32, i32, i32, i32, vp, vp, vp, LP, i32, vp]),
    "nvllm_rope_table": (i32, [vp, i32, i32, f64, vp]),
    "nvllm_silu_mul": (i32, [vp, vp, i32, i32, i32, vp]),
    "nvllm_embedding_gather": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "nvllm_gather_last_tokens": (i32, [vp, vp, vp, i32, i32, i32, vp]),
    "nvllm_argmax": (i32, [vp, vp, vp, i32, i32, i64, i64, i32, vp]),
    "nvllm_argmax_combine": (i32, [vp, vp, vp, i32, i32, vp]),
    "nvllm_assemble_vocab_shards": (i32, [vp, vp, i32, i32, i32, vp]),
    "nvllm_sample": (i32, [vp, vp, i32, i32, i64, vp, vp, vp, vp, vp, vp, vp]),
    "nvllm_comm_unique_id": (i32, [vp]),
    "nvllm_comm_init": (i32, [C.POINTER(vp), vp, i32, i32]),
    "nvllm_comm_destroy": (i32, [vp]),
    "nvllm_allreduce": (i32, [vp, vp, sz, i32, vp]),
    "nvllm_allgather": (i32, [vp, vp, vp, sz, vp]),
    "nvllm_comm_symm_alloc": (i32, [vp, sz, C.POINTER(vp)]),
    "nvllm_allreduce_add_rmsnorm": (i32, [vp, vp, vp, vp, vp, i32, i32, f32, i32, vp]),
    "nvllm_comm_error": (i32, [vp]),
    "nvllm_comm_ll_applies": (i32, [vp, i32, i32]),
    "nvllm_gemm_bf16_push": (i32, [vp, vp, vp, vp, i32, i32, i32, i64, i64, i64, vp]),
    "nvllm_allreduce_add_rmsnorm_pushed": (i32, [vp, vp, vp, vp, vp, i32, i32, f32, i32, vp]),
    "nvllm_init_weight": (i32, [vp, i64, u64, u64, f32, i32, vp]),
}


class NvllmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nvllm error {code}: {msg}")
        self.code = code


def lib():
    """Load libnvllm.so once.  Raises if it has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                               f"g.build()'` (make -C nano-vllm-rs_b200/csrc). There is no fallback path.")
        l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(l, name, None)
            if fn is None:
                continue  # reported by tests/test_abi.py; calling it raises AttributeError
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        raise NvllmError(rc, lib().nvllm_last_error().decode(errors="replace"))
    return rc
