"""-m gpu test of nvllm_paged_decode_attention_static, the split-K decode attention entry point that allocates nothing.

Some users of libnvllm.so reach it from a spot where they cannot get device memory: an operator hook that is handed the
tensors of one op and nothing more, a server whose memory pool is sealed after start-up, a C or Rust host with no CUDA
allocator.  For them the library keeps the f32 split-K partials in an 8 MiB array inside its own device image instead of a
caller-supplied workspace.  This file checks, through the C ABI (ctypes), that the variant gives the same answer as
nvllm_paged_decode_attention on two calls in a row of different shapes (the second one lands on what the first left in
the shared array), that it tells an oversized request apart with NVLLM_ERR_WORKSPACE, and that the two calls really did
not allocate.

The file name sorts last on purpose: `pytest -x` stops at the first failure, so a problem here cannot hide any other test.
It holds no state of its own and passes alone in a fresh process as well as after the rest of the suite."""
import numpy as np
import pytest
import torch

from tests.util import assert_close, cache_to_layout, make_paged_kv, randn_bf16, to_dev_bf16, to_np_f32

pytestmark = pytest.mark.gpu

# Same attention tolerance as tests/test_gpu_kernels.py; the two entry points run the same kernels on
# the same inputs, so they are in fact expected to agree bit for bit, which is asserted as well.
ATT_ATOL, ATT_RTOL = 1e-2, 1.6e-2

# (context_lens, H, Hkv, D, block_size, num_splits): both go through split-K + reduce, i.e. both use the scratch.
CALLS = [
    ([5, 300, 256, 257, 1024, 17, 0, 40], 16, 8, 128, 256, 3),
    ([2000, 33, 700], 8, 1, 128, 256, 4),
]


def _args(cuda, case):
    from nano_vllm_rs_b200._ffi import KvLayout
    ctx, H, Hkv, D, bs, splits = case
    rng = np.random.default_rng(len(ctx) * 31 + H + splits)
    kc, vc, tables, _, _ = make_paged_kv(rng, ctx, bs, Hkv, D)
    q = to_dev_bf16(randn_bf16(rng, (len(ctx), H, D)), cuda)
    lay = KvLayout.head_major(bs, Hkv, D)
    return dict(q=q, kc=cache_to_layout(kc, "head_major", cuda), vc=cache_to_layout(vc, "head_major", cuda),
                tables=torch.from_numpy(tables).to(cuda), ctx=torch.tensor(ctx, dtype=torch.int32, device=cuda), lay=lay,
                B=len(ctx), H=H, D=D, splits=splits, scale=float(1.0 / np.sqrt(D)))


def _err(lib):
    return lib.nvllm_last_error().decode(errors="replace")


def _call(lib, a, out, static, workspace=None):
    from nano_vllm_rs_b200.ops import BF16, _p, _stream
    common = (_p(out), _p(a["q"]), _p(a["kc"]), _p(a["vc"]), _p(a["tables"]), _p(a["ctx"]), a["B"], a["H"],
              a["tables"].shape[1], a["q"].stride(0), out.stride(0), a["lay"], a["scale"])
    if static:
        return lib.nvllm_paged_decode_attention_static(*common, a["splits"], BF16, _stream())
    return lib.nvllm_paged_decode_attention(*common, _p(workspace), workspace.numel(), a["splits"], BF16, _stream())


def test_static_scratch_decode_matches_workspace_decode_twice_in_a_row(cuda):
    from nano_vllm_rs_b200 import _ffi
    lib = _ffi.lib()
    assert lib.nvllm_paged_decode_static_scratch_bytes() == 8 << 20
    prepared = []
    for case in CALLS:  # every buffer either entry point touches exists before the calls under test
        a = _args(cuda, case)
        need = lib.nvllm_paged_decode_workspace_bytes(a["B"], a["H"], a["D"], a["splits"])
        assert 0 < need <= 8 << 20
        a["ws"] = torch.empty(need, dtype=torch.uint8, device=cuda)
        a["ref"] = torch.empty((a["B"], a["H"], a["D"]), dtype=torch.bfloat16, device=cuda)
        a["out"] = torch.full((a["B"], a["H"], a["D"]), float("nan"), dtype=torch.bfloat16, device=cuda)
        prepared.append(a)
    for a in prepared:
        assert _call(lib, a, a["ref"], static=False, workspace=a["ws"]) == 0, _err(lib)
    torch.cuda.synchronize()
    torch_bytes = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    for a in prepared:  # two calls back to back on one stream, no host sync between them
        assert _call(lib, a, a["out"], static=True) == 0, _err(lib)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_bytes
    assert free_after >= free_before, f"device memory shrank by {free_before - free_after} bytes across the static calls"
    for a, case in zip(prepared, CALLS):
        assert_close(to_np_f32(a["out"]), to_np_f32(a["ref"]), ATT_ATOL, ATT_RTOL, f"static-scratch decode {case}")
        assert torch.equal(a["out"], a["ref"]), f"static-scratch decode {case}: same kernels, must be bit-identical"


def test_static_scratch_decode_rejects_what_does_not_fit(cuda):
    from nano_vllm_rs_b200 import _ffi
    lib = _ffi.lib()
    a = _args(cuda, ([64, 64], 16, 8, 128, 256, 2))
    out = torch.zeros((a["B"], a["H"], a["D"]), dtype=torch.bfloat16, device=cuda)
    # the size check reads only the shape arguments: 4096 sequences x 16 heads x 8 splits x 130 floats = 272 MB > 8 MiB
    big = dict(a, B=4096, splits=8)
    assert lib.nvllm_paged_decode_workspace_bytes(4096, 16, 128, 8) > 8 << 20
    rc = _call(lib, big, out, static=True)
    assert rc == -4, rc  # NVLLM_ERR_WORKSPACE
    msg = _err(lib)
    assert "8388608" in msg and "workspace" in msg, msg
    torch.cuda.synchronize()
    assert float(out.float().abs().sum()) == 0.0  # nothing was launched
    # one split needs no scratch, whatever the batch: accepted (here on the small real batch)
    one = dict(a, splits=1)
    assert _call(lib, one, out, static=True) == 0, _err(lib)
    torch.cuda.synchronize()
