"""CPU tests of launch heuristics that are plain host arithmetic (no compute call): the split count of the paged decode
attention.  The expectations are the measured optima of profiles/r02_decode_split_sweep.jsonl (B200, 148 SMs, two resident
CTAs per SM): one split once the grid holds a wave; otherwise as many as fill the first wave, never below 256 tokens each."""
import pytest


@pytest.mark.parametrize("batch,heads,kv_heads,max_ctx,want", [
    (512, 4, 1, 2048, 1),      # Qwen3-8B TP=8 shard: 512 CTAs = 1.73 waves — one split measured 5.94 vs 5.16 TB/s for four
    (512, 8, 2, 2048, 1),      # TP=4 shard
    (512, 32, 8, 2048, 1),     # one GPU
    (256, 16, 8, 1024, 1),     # configs[1]
    (64, 8, 1, 33024, 4),      # Qwen3-32B TP=8, 64 x 32.5K: 4 x 64 = 256 CTAs fill the first wave (measured best)
    (1, 16, 8, 4096, 16),      # one sequence: bounded by 256 tokens per split
    (1, 16, 8, 100, 1),        # short context: nothing to split
    (8, 32, 8, 40960, 4),      # 64 base CTAs again
    (300, 16, 1, 8192, 1),     # g = 16 -> two q-head groups per kv head: 600 CTAs
])
def test_paged_decode_split_choice(batch, heads, kv_heads, max_ctx, want):
    from nano_vllm_rs_b200 import _ffi
    got = _ffi.lib().nvllm_paged_decode_choose_splits(batch, heads, kv_heads, max_ctx)
    assert got == want
    # the workspace the caller sizes with num_splits = 0 covers whatever the chooser may return
    need = _ffi.lib().nvllm_paged_decode_workspace_bytes(batch, heads, 128, got)
    cap = _ffi.lib().nvllm_paged_decode_workspace_bytes(batch, heads, 128, 0)
    assert need <= cap
