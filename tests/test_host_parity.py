"""CPU tests (no GPU): the C++ host layer behind include/nvllm_engine.h against the Python oracle, BIT-EXACT —
XXH64 digests, block ids / ref-counts / hashes, free-list order, block tables, slot mappings, context lengths,
cu_seqlens, prefill/decode order and preemption — every step of randomised and Appendix-B workloads."""
import json
import os

import numpy as np
import pytest

from nano_vllm_rs_b200 import engine as E
from oracle import host as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "xxh64.json")))


def test_compute_hash_golden_vectors_through_c_abi():
    for v in GOLD["compute_hash"]:
        prefix = int(v["prefix"], 16) if isinstance(v["prefix"], str) else v["prefix"]
        assert E.compute_hash(v["tokens"], prefix) == int(v["digest"], 16)
    rng = np.random.default_rng(0)
    for n in [1, 2, 3, 4, 5, 7, 16, 255, 256, 257, 1024]:
        t = rng.integers(-5, 151936, n)
        assert E.compute_hash(t) == H.compute_hash(t.tolist())
        assert E.compute_hash(t, 2 ** 64 - 3) == H.compute_hash(t.tolist(), 2 ** 64 - 3)


def _eq(a, b, what):
    if a is None or b is None:
        assert a is None and b is None, what
        return
    assert a.dtype == b.dtype, (what, a.dtype, b.dtype)
    assert np.array_equal(a, b), (what, a, b)


def lockstep(prompts, bs, num_blocks, max_tokens, max_batched, max_seqs, eos=None, seed=0, check_blocks=True):
    """Drive the C++ scheduler and the oracle with the same token stream; compare everything at every step."""
    ocfg = H.Config(kvcache_block_size=bs, num_kvcache_blocks=num_blocks, max_num_batched_tokens=max_batched,
                    max_num_seqs=max_seqs, eos_token_id=eos)
    osc = H.Scheduler(ocfg)
    ccfg = E.Config("tiny", kvcache_block_size=bs, num_kvcache_blocks=num_blocks, max_num_batched_tokens=max_batched,
                    max_num_seqs=max_seqs, eos_token_id=eos)
    csc = E.Scheduler(ccfg)
    oseqs = []
    for p, mt in zip(prompts, max_tokens):
        s = H.Sequence(p, H.SamplingParams(temperature=0.0, max_tokens=mt), block_size=bs)
        oseqs.append(s)
        osc.add_sequence(s)
        cid = csc.add_sequence(p, E.SamplingParams(temperature=0.0, max_tokens=mt))
        assert cid == len(oseqs) - 1
    id0 = oseqs[0].seq_id
    rng = np.random.default_rng(seed)
    steps = 0
    while not osc.is_finished():
        assert not csc.is_finished()
        batch, is_prefill = osc.schedule()
        o = H.prepare_prefill(batch, bs) if is_prefill else H.prepare_decode(batch, bs)
        c = csc.schedule()
        assert c.is_prefill == is_prefill, steps
        _eq(c.seq_ids, np.asarray([s.seq_id - id0 for s in batch], np.int64), "seq_ids")
        _eq(c.input_ids, o.input_ids, "input_ids"); _eq(c.positions, o.positions, "positions")
        _eq(c.slot_mapping, o.slot_mapping, "slot_mapping")
        _eq(c.cu_seqlens_q, o.cu_seqlens_q, "cu_q"); _eq(c.cu_seqlens_k, o.cu_seqlens_k, "cu_k")
        _eq(c.context_lens, o.context_lens, "context_lens"); _eq(c.block_tables, o.block_tables, "block_tables")
        if is_prefill:
            assert (c.max_seqlen_q, c.max_seqlen_k) == (o.max_seqlen_q, o.max_seqlen_k)
        if check_blocks:
            assert csc.free_list() == list(osc.block_manager.free_block_ids), steps
            for b in osc.block_manager.blocks:
                assert csc.block_info(b.block_id) == (b.ref_count, b.hash), (steps, b.block_id)
        toks = rng.integers(0, 50, len(batch)).tolist()   # small vocab: EOS and prefix collisions do happen
        osc.postprocess(batch, toks)
        csc.postprocess(toks)
        st = csc.get_stats()
        assert (st["waiting"], st["running"], st["finished"], st["preemptions"]) == (
            len(osc.waiting), len(osc.running), osc.num_finished, osc.num_preemptions), steps
        assert st["cached_blocks"] == len(osc.block_manager.hash_to_block_id)
        steps += 1
        assert steps < 100000
    assert csc.is_finished()
    for i, s in enumerate(oseqs):
        cs = csc.sequence(i)
        assert cs["token_ids"] == s.token_ids and cs["status"] == s.status and cs["num_prompt_tokens"] == s.num_prompt_tokens
    return steps


def test_reference_block_manager_cases_bs4():
    # block_manager.rs:426-506 scenarios, in lockstep
    lockstep([[1, 2, 3, 4, 5, 6, 7, 8, 9]], 4, 10, [3], 1000, 10)
    lockstep([[1, 2, 3, 4, 5, 6, 7, 8], [1, 2, 3, 4, 9, 10, 11, 12]], 4, 10, [2, 2], 1000, 10)
    lockstep([[1, 2, 3]], 4, 10, [6], 1000, 10)


@pytest.mark.parametrize("seed", range(6))
def test_random_workloads_with_shared_prefixes_eos_and_preemption(seed):
    rng = np.random.default_rng(seed)
    bs = [4, 16, 16, 32, 4, 16][seed]
    n = int(rng.integers(3, 20))
    shared = rng.integers(0, 50, int(rng.integers(0, 5)) * bs).tolist()
    prompts = []
    for _ in range(n):
        own = rng.integers(0, 50, int(rng.integers(1, 4 * bs))).tolist()
        prompts.append((shared + own) if rng.random() < 0.6 else own)
    need = max((len(p) + 40 + bs - 1) // bs for p in prompts)
    num_blocks = int(need + rng.integers(0, 2 * need))          # tight: forces preemption and recompute
    max_tokens = [int(rng.integers(1, 40)) for _ in prompts]
    steps = lockstep(prompts, bs, num_blocks, max_tokens, max_batched=int(rng.integers(8 * bs, 40 * bs)),
                     max_seqs=int(rng.integers(2, 12)), eos=7 if seed % 2 else None, seed=seed)
    assert steps > 0


def test_cfg1_trace_in_lockstep():
    rng = np.random.default_rng(1234)
    prompts = [rng.integers(0, 151936, 128).tolist() for _ in range(4)]
    assert lockstep(prompts, 256, 1000, [65] * 4, 32768, 512) == 65


def test_cfg2_trace_scaled_in_lockstep():
    rng = np.random.default_rng(1234)
    prompts = [rng.integers(0, 151936, 512).tolist() for _ in range(8)]
    assert lockstep(prompts, 256, 32, [512] * 8, 1024, 512, check_blocks=False) == 4 + 511


def test_cfg4_shared_prefix_scaled_in_lockstep():
    rng = np.random.default_rng(4321)
    prefix = rng.integers(0, 151936, 64).tolist()
    prompts = [prefix + rng.integers(0, 151936, 16).tolist() for _ in range(6)]
    lockstep(prompts, 16, 64, [3] * 6, 200, 512)


def test_config_defaults_and_validate():
    c = E.Config("qwen3-0.6b")
    assert (c.max_num_batched_tokens, c.max_num_seqs, c.max_model_len, c.kvcache_block_size) == (32768, 512, 4096, 256)
    assert (c.hidden_size, c.num_hidden_layers, c.num_attention_heads, c.num_key_value_heads, c.head_dim) == (1024, 28, 16, 8, 128)
    c.validate()
    r = E.Config("reference-default")      # qwen3.rs:70-89
    assert (r.hidden_size, r.intermediate_size, r.num_hidden_layers, r.num_key_value_heads, r.qk_norm) == (4096, 11008, 32, 32, 0)
    with pytest.raises(Exception):
        E.Config("qwen3-8b", tensor_parallel_size=9).validate()     # config.rs:99-101
    with pytest.raises(Exception):
        E.Config("qwen3-8b", tensor_parallel_size=3).validate()     # qwen3.rs:111-113
    E.Config("qwen3-32b", tensor_parallel_size=8, tp_rank=7, max_model_len=32768 + 64).validate()


def test_cfg2_full_size_trace_appendix_b():
    """BASELINE configs[1] at FULL size through the C++ scheduler: 256 x 512-token prompts,
    512 new tokens, block 256 — 4 prefill steps of 64 sequences, seq j owns blocks [2j, 2j+1], takes block 512+j at the
    first decode step and 768+j at length 769; 1024 blocks in total; slot mapping / context lengths follow."""
    rng = np.random.default_rng(1234)
    B, L, NEW, bs = 256, 512, 512, 256
    cfg = E.Config("qwen3-0.6b", kvcache_block_size=bs, num_kvcache_blocks=1024, max_num_batched_tokens=32768, max_num_seqs=256)
    sc = E.Scheduler(cfg)
    for _ in range(B):
        sc.add_sequence(rng.integers(0, 151936, L), E.SamplingParams(temperature=0.0, max_tokens=NEW, ignore_eos=True))
    for step in range(4):
        b = sc.schedule()
        assert b.is_prefill and len(b.seq_ids) == 64 and len(b.input_ids) == 32768 and b.block_tables is None
        j0 = 64 * step
        assert b.slot_mapping[0] == 2 * j0 * bs and b.slot_mapping[-1] == (2 * (j0 + 63) + 1) * bs + bs - 1
        assert np.array_equal(b.cu_seqlens_q, np.arange(0, 32768 + 1, L, dtype=np.int32))
        sc.postprocess(np.zeros(64, np.int64))
    st = sc.get_stats()
    assert st["waiting"] == 0 and st["running"] == B and st["used_blocks"] == 512 and st["cached_blocks"] == 512
    for k in range(NEW - 1):                       # the prefill already produced completion #1
        b = sc.schedule()
        n = L + 1 + k                              # sequence length at this decode forward
        assert not b.is_prefill and len(b.seq_ids) == B
        assert (b.context_lens == n).all() and (b.positions == n - 1).all()
        nblk = (n + bs - 1) // bs
        assert b.block_tables.shape == (B, nblk)
        j = np.arange(B)
        assert np.array_equal(b.block_tables[:, 0], 2 * j) and np.array_equal(b.block_tables[:, 1], 2 * j + 1)
        assert np.array_equal(b.block_tables[:, 2], 512 + j)
        if nblk == 4:
            assert np.array_equal(b.block_tables[:, 3], 768 + j)
        last = b.block_tables[:, nblk - 1].astype(np.int64)
        assert np.array_equal(b.slot_mapping.astype(np.int64), last * bs + (n - 1) % bs)
        sc.postprocess(np.full(B, 7, np.int64))
    st = sc.get_stats()
    assert sc.is_finished() and st["finished"] == B and st["preemptions"] == 0 and st["used_blocks"] == 0
    assert st["prefill_batches"] == 4 and st["decode_batches"] == NEW - 1


def test_cfg4_full_size_shared_prefix_trace_appendix_b():
    """BASELINE configs[3] at FULL size (Appendix B, cfg4): 256 sequences = one shared 4096-token prefix + 256 unique tokens;
    17 full blocks each; sequence 0 allocates blocks 0..16, sequence k >= 1 hits blocks 0..15 (ref_count -> k + 1,
    num_cached_tokens = 4096) and misses the 17th (chained hash differs) -> new block 16 + k; 7 sequences per prefill step
    (admission counts the full 4352 tokens), 37 prefill steps, 272 blocks after prefill, + 256 at the first decode step."""
    rng = np.random.default_rng(4321)
    bs, B = 256, 256
    prefix = rng.integers(0, 151936, 4096)
    cfg = E.Config("qwen3-8b", kvcache_block_size=bs, num_kvcache_blocks=600, max_num_batched_tokens=32768, max_num_seqs=256,
                   max_model_len=8192)
    sc = E.Scheduler(cfg)
    for _ in range(B):
        sc.add_sequence(np.concatenate([prefix, rng.integers(0, 151936, 256)]),
                        E.SamplingParams(temperature=0.0, max_tokens=2, ignore_eos=True))
    steps, seen = 0, 0
    while sc.get_stats()["waiting"] > 0:
        b = sc.schedule()
        assert b.is_prefill
        n = len(b.seq_ids)
        assert n == (7 if seen + 7 <= B else B - seen)
        for i in range(n):
            k = seen + i
            q = b.cu_seqlens_q[i + 1] - b.cu_seqlens_q[i]
            assert b.cu_seqlens_k[i + 1] - b.cu_seqlens_k[i] == 4352
            assert q == (4352 if k == 0 else 256)            # only the uncached suffix is computed
            if b.block_tables is not None:
                want = list(range(16)) + [16 + k]
                assert b.block_tables[i, :17].tolist() == want
        seen += n
        sc.postprocess(np.zeros(n, np.int64))
        steps += 1
    assert steps == 37
    st = sc.get_stats()
    assert st["used_blocks"] == 272 and st["cached_blocks"] == 272
    assert sc.block_info(0)[0] == B and sc.block_info(15)[0] == B and sc.block_info(16)[0] == 1 and sc.block_info(17)[0] == 1
    b = sc.schedule()                                        # first decode step: 4353 % 256 == 1 -> one new block each
    assert not b.is_prefill and (b.context_lens == 4353).all()
    assert np.array_equal(b.block_tables[:, 17], 272 + np.arange(B))
    assert sc.get_stats()["used_blocks"] == 272 + 256
    sc.postprocess(np.zeros(B, np.int64))
    assert sc.is_finished() and sc.get_stats()["used_blocks"] == 0


def test_max_model_len_is_enforced_at_admission():
    """nothing bounded prompt + max_tokens by max_model_len (the reference carries the field, config.rs:27,60, and
    never reads it): a sequence could outgrow the runner's block tables mid-run.  Now the prompt is refused (no state
    change) when it cannot fit, and max_tokens is clamped so the sequence ends with len == max_model_len + 1 at the latest —
    in lockstep with the oracle scheduler given the clamped max_tokens explicitly."""
    bs, mml = 16, 64
    cfg = E.Config("tiny", kvcache_block_size=bs, num_kvcache_blocks=32, max_model_len=mml, max_num_batched_tokens=256)
    sc = E.Scheduler(cfg)
    with pytest.raises(E._ffi.NvllmError) as ei:
        sc.add_sequence(list(range(mml + 1)), E.SamplingParams(temperature=0.0, max_tokens=4))
    assert ei.value.code == -1 and sc.get_stats()["total_sequences"] == 0 and sc.is_finished()
    with pytest.raises(E._ffi.NvllmError):
        E.Scheduler(E.Config("tiny", kvcache_block_size=bs, num_kvcache_blocks=32, max_model_len=4096,
                             max_num_batched_tokens=128)).add_sequence(list(range(129)), E.SamplingParams(max_tokens=1))
    # prompt of 60 tokens, max_tokens 64 -> clamped to 64 - 60 + 1 = 5 completions; the longest context ever scheduled is 64
    prompts = [list(range(60)), list(range(100, 164)), list(range(7))]
    asked = [64, 64, 200]
    clamped = [5, 1, 58]
    ids = [sc.add_sequence(p, E.SamplingParams(temperature=0.0, max_tokens=m, ignore_eos=True)) for p, m in zip(prompts, asked)]
    osc = H.Scheduler(H.Config(kvcache_block_size=bs, num_kvcache_blocks=32, max_num_batched_tokens=256))
    oseqs = [H.Sequence(p, H.SamplingParams(temperature=0.0, max_tokens=m, ignore_eos=True), block_size=bs)
             for p, m in zip(prompts, clamped)]
    for s in oseqs:
        osc.add_sequence(s)
    longest = 0
    while not osc.is_finished():
        batch, is_prefill = osc.schedule()
        o = H.prepare_prefill(batch, bs) if is_prefill else H.prepare_decode(batch, bs)
        c = sc.schedule()
        assert c.is_prefill == is_prefill
        _eq(c.slot_mapping, o.slot_mapping, "slot_mapping"); _eq(c.block_tables, o.block_tables, "block_tables")
        if not is_prefill:
            longest = max(longest, int(c.context_lens.max()))
            assert c.block_tables.shape[1] <= mml // bs
        toks = [3] * len(batch)
        osc.postprocess(batch, toks); sc.postprocess(toks)
    assert sc.is_finished() and longest == mml
    assert [len(sc.sequence(i)["token_ids"]) - len(p) for i, p in zip(ids, prompts)] == clamped


def test_release_sequence_recycles_the_slab_and_never_reuses_ids():
    sc = E.Scheduler(E.Config("tiny", kvcache_block_size=16, num_kvcache_blocks=32))
    a = sc.add_sequence([1, 2, 3], E.SamplingParams(temperature=0.0, max_tokens=1))
    b = sc.add_sequence([4, 5], E.SamplingParams(temperature=0.0, max_tokens=3))
    with pytest.raises(E._ffi.NvllmError):
        sc.release_sequence(a)                    # still waiting
    sc.schedule(); sc.postprocess([9, 9])          # prefill: a finishes (max_tokens 1)
    assert sc.sequence(a)["status"] == "Finished" and sc.sequence(b)["status"] == "Running"
    sc.release_sequence(a)
    with pytest.raises(KeyError):
        sc.sequence(a)
    with pytest.raises(E._ffi.NvllmError):
        sc.release_sequence(a)
    c = sc.add_sequence([6, 7, 8, 9], E.SamplingParams(temperature=0.0, max_tokens=2))   # takes a's slab slot, a NEW id
    assert c == 2 and sc.sequence(c)["token_ids"] == [6, 7, 8, 9] and sc.sequence(b)["token_ids"] == [4, 5, 9]
    while not sc.is_finished():
        bt = sc.schedule()
        sc.postprocess([1] * len(bt.seq_ids))
    assert sc.sequence(b)["token_ids"] == [4, 5, 9, 1, 1] and sc.sequence(c)["token_ids"] == [6, 7, 8, 9, 1, 1]
    assert sc.get_stats()["used_blocks"] == 0 and sc.get_stats()["finished"] == 3


def test_config_validate_mirrors_the_kernel_constraints():
    for bs in (48, 80, 192):          # % 16 == 0 but neither a divisor nor a multiple of the prefill kernel's 128-key tile
        with pytest.raises(Exception):
            E.Config("tiny", kvcache_block_size=bs).validate()
    for bs in (16, 32, 64, 128, 256, 512):
        E.Config("tiny", kvcache_block_size=bs).validate()
    with pytest.raises(Exception):
        E.Config("tiny", head_dim=64).validate()
