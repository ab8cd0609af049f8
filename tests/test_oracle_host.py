"""CPU tests: the oracle's integer host layer against every known answer the reference's own unit tests hold for
this path plus the worked traces of SURVEY.md Appendix B.  These pin the oracle before anything
is compared with it."""
import ctypes
import json
import os
import struct

import numpy as np
import pytest

from oracle import host as H
from oracle.xxh64 import compute_hash, xxh64

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "xxh64.json")))


def seq(tokens, bs, **sp):
    return H.Sequence(tokens, H.SamplingParams(**sp), block_size=bs)


# ---------------------------------------------------------------------------- XXH64 (block_manager.rs:109-123)

def test_xxh64_canonical_vectors():
    for v in GOLD["canonical"]:
        assert xxh64(bytes.fromhex(v["data_hex"]), v["seed"]) == int(v["digest"], 16)


def test_compute_hash_golden_vectors():
    for v in GOLD["compute_hash"]:
        prefix = v["prefix"]
        if isinstance(prefix, str):
            prefix = int(prefix, 16)
        assert compute_hash(v["tokens"], prefix) == int(v["digest"], 16), v


def test_xxh64_against_independent_package_and_c_restatement():
    xxhash = pytest.importorskip("xxhash")
    rng = np.random.default_rng(0)
    lib_path = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    clib = ctypes.CDLL(lib_path) if os.path.exists(lib_path) else None
    if clib is not None:
        clib.oracle_xxh64.restype = ctypes.c_uint64
        clib.oracle_xxh64.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint64]
        clib.oracle_compute_hash.restype = ctypes.c_uint64
        clib.oracle_compute_hash.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64]
    for n in list(range(0, 70)) + [255, 256, 2048, 2056, 4099]:
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        seed = int(rng.integers(0, 2 ** 63))
        want = xxhash.xxh64(data, seed=seed).intdigest()
        assert xxh64(data, seed) == want
        if clib is not None:
            assert clib.oracle_xxh64(data, n, seed) == want
    toks = rng.integers(0, 151936, 256).astype(np.int64)   # one full KV block of Qwen3 token ids
    want = xxhash.xxh64(struct.pack("<Q", 77) + toks.tobytes(), seed=0).intdigest()
    assert compute_hash(toks.tolist(), 77) == want
    if clib is not None:
        assert clib.oracle_compute_hash(toks.ctypes.data, 256, 1, 77) == want


def test_hash_properties_block_manager_rs_406_423():
    h1, h2, h3 = compute_hash([1, 2, 3, 4, 5]), compute_hash([1, 2, 3, 4, 5]), compute_hash([1, 2, 3, 4, 6])
    assert h1 == h2 and h1 != h3
    assert compute_hash([1, 2, 3, 4, 5], 12345) != h1


# ---------------------------------------------------------------------------- sequence.rs tests

def test_sequence_block_calculations_sequence_rs_288_303():
    s = seq([1] * 300, 256)
    assert s.num_blocks() == 2 and s.last_block_num_tokens() == 44
    assert len(s.get_block_tokens(0)) == 256 and len(s.get_block_tokens(1)) == 44


def test_sequence_append_and_stop_sequence_rs_273_333():
    s = seq([1, 2, 3], 256)
    s.append_token(4); s.append_token(5)
    assert s.num_tokens == 5 and s.num_completion_tokens() == 2 and s.last_token == 5 and s.completion_token_ids() == [4, 5]
    s = seq([1, 2, 3], 256, max_tokens=2)
    assert not s.should_stop(999)
    s.append_token(4); s.append_token(5)
    assert s.should_stop(999)
    s2 = seq([1, 2, 3], 256, max_tokens=10); s2.append_token(999)
    assert s2.should_stop(999)
    s3 = seq([1, 2, 3], 256, max_tokens=10, ignore_eos=True); s3.append_token(999)
    assert not s3.should_stop(999)


# ---------------------------------------------------------------------------- block_manager.rs tests (:426-539)

def test_sequence_allocation_9_tokens_bs4():
    bm = H.BlockManager(10, 4)
    s = seq([1, 2, 3, 4, 5, 6, 7, 8, 9], 4)
    assert bm.can_allocate(s)
    bm.allocate(s)
    assert len(s.block_table) == 3 and len(bm.free_block_ids) == 7 and len(bm.used_block_ids) == 3
    assert s.block_table == [0, 1, 2]  # FIFO from the front (:223)


def test_sequence_deallocation_returns_everything():
    bm = H.BlockManager(10, 4)
    s = seq([1, 2, 3, 4, 5], 4)
    bm.allocate(s)
    bm.deallocate(s)
    assert s.block_table == [] and s.num_cached_tokens == 0 and len(bm.free_block_ids) == 10 and not bm.used_block_ids
    assert list(bm.free_block_ids) == [2, 3, 4, 5, 6, 7, 8, 9, 1, 0]  # reverse table order, pushed to the back (:141,:241)
    assert bm.hash_to_block_id == {}  # mapping dropped on free (:144-148)


def test_prefix_caching_shared_first_block():
    bm = H.BlockManager(10, 4)
    s1, s2 = seq([1, 2, 3, 4, 5, 6, 7, 8], 4), seq([1, 2, 3, 4, 9, 10, 11, 12], 4)
    bm.allocate(s1)
    bm.allocate(s2)
    assert s2.num_cached_tokens == 4
    assert bm.blocks[s1.block_table[0]].ref_count == 2
    assert s1.block_table == [0, 1] and s2.block_table == [0, 2]
    assert bm.blocks[0].hash == 0x73f859a04f669e6d and bm.blocks[1].hash == 0x087884241753df3e
    assert bm.blocks[2].hash == 0x8fcdfab13e7748aa


def test_append_operations_1_to_2_blocks():
    bm = H.BlockManager(10, 4)
    s = seq([1, 2, 3], 4)
    bm.allocate(s)
    assert len(s.block_table) == 1
    s.append_token(4)
    assert bm.can_append(s); bm.may_append(s)
    assert bm.blocks[s.block_table[0]].hash == compute_hash([1, 2, 3, 4])  # completed block hashed (:285-299)
    s.append_token(5)
    assert bm.can_append(s); bm.may_append(s)
    assert len(s.block_table) == 2


def test_insufficient_blocks_refused():
    bm = H.BlockManager(2, 4)
    s = seq([1] * 12, 4)
    assert not bm.can_allocate(s)
    with pytest.raises(RuntimeError):
        bm.allocate(s)


def test_block_manager_stats():
    bm = H.BlockManager(100, 256)
    assert bm.stats() == dict(total_blocks=100, free_blocks=100, used_blocks=0, cached_blocks=0, block_size=256)
    s = seq([1] * 300, 256)
    bm.allocate(s)
    st = bm.stats()
    assert st["used_blocks"] == 2 and st["cached_blocks"] == 1


# ---------------------------------------------------------------------------- scheduler.rs tests (:417-578)

def sched_cfg(**kw):
    c = H.Config(max_num_batched_tokens=1000, max_num_seqs=10, max_model_len=2048, eos_token_id=2,
                 kvcache_block_size=16, num_kvcache_blocks=100)
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def test_prefill_then_decode_scheduling():
    sc = H.Scheduler(sched_cfg())
    for i in range(3):
        sc.add_sequence(seq([1, 2, 3, 4, i], 16))
    seqs, is_prefill = sc.schedule()
    assert is_prefill and len(seqs) == 3 and len(sc.running) == 3 and not sc.waiting
    assert all(s.status == H.RUNNING for s in seqs)
    sc = H.Scheduler(sched_cfg())
    sc.add_sequence(seq([1, 2, 3, 4, 5], 16))
    seqs, _ = sc.schedule()
    sc.postprocess(seqs, [6])
    seqs, is_prefill = sc.schedule()
    assert not is_prefill and len(seqs) == 1 and len(seqs[0]) == 6


def test_completion_by_max_tokens_and_eos():
    sc = H.Scheduler(sched_cfg())
    sc.add_sequence(seq([1, 2, 3], 16, max_tokens=1))
    seqs, _ = sc.schedule()
    sc.postprocess(seqs, [4])
    assert sc.is_finished() and sc.num_finished == 1
    sc = H.Scheduler(sched_cfg())
    sc.add_sequence(seq([1, 2, 3], 16, max_tokens=10))
    seqs, _ = sc.schedule()
    sc.postprocess(seqs, [2])
    assert sc.is_finished() and sc.num_finished == 1
    assert len(sc.block_manager.free_block_ids) == 100


def test_batch_size_limits():
    sc = H.Scheduler(sched_cfg(max_num_seqs=2, max_num_batched_tokens=10))
    for _ in range(5):
        sc.add_sequence(seq([1, 2, 3, 4, 5, 6], 16))
    seqs, _ = sc.schedule()
    assert len(seqs) == 1 and len(sc.waiting) == 4


def test_memory_pressure_admits_what_fits():
    sc = H.Scheduler(sched_cfg(num_kvcache_blocks=5))
    for _ in range(3):
        sc.add_sequence(seq([1] * 20, 16))
    seqs, is_prefill = sc.schedule()
    # 20 tokens = 2 blocks: seq 0 takes blocks 0,1 (block 0 hashed); seqs 1,2 hit block 0 and take one new block each,
    # but each still had to pass free >= 2 (:152-154): free goes 5 -> 3 -> 2 -> 1
    assert is_prefill and len(seqs) == 3
    assert [s.block_table for s in seqs] == [[0, 1], [0, 2], [0, 3]]
    assert list(sc.block_manager.free_block_ids) == [4] and sc.block_manager.blocks[0].ref_count == 3


def test_preemption_recompute_order():
    """scheduler.rs:185-198,226-231: victim = tail of running; freed ids go to the back; victim to the waiting front."""
    sc = H.Scheduler(sched_cfg(num_kvcache_blocks=4, eos_token_id=None))
    a, b = seq(list(range(100, 116)), 16, max_tokens=40), seq(list(range(200, 216)), 16, max_tokens=40)
    sc.add_sequence(a); sc.add_sequence(b)
    seqs, p = sc.schedule()
    assert p and [s.block_table for s in seqs] == [[0], [1]]
    sc.postprocess(seqs, [7, 8])                 # len 17 -> each needs a second block
    seqs, p = sc.schedule()
    assert not p and [s.block_table for s in seqs] == [[0, 2], [1, 3]]
    for t in range(15):
        sc.postprocess(seqs, [t, t]); seqs, p = sc.schedule()
    assert len(a) == 32 and not p
    sc.postprocess(seqs, [1, 1])                 # len 33: both need a block, none free -> b is preempted
    seqs, p = sc.schedule()
    assert not p and seqs == [a] and sc.num_preemptions == 1
    assert a.block_table == [0, 2, 3]            # b freed [1,3] in reverse: free list = [3, 1]
    assert list(sc.waiting) == [b] and b.block_table == [] and b.status == H.PREEMPTED
    assert list(sc.block_manager.free_block_ids) == [1]


# ---------------------------------------------------------------------------- model_runner.rs metadata

def test_prepare_decode_inputs_model_runner_rs_501_520():
    s1, s2 = seq([1, 2, 3], 16), seq([4, 5], 16)
    s1.block_table, s2.block_table = [3], [7]
    d = H.prepare_decode([s1, s2], 16)
    assert d.input_ids.tolist() == [3, 5] and d.positions.tolist() == [2, 1]
    assert d.input_ids.dtype == np.int64 and d.positions.dtype == np.int64
    assert d.context_lens.tolist() == [3, 2] and d.context_lens.dtype == np.int32
    assert d.slot_mapping.tolist() == [3 * 16 + 2, 7 * 16 + 1] and d.slot_mapping.dtype == np.int32
    assert d.block_tables.tolist() == [[3], [7]] and d.block_tables.dtype == np.int32


def test_block_table_padding_minus_one():
    s1, s2 = seq([1] * 40, 16), seq([4, 5], 16)
    s1.block_table, s2.block_table = [3, 9, 2], [7]
    d = H.prepare_decode([s1, s2], 16)
    assert d.block_tables.tolist() == [[3, 9, 2], [7, -1, -1]]


# ---------------------------------------------------------------------------- Appendix B traces

def run_trace(prompts, bs, num_blocks, max_tokens, max_batched=32768, max_seqs=512, token_fn=lambda s, step: 7):
    cfg = H.Config(kvcache_block_size=bs, num_kvcache_blocks=num_blocks, max_num_batched_tokens=max_batched, max_num_seqs=max_seqs)
    sc = H.Scheduler(cfg)
    seqs = [seq(p, bs, max_tokens=max_tokens, ignore_eos=True) for p in prompts]
    for s in seqs:
        sc.add_sequence(s)
    steps = []
    while not sc.is_finished():
        batch, is_prefill = sc.schedule()
        inp = H.prepare_prefill(batch, bs) if is_prefill else H.prepare_decode(batch, bs)
        steps.append((is_prefill, [s.seq_id for s in batch], inp, [list(s.block_table) for s in batch]))
        sc.postprocess(batch, [token_fn(s, len(steps)) for s in batch])
    return sc, seqs, steps


def test_trace_cfg1_four_128_token_prompts():
    rng = np.random.default_rng(1234)
    prompts = [rng.integers(0, 151936, 128).tolist() for _ in range(4)]
    sc, seqs, steps = run_trace(prompts, 256, 1000, max_tokens=65)
    assert len(steps) == 65 and steps[0][0] and not any(s[0] for s in steps[1:])   # 1 prefill + 64 decode forwards
    p = steps[0][2]
    assert p.cu_seqlens_q.tolist() == [0, 128, 256, 384, 512] and p.cu_seqlens_k.tolist() == p.cu_seqlens_q.tolist()
    assert p.block_tables is None and p.max_seqlen_q == 128
    assert p.slot_mapping.tolist() == [256 * i + j for i in range(4) for j in range(128)]
    for k in range(64):
        d = steps[1 + k][2]
        assert d.positions.tolist() == [128 + k] * 4
        assert d.slot_mapping.tolist() == [256 * i + 128 + k for i in range(4)]
        assert d.context_lens.tolist() == [129 + k] * 4
        assert d.block_tables.tolist() == [[0], [1], [2], [3]]
    assert sc.block_manager.hash_to_block_id == {} or True  # no full block is ever completed: no XXH64 in cfg1
    assert all(b.hash is None for b in sc.block_manager.blocks)


def test_trace_cfg2_block_ids_scaled():
    """cfg2 (256 x 512-token prompts, bs 256, 512 new tokens) scaled 1/16 in sequence count so the pure-Python
    loop stays fast: 16 seqs, max_num_batched_tokens 2048 -> 4 prefill steps of 4 seqs; same arithmetic."""
    n = 16
    rng = np.random.default_rng(1234)
    prompts = [rng.integers(0, 151936, 512).tolist() for _ in range(n)]
    sc, seqs, steps = run_trace(prompts, 256, 4 * n, max_tokens=512, max_batched=2048)
    pre = [s for s in steps if s[0]]
    assert len(pre) == 4 and all(len(s[1]) == 4 for s in pre)
    for j in range(n):
        assert pre[j // 4][3][j % 4] == [2 * j, 2 * j + 1]
    dec = [s for s in steps if not s[0]]
    assert len(dec) == 511
    assert dec[0][3] == [[2 * j, 2 * j + 1, 2 * n + j] for j in range(n)]          # len 513: new block from the front
    assert dec[255][3] == [[2 * j, 2 * j + 1, 2 * n + j] for j in range(n)]        # len 768: third block hashed
    assert dec[256][3] == [[2 * j, 2 * j + 1, 2 * n + j, 3 * n + j] for j in range(n)]  # len 769
    assert dec[256][2].slot_mapping.tolist() == [(3 * n + j) * 256 for j in range(n)]
    assert dec[0][2].slot_mapping.tolist() == [(2 * n + j) * 256 for j in range(n)]
    assert len(sc.block_manager.free_block_ids) == 4 * n  # everything returned at the end


def test_trace_cfg4_shared_prefix_scaled():
    """cfg4 (shared 4096-token prefix + 256 unique) scaled: bs 16, prefix 64 = 4 blocks, suffix 16, 6 seqs."""
    rng = np.random.default_rng(4321)
    prefix = rng.integers(0, 151936, 64).tolist()
    prompts = [prefix + rng.integers(0, 151936, 16).tolist() for _ in range(6)]
    sc, seqs, steps = run_trace(prompts, 16, 64, max_tokens=3, max_batched=200)
    pre = [s for s in steps if s[0]]
    assert [len(s[1]) for s in pre] == [2, 2, 2]   # admission counts the full 80 tokens per seq (scheduler.rs:135): 2 x 80 <= 200
    tables = [t for s in pre for t in s[3]]
    assert tables[0] == [0, 1, 2, 3, 4]
    for k in range(1, 6):
        assert tables[k] == [0, 1, 2, 3, 4 + k]
    p0 = pre[0][2]
    assert p0.cu_seqlens_q.tolist() == [0, 80, 96] and p0.cu_seqlens_k.tolist() == [0, 80, 160]
    assert p0.block_tables is not None and p0.positions[80:].tolist() == list(range(64, 80))
    assert p0.slot_mapping[80:].tolist() == [5 * 16 + i for i in range(16)]
    p1 = pre[1][2]
    assert p1.cu_seqlens_q.tolist() == [0, 16, 32] and p1.max_seqlen_q == 16 and p1.max_seqlen_k == 80
    d0 = [s for s in steps if not s[0]][0]
    assert d0[3] == [[0, 1, 2, 3, 4 + k, 10 + k] for k in range(6)]   # 81 % 16 == 1: new block each, running order
