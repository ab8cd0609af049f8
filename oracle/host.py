"""Pure-Python restatement of the reference's integer host layer (TEST INFRASTRUCTURE — see oracle/__init__.py).

Everything here must be reproduced BIT-EXACTLY by the C++ host (csrc/host/): block ids, ref-counts, hashes,
block tables, slot mappings, context lengths, cu_seqlens, scheduling order.  Citations are file:line under
/root/reference/src/engine/.  Where the reference is internally inconsistent or does not compile, the INTENT
stated in SURVEY.md §8 (notes N1-N3, row a14) is followed and marked "intent:"."""
from collections import deque
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from .xxh64 import compute_hash

# ---------------------------------------------------------------------------- sampling_params.rs / config.rs


@dataclass
class SamplingParams:
    """sampling_params.rs:10-88 (plain data)."""
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False
    top_p: Optional[float] = None
    top_k: Optional[int] = None
    repetition_penalty: Optional[float] = None

    def is_greedy(self):
        return self.temperature == 0.0


@dataclass
class Config:
    """config.rs:16-71 defaults (only the fields the path reads)."""
    max_num_batched_tokens: int = 32768
    max_num_seqs: int = 512
    max_model_len: int = 4096
    tensor_parallel_size: int = 1
    enforce_eager: bool = False
    eos_token_id: Optional[int] = None
    kvcache_block_size: int = 256
    num_kvcache_blocks: Optional[int] = None  # None -> 1000 (scheduler.rs:72, model_runner.rs:84)


# ---------------------------------------------------------------------------- sequence.rs

WAITING, RUNNING, FINISHED, PREEMPTED, ERROR = "Waiting", "Running", "Finished", "Preempted", "Error"


class Sequence:
    """sequence.rs:50-237.  intent: block_size follows Config (the reference hard-codes 256, sequence.rs:99,
    which contradicts its own BlockManager tests with block size 4, block_manager.rs:426-438)."""
    _counter = 0  # sequence.rs:12 global AtomicU64

    def __init__(self, prompt_token_ids, sampling_params=None, block_size=256):
        self.seq_id = Sequence._counter  # sequence.rs:85
        Sequence._counter += 1
        self.status = WAITING
        self.token_ids = [int(t) for t in prompt_token_ids]
        self.last_token = self.token_ids[-1] if self.token_ids else 0  # sequence.rs:87
        self.num_tokens = len(self.token_ids)
        self.num_prompt_tokens = self.num_tokens
        self.num_cached_tokens = 0
        self.block_table: List[int] = []
        self.sampling_params = sampling_params or SamplingParams()
        self.block_size = block_size

    def __len__(self):  # sequence.rs:104
        return self.num_tokens

    def num_completion_tokens(self):  # sequence.rs:135
        return self.num_tokens - self.num_prompt_tokens

    def completion_token_ids(self):
        return self.token_ids[self.num_prompt_tokens:]

    def append_token(self, token_id):  # sequence.rs:150-154
        self.token_ids.append(int(token_id))
        self.last_token = int(token_id)
        self.num_tokens += 1

    def num_blocks(self):  # sequence.rs:157-159
        return (self.num_tokens + self.block_size - 1) // self.block_size

    def num_cached_blocks(self):  # sequence.rs:162-164
        return self.num_cached_tokens // self.block_size

    def last_block_num_tokens(self):  # sequence.rs:167-174
        rem = self.num_tokens % self.block_size
        return self.block_size if (rem == 0 and self.num_tokens > 0) else rem

    def get_block_tokens(self, block_idx):  # sequence.rs:177-186
        start = block_idx * self.block_size
        end = min((block_idx + 1) * self.block_size, self.num_tokens)
        return [] if start >= self.num_tokens else self.token_ids[start:end]

    def should_stop(self, eos_token_id):  # sequence.rs:189-205
        if self.num_completion_tokens() >= self.sampling_params.max_tokens:
            return True
        if not self.sampling_params.ignore_eos and eos_token_id is not None and self.last_token == eos_token_id:
            return True
        return False

    def is_finished(self):
        return self.status in (FINISHED, ERROR)


# ---------------------------------------------------------------------------- block_manager.rs


class Block:
    """block_manager.rs:12-65."""
    __slots__ = ("block_id", "ref_count", "hash", "token_ids")

    def __init__(self, block_id):
        self.block_id, self.ref_count, self.hash, self.token_ids = block_id, 0, None, []

    def update(self, h, token_ids):  # :38-41
        self.hash, self.token_ids = h, list(token_ids)

    def reset(self):  # :44-48
        self.ref_count, self.hash, self.token_ids = 1, None, []


class BlockManager:
    """block_manager.rs:69-331.  Quirks kept on purpose (SURVEY.md §8 note N3): FIFO free list taken from the
    front, freed ids appended at the back in reverse table order; the hash->block entry is dropped when a block
    is freed; a cache hit does not re-`update` the block; can_allocate demands free >= num_blocks(seq) even when
    most blocks will be hits."""

    def __init__(self, num_blocks, block_size):
        assert num_blocks > 0 and block_size > 0  # :92-93
        self.num_blocks, self.block_size = num_blocks, block_size
        self.blocks = [Block(i) for i in range(num_blocks)]
        self.hash_to_block_id = {}
        self.free_block_ids = deque(range(num_blocks))
        self.used_block_ids = set()

    compute_hash = staticmethod(compute_hash)  # :109-123

    def _allocate_block(self, block_id):  # :126-134
        blk = self.blocks[block_id]
        assert blk.ref_count == 0, f"Block {block_id} is not free"
        blk.reset()
        self.free_block_ids.remove(block_id)  # retain(|id| id != block_id)
        self.used_block_ids.add(block_id)
        return blk

    def _deallocate_block(self, block_id):  # :137-149
        blk = self.blocks[block_id]
        assert blk.ref_count == 0
        self.used_block_ids.discard(block_id)
        self.free_block_ids.append(block_id)
        if blk.hash is not None and self.hash_to_block_id.get(blk.hash) == block_id:
            del self.hash_to_block_id[blk.hash]

    def can_allocate(self, seq):  # :152-154
        return len(self.free_block_ids) >= seq.num_blocks()

    def _allocate_new_block(self, h, token_ids):  # :222-237
        if not self.free_block_ids:
            raise RuntimeError("No free blocks available")
        block_id = self.free_block_ids[0]
        blk = self._allocate_block(block_id)
        if h is not None:
            blk.update(h, token_ids)
            self.hash_to_block_id[h] = block_id
        else:
            blk.token_ids = list(token_ids)
        return block_id

    def allocate(self, seq):  # :157-219
        if seq.block_table:
            raise RuntimeError("Sequence already has allocated blocks")
        if not self.can_allocate(seq):
            raise RuntimeError("Not enough free blocks to allocate sequence")
        prefix_hash, cache_miss = None, False
        for block_idx in range(seq.num_blocks()):
            toks = seq.get_block_tokens(block_idx)
            cur = self.compute_hash(toks, prefix_hash) if len(toks) == self.block_size else None  # :173-177
            if cur is not None:
                existing = self.hash_to_block_id.get(cur)
                if existing is not None:
                    if not cache_miss and self.blocks[existing].token_ids == toks:  # :185
                        seq.num_cached_tokens += self.block_size  # :187
                        if existing in self.used_block_ids:
                            self.blocks[existing].ref_count += 1  # :191
                        else:
                            self._allocate_block(existing)  # :195
                        block_id = existing
                    else:
                        cache_miss = True
                        block_id = self._allocate_new_block(cur, toks)  # :200-201
                else:
                    cache_miss = True
                    block_id = self._allocate_new_block(cur, toks)  # :205-206
            else:
                cache_miss = True
                block_id = self._allocate_new_block(None, toks)  # :210-211
            seq.block_table.append(block_id)  # :214
            prefix_hash = cur  # :215

    def deallocate(self, seq):  # :240-252
        for block_id in reversed(seq.block_table):
            blk = self.blocks[block_id]
            assert blk.ref_count > 0
            blk.ref_count -= 1
            if blk.ref_count == 0:
                self._deallocate_block(block_id)
        seq.num_cached_tokens = 0
        seq.block_table = []

    def can_append(self, seq):  # :255-262
        if len(seq) % self.block_size == 1:
            return len(self.free_block_ids) > 0
        return True

    def may_append(self, seq):  # :265-304
        if not seq.block_table:
            raise RuntimeError("Sequence has no allocated blocks")
        last_idx = len(seq.block_table) - 1
        last_id = seq.block_table[last_idx]
        last = self.blocks[last_id]
        if len(seq) % self.block_size == 1:
            if last.hash is not None:  # :276
                if not self.free_block_ids:
                    raise RuntimeError("No free blocks for append")
                new_id = self.free_block_ids[0]  # :278
                self._allocate_block(new_id)
                seq.block_table.append(new_id)
        elif len(seq) % self.block_size == 0:
            if last.hash is None:  # :287
                toks = seq.get_block_tokens(seq.num_blocks() - 1)
                prefix = self.blocks[seq.block_table[last_idx - 1]].hash if len(seq.block_table) > 1 else None
                h = self.compute_hash(toks, prefix)
                last.update(h, toks)
                self.hash_to_block_id[h] = last_id

    def stats(self):  # :307-315
        return dict(total_blocks=self.num_blocks, free_blocks=len(self.free_block_ids),
                    used_blocks=len(self.used_block_ids), cached_blocks=len(self.hash_to_block_id),
                    block_size=self.block_size)


# ---------------------------------------------------------------------------- scheduler.rs


class Scheduler:
    """scheduler.rs:14-334.  The reference clones sequences between queues (:164,215,268) and re-inserts the
    updated copy by seq_id; with Python object identity that is the same state machine."""

    def __init__(self, config: Config):
        self.max_num_seqs = config.max_num_seqs
        self.max_num_batched_tokens = config.max_num_batched_tokens
        self.eos_token_id = config.eos_token_id
        self.block_manager = BlockManager(config.num_kvcache_blocks or 1000, config.kvcache_block_size)  # :71-74
        self.waiting = deque()
        self.running = deque()
        self.num_preemptions = 0
        self.num_finished = 0

    def is_finished(self):  # :88-90
        return not self.waiting and not self.running

    def add_sequence(self, seq):  # :93-98
        seq.status = WAITING
        self.waiting.append(seq)

    def schedule(self):  # :103-116 -> (sequences, is_prefill)
        seqs = self._try_schedule_prefill()
        if seqs:
            return seqs, True
        return self._try_schedule_decode(), False

    def _try_schedule_prefill(self):  # :119-168
        scheduled, num_seqs, num_batched = [], 0, 0
        while self.waiting:
            seq = self.waiting[0]
            if num_seqs >= self.max_num_seqs:  # :131
                break
            seq_tokens = len(seq) - seq.num_cached_tokens  # :135 (before allocate: cached tokens never discounted)
            if num_batched + seq_tokens > self.max_num_batched_tokens:  # :136 strict >
                break
            if not self.block_manager.can_allocate(seq):  # :141
                break
            self.waiting.popleft()
            self.block_manager.allocate(seq)  # :149
            num_seqs += 1
            num_batched += seq_tokens
            seq.status = RUNNING
            scheduled.append(seq)
        for seq in scheduled:  # :163-165
            self.running.append(seq)
        return scheduled

    def _try_schedule_decode(self):  # :171-223
        scheduled, num_seqs, resched = [], 0, []
        while self.running:
            seq = self.running.popleft()  # :178
            if num_seqs >= self.max_num_seqs:
                resched.append(seq)
                continue
            preempted_self = False
            while not self.block_manager.can_append(seq):  # :185
                if self.running:
                    self._preempt(self.running.pop())  # :186-188 victim = tail of running
                elif scheduled:
                    self._preempt(scheduled.pop())  # :189-192
                else:
                    self._preempt(seq)  # :193-196
                    preempted_self = True
                    break
            # :201 re-tests can_append on a moved value; intent: a self-preempted sequence is not scheduled
            if not preempted_self and self.block_manager.can_append(seq):
                num_seqs += 1
                self.block_manager.may_append(seq)  # :203
                scheduled.append(seq)
        for seq in reversed(resched):  # :209-211
            self.running.appendleft(seq)
        for seq in reversed(scheduled):  # :214-216
            self.running.appendleft(seq)
        if not scheduled:
            raise RuntimeError("No sequences could be scheduled for decode")  # :218-220
        return scheduled

    def _preempt(self, seq):  # :226-231
        seq.status = PREEMPTED
        self.block_manager.deallocate(seq)
        self.waiting.appendleft(seq)
        self.num_preemptions += 1

    def postprocess(self, sequences, token_ids):  # :234-257
        if len(sequences) != len(token_ids):
            raise RuntimeError("Mismatch between sequences and token_ids length")
        finished = []
        for seq, tok in zip(sequences, token_ids):
            seq.append_token(tok)
            if seq.should_stop(self.eos_token_id):
                seq.status = FINISHED
                self.block_manager.deallocate(seq)
                self.running = deque(s for s in self.running if s.seq_id != seq.seq_id)  # :260-262
                self.num_finished += 1
                finished.append(seq)
        return finished


# ---------------------------------------------------------------------------- model_runner.rs (metadata only)


@dataclass
class StepInputs:
    """What ModelRunner::prepare_inputs + create_context hand to the model (model_runner.rs:159-300) — the
    fields of utils/context.rs:11-35 plus ids / positions.  dtypes: ids, positions i64; the rest i32."""
    is_prefill: bool
    input_ids: np.ndarray
    positions: np.ndarray
    slot_mapping: np.ndarray
    cu_seqlens_q: Optional[np.ndarray] = None
    cu_seqlens_k: Optional[np.ndarray] = None
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    context_lens: Optional[np.ndarray] = None
    block_tables: Optional[np.ndarray] = None


def _padded_tables(sequences, width=None):
    """model_runner.rs:280-293: pad every block table with -1 to the longest one."""
    max_blocks = max([len(s.block_table) for s in sequences] + [1]) if width is None else width
    t = np.full((len(sequences), max_blocks), -1, dtype=np.int32)
    for i, s in enumerate(sequences):
        t[i, : len(s.block_table)] = s.block_table
    return t


def prepare_prefill(sequences, block_size) -> StepInputs:
    """prepare_prefill_inputs + create_prefill_context, model_runner.rs:172-193,222-263.
    intent: the reference's slot_mapping is a placeholder (:241-243) and it ignores
    num_cached_tokens (:177); intended: only the uncached suffix [start, len) is computed, positions are absolute,
    slot = table[p / bs] * bs + p % bs, cu_seqlens_k counts the whole sequence, and block tables are passed iff some
    sequence has a cached prefix (attention.rs:85-91 branch).  start = min(num_cached_tokens, len - 1): a fully
    cached prompt still recomputes its last token so that the step yields logits (idempotent rewrite of that slot)."""
    ids, pos, slots, cu_q, cu_k = [], [], [], [0], [0]
    max_q = max_k = 0
    any_cached = False
    for s in sequences:
        n = len(s)
        start = min(s.num_cached_tokens, n - 1)
        any_cached |= start > 0
        ids.extend(s.token_ids[start:n])
        pos.extend(range(start, n))
        for p in range(start, n):
            slots.append(s.block_table[p // block_size] * block_size + p % block_size)
        cu_q.append(cu_q[-1] + (n - start))
        cu_k.append(cu_k[-1] + n)
        max_q, max_k = max(max_q, n - start), max(max_k, n)
    return StepInputs(True, np.asarray(ids, np.int64), np.asarray(pos, np.int64), np.asarray(slots, np.int32),
                      np.asarray(cu_q, np.int32), np.asarray(cu_k, np.int32), max_q, max_k, None,
                      _padded_tables(sequences) if any_cached else None)


def prepare_decode(sequences, block_size) -> StepInputs:
    """prepare_decode_inputs + create_decode_context, model_runner.rs:196-210,266-300: last token, position
    len-1, context_lens = len, block tables padded with -1.  intent: slot = table.last() * bs +
    last_block_num_tokens - 1 (the reference's `0..batch` is a placeholder, :270)."""
    ids = [s.last_token for s in sequences]
    pos = [len(s) - 1 for s in sequences]
    slots = [s.block_table[-1] * block_size + s.last_block_num_tokens() - 1 for s in sequences]
    ctx = [len(s) for s in sequences]
    return StepInputs(False, np.asarray(ids, np.int64), np.asarray(pos, np.int64), np.asarray(slots, np.int32),
                      context_lens=np.asarray(ctx, np.int32), block_tables=_padded_tables(sequences))
