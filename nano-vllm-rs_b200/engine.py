"""Python mirror of the reference's public API over the C++ host runtime (include/nvllm_engine.h).

Names, fields and defaults follow the reference crate: Config (config.rs:16-186),
SamplingParams (sampling_params.rs:10-120), Scheduler (scheduler.rs), ModelRunner (model_runner.rs), LLMEngine /
LLMEngineBuilder (llm_engine.rs:39-483).  All state and logic live in libnvllm.so; this file only marshals."""
import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import _ffi
from ._ffi import check, f32, f64, i32, i64, u64, vp


class CConfig(C.Structure):
    """nvllm_config."""
    _fields_ = [("max_num_batched_tokens", i32), ("max_num_seqs", i32), ("max_model_len", i32),
                ("tensor_parallel_size", i32), ("enforce_eager", i32), ("kvcache_block_size", i32),
                ("num_kvcache_blocks", i32), ("tp_rank", i32), ("eos_token_id", i64),
                ("vocab_size", i32), ("hidden_size", i32), ("intermediate_size", i32), ("num_hidden_layers", i32),
                ("num_attention_heads", i32), ("num_key_value_heads", i32), ("head_dim", i32),
                ("max_position_embeddings", i32), ("tie_word_embeddings", i32), ("qk_norm", i32),
                ("kv_layout", i32), ("sampling_seed", i32), ("rms_norm_eps", f32), ("reserved_f", f32), ("rope_theta", f64)]


class CSamplingParams(C.Structure):
    """nvllm_sampling_params."""
    _fields_ = [("temperature", f32), ("max_tokens", i32), ("ignore_eos", i32), ("top_k", i32), ("top_p", f32),
                ("repetition_penalty", f32)]


class CBatch(C.Structure):
    """nvllm_batch."""
    _fields_ = [("is_prefill", i32), ("num_seqs", i32), ("num_tokens", i32), ("max_seqlen_q", i32),
                ("max_seqlen_k", i32), ("max_blocks", i32), ("has_block_tables", i32), ("reserved", i32),
                ("seq_ids", C.POINTER(i64)), ("input_ids", C.POINTER(i64)), ("positions", C.POINTER(i64)),
                ("slot_mapping", C.POINTER(i32)), ("cu_seqlens_q", C.POINTER(i32)), ("cu_seqlens_k", C.POINTER(i32)),
                ("context_lens", C.POINTER(i32)), ("block_tables", C.POINTER(i32)), ("temperatures", C.POINTER(f32)),
                ("top_k", C.POINTER(i32)), ("top_p", C.POINTER(f32))]


class CSchedStats(C.Structure):
    _fields_ = [(n, i64) for n in ("total_sequences", "waiting", "running", "finished", "preemptions",
                                   "prefill_batches", "decode_batches", "total_blocks", "free_blocks", "used_blocks",
                                   "cached_blocks")]


ENGINE_PROTOTYPES = {
    "nvllm_config_init": (i32, [C.POINTER(CConfig), C.c_char_p]),
    "nvllm_config_validate": (i32, [C.POINTER(CConfig)]),
    "nvllm_sampling_params_init": (i32, [C.POINTER(CSamplingParams)]),
    "nvllm_compute_hash": (u64, [vp, C.c_size_t, i32, u64]),
    "nvllm_scheduler_create": (i32, [C.POINTER(vp), C.POINTER(CConfig)]),
    "nvllm_scheduler_destroy": (i32, [vp]),
    "nvllm_scheduler_add_sequence": (i64, [vp, vp, i32, C.POINTER(CSamplingParams)]),
    "nvllm_scheduler_is_finished": (i32, [vp]),
    "nvllm_scheduler_schedule": (i32, [vp, C.POINTER(CBatch)]),
    "nvllm_scheduler_postprocess": (i32, [vp, vp, i32]),
    "nvllm_scheduler_stats": (i32, [vp, C.POINTER(CSchedStats)]),
    "nvllm_scheduler_block_info": (i32, [vp, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(u64)]),
    "nvllm_scheduler_free_list": (i32, [vp, vp, i32]),
    "nvllm_scheduler_sequence": (i32, [vp, i64, vp, i32, C.POINTER(i32), C.POINTER(i32)]),
    "nvllm_scheduler_release_sequence": (i32, [vp, i64]),
    "nvllm_runner_create": (i32, [C.POINTER(vp), C.POINTER(CConfig), vp]),
    "nvllm_runner_destroy": (i32, [vp]),
    "nvllm_runner_init_weights": (i32, [vp, u64, f32]),
    "nvllm_runner_load_weight": (i32, [vp, C.c_char_p, vp, i64, i64]),
    "nvllm_runner_execute": (i32, [vp, C.POINTER(CBatch), vp, vp]),
    "nvllm_runner_execute_async": (i32, [vp, C.POINTER(CBatch)]),
    "nvllm_runner_fetch_tokens": (i32, [vp, vp, i32]),
    "nvllm_runner_stream": (vp, [vp]),
    "nvllm_runner_profile": (i32, [vp, i32]),
    "nvllm_runner_profile_report": (i64, [vp, C.c_char_p, C.c_size_t]),
    "nvllm_runner_last_launches": (i64, [vp]),
    "nvllm_runner_kv_cache": (i32, [vp, i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(_ffi.KvLayout)]),
    "nvllm_runner_time_decode_attention": (i32, [vp, i32, C.POINTER(f32), C.POINTER(i64)]),
    "nvllm_runner_replay_decode": (i32, [vp, i32, C.POINTER(f32)]),
    "nvllm_engine_create": (i32, [C.POINTER(vp), C.POINTER(CConfig), vp]),
    "nvllm_engine_destroy": (i32, [vp]),
    "nvllm_engine_scheduler": (vp, [vp]),
    "nvllm_engine_runner": (vp, [vp]),
    "nvllm_engine_add_request": (i64, [vp, vp, i32, C.POINTER(CSamplingParams)]),
    "nvllm_engine_step": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "nvllm_engine_is_finished": (i32, [vp]),
}

_bound = False


def lib():
    global _bound
    l = _ffi.lib()
    if not _bound:
        for name, (res, args) in ENGINE_PROTOTYPES.items():
            fn = getattr(l, name, None)
            if fn is not None:
                fn.restype, fn.argtypes = res, args
        _bound = True
    return l


# ---------------------------------------------------------------------------- Config / SamplingParams


@dataclass
class SamplingParams:
    """sampling_params.rs:10-88."""
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False
    top_p: Optional[float] = None
    top_k: Optional[int] = None
    repetition_penalty: Optional[float] = None

    def is_greedy(self):
        return self.temperature == 0.0

    def to_c(self):
        return CSamplingParams(self.temperature, self.max_tokens, int(self.ignore_eos), self.top_k or 0,
                               self.top_p or 0.0, self.repetition_penalty or 0.0)


class Config:
    """config.rs:16-186 + Qwen3Config (qwen3.rs:26-125).  `arch` picks a named architecture's dimensions;
    any field can be overridden by keyword (same names as the reference's fields)."""

    def __init__(self, arch="qwen3-0.6b", **overrides):
        self.c = CConfig()
        check(lib().nvllm_config_init(C.byref(self.c), arch.encode()))
        self.arch = arch
        for k, v in overrides.items():
            self._set(k, v)

    def _set(self, k, v):
        if k == "eos_token_id":
            v = -1 if v is None else v
        if k == "num_kvcache_blocks":
            v = 0 if v is None else v
        if not any(k == f[0] for f in CConfig._fields_):
            raise AttributeError(f"Config has no field {k}")
        setattr(self.c, k, int(v) if not isinstance(v, float) else v)

    def __getattr__(self, k):
        if k in ("c", "arch"):
            raise AttributeError(k)
        return getattr(self.c, k)

    # builder methods of config.rs:147-185
    def with_max_num_batched_tokens(self, v): self._set("max_num_batched_tokens", v); return self
    def with_max_num_seqs(self, v): self._set("max_num_seqs", v); return self
    def with_max_model_len(self, v): self._set("max_model_len", v); return self
    def with_tensor_parallel_size(self, v): self._set("tensor_parallel_size", v); return self
    def with_enforce_eager(self, v): self._set("enforce_eager", v); return self

    def validate(self):
        check(lib().nvllm_config_validate(C.byref(self.c)))
        return self


def compute_hash(token_ids, prefix_hash=None):
    """BlockManager::compute_hash (block_manager.rs:109-123) through the C ABI."""
    a = np.ascontiguousarray(token_ids, dtype=np.int64)
    return lib().nvllm_compute_hash(a.ctypes.data, a.size, 0 if prefix_hash is None else 1, prefix_hash or 0)


# ---------------------------------------------------------------------------- Scheduler


class Batch:
    """Host view of one scheduled step (numpy copies of an nvllm_batch) + the raw struct for the runner."""

    def __init__(self, cb: CBatch):
        self.c = cb
        n, t = cb.num_seqs, cb.num_tokens
        arr = lambda p, k, dt: np.ctypeslib.as_array(p, (k,)).astype(dt, copy=True) if (k and p) else np.zeros(0, dt)
        self.is_prefill = bool(cb.is_prefill)
        self.seq_ids = arr(cb.seq_ids, n, np.int64)
        self.input_ids = arr(cb.input_ids, t, np.int64)
        self.positions = arr(cb.positions, t, np.int64)
        self.slot_mapping = arr(cb.slot_mapping, t, np.int32)
        self.max_seqlen_q, self.max_seqlen_k = cb.max_seqlen_q, cb.max_seqlen_k
        self.cu_seqlens_q = arr(cb.cu_seqlens_q, n + 1, np.int32) if cb.is_prefill else None
        self.cu_seqlens_k = arr(cb.cu_seqlens_k, n + 1, np.int32) if cb.is_prefill else None
        self.context_lens = None if cb.is_prefill else arr(cb.context_lens, n, np.int32)
        self.block_tables = (arr(cb.block_tables, n * cb.max_blocks, np.int32).reshape(n, cb.max_blocks)
                             if cb.has_block_tables else None)
        self.temperatures = arr(cb.temperatures, n, np.float32)
        self.top_k = arr(cb.top_k, n, np.int32)
        self.top_p = arr(cb.top_p, n, np.float32)

    def seq_lens(self):
        """Tokens of every scheduled sequence (prompt + completion so far): the `n` of the sampler's row seed."""
        return np.diff(self.cu_seqlens_k).astype(np.int64) if self.is_prefill else self.context_lens.astype(np.int64)


class Scheduler:
    """Scheduler (scheduler.rs) + BlockManager (block_manager.rs) + Sequences, in C++ (csrc/host/)."""

    def __init__(self, config: Config, _handle=None):
        self._owned = _handle is None
        if _handle is None:
            h = vp()
            check(lib().nvllm_scheduler_create(C.byref(h), C.byref(config.c)))
            _handle = h
        self.h = _handle

    def __del__(self):
        if getattr(self, "_owned", False) and self.h:
            lib().nvllm_scheduler_destroy(self.h)
            self.h = None

    def add_sequence(self, prompt_token_ids, sampling_params: SamplingParams = None) -> int:
        a = np.ascontiguousarray(prompt_token_ids, dtype=np.int64)
        sp = (sampling_params or SamplingParams()).to_c()
        r = lib().nvllm_scheduler_add_sequence(self.h, a.ctypes.data, a.size, C.byref(sp))
        if r < 0:
            check(int(r))
        return r

    def is_finished(self):
        return bool(lib().nvllm_scheduler_is_finished(self.h))

    def schedule(self) -> Batch:
        cb = CBatch()
        check(lib().nvllm_scheduler_schedule(self.h, C.byref(cb)))
        return Batch(cb)

    def postprocess(self, token_ids):
        a = np.ascontiguousarray(token_ids, dtype=np.int64)
        check(lib().nvllm_scheduler_postprocess(self.h, a.ctypes.data, a.size))

    def get_stats(self):
        st = CSchedStats()
        check(lib().nvllm_scheduler_stats(self.h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in CSchedStats._fields_}

    def block_info(self, block_id):
        rc, hh, hv = i32(), i32(), u64()
        check(lib().nvllm_scheduler_block_info(self.h, block_id, C.byref(rc), C.byref(hh), C.byref(hv)))
        return rc.value, (hv.value if hh.value else None)

    def free_list(self):
        n = lib().nvllm_scheduler_free_list(self.h, None, 0)
        a = np.zeros(max(n, 1), np.int32)
        lib().nvllm_scheduler_free_list(self.h, a.ctypes.data, n)
        return a[:n].tolist()

    def release_sequence(self, seq_id):
        """Frees a FINISHED sequence's storage (its outputs have been handed to the caller); sequence(seq_id) raises
        KeyError afterwards."""
        check(lib().nvllm_scheduler_release_sequence(self.h, seq_id))

    def sequence(self, seq_id):
        npt, st = i32(), i32()
        n = lib().nvllm_scheduler_sequence(self.h, seq_id, None, 0, C.byref(npt), C.byref(st))
        if n < 0:
            raise KeyError(seq_id)
        a = np.zeros(max(n, 1), np.int64)
        lib().nvllm_scheduler_sequence(self.h, seq_id, a.ctypes.data, n, None, None)
        return dict(token_ids=a[:n].tolist(), num_prompt_tokens=npt.value,
                    status=["Waiting", "Running", "Finished", "Preempted"][st.value])


# ---------------------------------------------------------------------------- ModelRunner / LLMEngine


class ModelRunner:
    """ModelRunner (model_runner.rs:19-448) + Qwen3Model (qwen3.rs): C++ host runtime over the sm_100a kernels."""

    def __init__(self, config: Config, comm=None, _handle=None):
        self.config = config
        self._owned = _handle is None
        if _handle is None:
            h = vp()
            check(lib().nvllm_runner_create(C.byref(h), C.byref(config.c), comm))
            _handle = h
        self.h = _handle
        self.vshard = config.vocab_size // config.tensor_parallel_size

    def __del__(self):
        if getattr(self, "_owned", False) and self.h:
            lib().nvllm_runner_destroy(self.h)
            self.h = None

    def init_weights(self, seed=1234, std=0.02):
        check(lib().nvllm_runner_init_weights(self.h, seed, std))

    def load_weight(self, name, bf16_bits):
        """bf16_bits: uint16 numpy array [rows, cols] (or [n] for norm weights) holding bf16 bit patterns."""
        a = np.ascontiguousarray(bf16_bits, dtype=np.uint16)
        rows, cols = (1, a.shape[0]) if a.ndim == 1 else a.shape
        check(lib().nvllm_runner_load_weight(self.h, name.encode(), a.ctypes.data, rows, cols))

    def execute_model(self, batch: Batch, return_logits=False):
        """execute_model + sample_tokens (model_runner.rs:105-156): returns the sampled token ids [num_seqs] (and logits)."""
        n = batch.c.num_seqs
        toks = np.zeros(n, np.int64)
        logits = np.zeros((n, self.vshard), np.float32) if return_logits else None
        check(lib().nvllm_runner_execute(self.h, C.byref(batch.c), toks.ctypes.data,
                                         logits.ctypes.data if return_logits else None))
        return (toks, logits) if return_logits else toks

    def profile(self, enable=True):
        """Start (clearing the accumulators) / stop the per-launch CUDA-event profile of eager steps."""
        check(lib().nvllm_runner_profile(self.h, int(enable)))

    def profile_report(self):
        """{kernel name: (launches, total_us)} accumulated since profile(True)."""
        n = lib().nvllm_runner_profile_report(self.h, None, 0)
        buf = C.create_string_buffer(int(n) + 1)
        lib().nvllm_runner_profile_report(self.h, buf, n + 1)
        out = {}
        for line in buf.value.decode().splitlines():
            name, calls, us = line.split("\t")
            out[name] = (int(calls), float(us))
        return out

    def last_launches(self):
        return lib().nvllm_runner_last_launches(self.h)

    def replay_decode(self, iters):
        """Device-only time (ms, CUDA events) of `iters` replays of the last decode step, inputs resident in HBM."""
        ms = f32()
        check(lib().nvllm_runner_replay_decode(self.h, iters, C.byref(ms)))
        return ms.value

    def time_decode_attention(self, iters=20):
        ms, nbytes = f32(), i64()
        check(lib().nvllm_runner_time_decode_attention(self.h, iters, C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value


@dataclass
class SequenceOutput:
    """SequenceOutput (sequence.rs:30-46)."""
    seq_id: int
    text: str
    token_ids: List[int]
    completion_token_ids: List[int]
    num_prompt_tokens: int
    num_completion_tokens: int
    status: str


class LLMEngine:
    """LLMEngine (llm_engine.rs:23-362), synchronous: add_request / step / generate / generate_stream / get_stats /
    health_check / shutdown over token ids.  `tokenizer` (optional): any object with encode(str) -> ids (list or an object
    with .ids, as the HF `tokenizers` crate the reference binds, llm_engine.rs:200-230) and decode(ids) -> str; with it,
    prompts may be strings and SequenceOutput.text is filled.  The reference's tokio plumbing (Arc<Mutex<..>>, mpsc
    channel, spawn) has no counterpart: one caller thread drives the engine, streaming is a Python generator."""

    def __init__(self, config: Config, comm=None, seed=1234, std=0.02, init_weights=True, tokenizer=None):
        config.validate()
        self.config = config
        self.tokenizer = tokenizer
        self.is_running = True
        self._live = {}     # seq_id -> tokens already reported (for step outputs / streaming)
        h = vp()
        check(lib().nvllm_engine_create(C.byref(h), C.byref(config.c), comm))
        self.h = h
        self.scheduler = Scheduler(config, _handle=vp(lib().nvllm_engine_scheduler(h)))
        self.model_runner = ModelRunner(config, _handle=vp(lib().nvllm_engine_runner(h)))
        if init_weights:
            self.model_runner.init_weights(seed, std)

    def __del__(self):
        if getattr(self, "h", None):
            lib().nvllm_engine_destroy(self.h)
            self.h = None

    # ---- tokenisation (llm_engine.rs:200-230: create_sequences) ----
    def _encode(self, prompt):
        if isinstance(prompt, str):
            if self.tokenizer is None:
                raise ValueError("string prompt given but the engine has no tokenizer")
            enc = self.tokenizer.encode(prompt)
            return list(getattr(enc, "ids", enc))
        return list(prompt)

    def _decode(self, ids):
        return self.tokenizer.decode(list(ids)) if self.tokenizer is not None else ""

    def add_request(self, prompt, sampling_params: SamplingParams = None) -> int:
        if not self.is_running:
            raise RuntimeError("engine has been shut down")
        sid = self.scheduler.add_sequence(self._encode(prompt), sampling_params)
        self._live[sid] = None
        return sid

    def step(self):
        """LLMEngine::step (llm_engine.rs:155-197). Returns (is_prefill, num_seqs, num_tokens)."""
        p, s, t = i32(), i32(), i32()
        check(lib().nvllm_engine_step(self.h, C.byref(p), C.byref(s), C.byref(t)))
        return bool(p.value), s.value, t.value

    def is_finished(self):
        return bool(lib().nvllm_engine_is_finished(self.h))

    def _output(self, sid, s):
        n = s["num_prompt_tokens"]
        comp = s["token_ids"][n:]
        return SequenceOutput(sid, self._decode(comp), s["token_ids"], comp, n, len(comp), s["status"])

    def step_outputs(self) -> List[SequenceOutput]:
        """One step with the reference's return value (llm_engine.rs:155-197): outputs of the sequences that FINISHED in it."""
        self.step()
        done = []
        for sid in list(self._live):
            s = self.scheduler.sequence(sid)
            if s["status"] == "Finished":
                done.append(self._output(sid, s))
                del self._live[sid]
                self.scheduler.release_sequence(sid)   # handed to the caller: the engine forgets it (llm_engine.rs:186-193)
        return done

    def generate(self, prompts, sampling_params: SamplingParams = None):
        """LLMEngine::generate (llm_engine.rs:71-98): completion token ids per prompt, in prompt order."""
        ids = [self.add_request(p, sampling_params) for p in prompts]
        while not self.is_finished():
            self.step()
        out = []
        for i in ids:
            s = self.scheduler.sequence(i)
            self._live.pop(i, None)
            out.append(s["token_ids"][s["num_prompt_tokens"]:])
            self.scheduler.release_sequence(i)
        return out

    def generate_outputs(self, prompts, sampling_params: SamplingParams = None) -> List[SequenceOutput]:
        """generate with the reference's return type (Vec<SequenceOutput>, in prompt order)."""
        ids = [self.add_request(p, sampling_params) for p in prompts]
        while not self.is_finished():
            self.step()
        outs = []
        for i in ids:
            self._live.pop(i, None)
            outs.append(self._output(i, self.scheduler.sequence(i)))
            self.scheduler.release_sequence(i)
        return outs

    def generate_stream(self, prompts, sampling_params: SamplingParams = None):
        """LLMEngine::generate_stream (llm_engine.rs:101-128) as a generator: after every step yields one SequenceOutput per
        sequence that produced a token in it (cumulative completion so far; status "Finished" on the last one)."""
        ids = [self.add_request(p, sampling_params) for p in prompts]
        seen = {i: 0 for i in ids}
        while seen:
            self.step()
            for i in list(seen):
                s = self.scheduler.sequence(i)
                n = len(s["token_ids"]) - s["num_prompt_tokens"]
                if n > seen[i] or s["status"] == "Finished":
                    yield self._output(i, s)
                seen[i] = n
                if s["status"] == "Finished":
                    del seen[i]
                    self._live.pop(i, None)
                    self.scheduler.release_sequence(i)

    def get_stats(self):
        """EngineStats (llm_engine.rs:312-327,375-400): scheduler counters + block-memory utilisation (percent)."""
        st = self.scheduler.get_stats()
        total = max(1, st["total_blocks"])
        return {"scheduler": st,
                "memory": {"total_blocks": st["total_blocks"], "free_blocks": st["free_blocks"], "used_blocks": st["used_blocks"],
                           "utilization": 100.0 * st["used_blocks"] / total},
                "is_running": self.is_running}

    def health_check(self):
        """HealthStatus (llm_engine.rs:330-342,404-416): healthy while block utilisation is below 95 %."""
        st = self.get_stats()
        return {"is_healthy": st["memory"]["utilization"] < 95.0, "memory_pressure": st["memory"]["utilization"],
                "active_sequences": st["scheduler"]["running"], "waiting_sequences": st["scheduler"]["waiting"]}

    def shutdown(self):
        """LLMEngine::shutdown (llm_engine.rs:345-357): stop accepting requests (queued work is dropped with the engine)."""
        self.is_running = False


class LLMEngineBuilder:
    """LLMEngineBuilder (llm_engine.rs:419-483)."""

    def __init__(self, arch="qwen3-0.6b"):
        self._cfg = Config(arch)

    def max_num_seqs(self, v): self._cfg._set("max_num_seqs", v); return self
    def max_num_batched_tokens(self, v): self._cfg._set("max_num_batched_tokens", v); return self
    def tensor_parallel_size(self, v): self._cfg._set("tensor_parallel_size", v); return self
    def enforce_eager(self, v): self._cfg._set("enforce_eager", v); return self
    def num_kvcache_blocks(self, v): self._cfg._set("num_kvcache_blocks", v); return self
    def build(self, **kw): return LLMEngine(self._cfg, **kw)
