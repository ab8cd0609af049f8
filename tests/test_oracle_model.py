"""CPU test: the oracle MODEL (oracle/qwen3.py: paged KV cache, prefill then decode steps driven by the oracle scheduler,
bf16 rounding at tensor boundaries) against logits of an independent implementation of the named architecture — Hugging
Face transformers' Qwen3ForCausalLM in fp32 (fixture tests/golden/qwen3_hf.json, generator make_qwen3_hf_golden.py).
The reference pins none of this (no numeric golden vectors; it lacks q/k-norm and an explicit head_dim)."""
import importlib.util
import json
import os

import numpy as np

from oracle import host as H
from oracle import qwen3 as Q

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "qwen3_hf.json")))
GOLD_WIDE = json.load(open(os.path.join(HERE, "golden", "qwen3_hf_wide.json")))   # two layers at the real Qwen3-0.6B widths
# same bound as the GPU-vs-oracle tests (tests/test_gpu_engine.py): bf16 rounding noise growing with sqrt(depth)
LOGIT_SIGMA_PER_SQRT_LAYER, LOGIT_RTOL = 0.035, 1.6e-2


def _weights(which="tiny"):
    spec = importlib.util.spec_from_file_location("mk", os.path.join(HERE, "golden", "make_qwen3_hf_golden.py"))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    gold = GOLD if which == "tiny" else GOLD_WIDE
    assert mk.CFGS[which][0] == gold["config"] and mk.SEED == gold["seed"]
    return mk.make_weights(which)


import pytest  # noqa: E402


@pytest.mark.parametrize("which", ["tiny", "wide"])
def test_oracle_model_matches_hf_qwen3_prefill_and_decode(which):
    GOLD = globals()["GOLD"] if which == "tiny" else GOLD_WIDE
    # bf16 rounding noise of the oracle vs HF's fp32: rms 0.005 sigma (tiny) / 0.011 sigma (width 1024, K up to 3072); the
    # bound is on the MAX over all logits (4-5 rms), so the wide fixture states 0.05 * sqrt(layers) (observed max 0.053)
    LOGIT_SIGMA_PER_SQRT_LAYER = 0.035 if which == "tiny" else 0.05
    cfg = dict(GOLD["config"], qk_norm=True)
    bs = 16
    oracle = Q.Qwen3Oracle(cfg, _weights(which), num_blocks=8, block_size=bs)
    sched = H.Scheduler(H.Config(kvcache_block_size=bs, num_kvcache_blocks=8, max_num_batched_tokens=256, max_num_seqs=8))
    n_dec = len(GOLD["forced"][0])
    seqs = [H.Sequence(p, H.SamplingParams(temperature=0.0, max_tokens=n_dec + 1, ignore_eos=True), block_size=bs)
            for p in GOLD["prompts"]]
    for s in seqs:
        sched.add_sequence(s)
    worst = 0.0
    for step in range(n_dec + 1):
        batch, is_prefill = sched.schedule()
        assert is_prefill == (step == 0) and len(batch) == len(seqs)
        inp = H.prepare_prefill(batch, bs) if is_prefill else H.prepare_decode(batch, bs)
        logits = oracle.forward(inp)
        ref = np.asarray(GOLD["steps"][step], dtype=np.float64)
        sig = ref.std(axis=-1, keepdims=True)
        tol = LOGIT_SIGMA_PER_SQRT_LAYER * np.sqrt(cfg["num_hidden_layers"]) * sig + LOGIT_RTOL * np.abs(ref)
        err = np.abs(logits - ref)
        worst = max(worst, float((err / sig).max()))
        assert (err <= tol).all(), f"step {step}: oracle vs HF Qwen3 off by {(err / sig).max():.4f} sigma"
        for i in range(len(seqs)):   # greedy token equal wherever HF's own top-1 / top-2 margin exceeds the tolerance
            top2 = np.sort(ref[i])[-2:]
            if top2[1] - top2[0] > 2 * tol[i].max():
                assert int(np.argmax(logits[i])) == int(np.argmax(ref[i]))
        forced = [f[step] if step < n_dec else 0 for f in GOLD["forced"]]
        sched.postprocess(batch, forced)
    print("max |oracle - HF| =", worst, "sigma")
    assert worst < (0.05 if which == "tiny" else 0.075)   # observed ~0.01-0.02: pure bf16 rounding noise; a wiring error shows up as O(1) sigma


def test_wiring_errors_would_be_caught():
    """Sensitivity of the check above: dropping the q/k-norm (the reference's own wiring) moves the logits by
    far more than the tolerance — the fixture does discriminate between the two architectures."""
    cfg = dict(GOLD["config"], qk_norm=False)
    bs = 16
    oracle = Q.Qwen3Oracle(cfg, _weights(), num_blocks=8, block_size=bs)
    s = H.Sequence(GOLD["prompts"][1], H.SamplingParams(temperature=0.0, max_tokens=1), block_size=bs)
    s.block_table = [0]
    logits = oracle.forward(H.prepare_prefill([s], bs))
    ref = np.asarray(GOLD["steps"][0][1])
    assert np.abs(logits[0] - ref).max() / ref.std() > 0.3


def test_oracle_prefix_cached_prefill_matches_hf_full_recompute():
    """SURVEY.md §8 a5 (flash_attention_varlen_with_cache, attention.rs:211-222): the second prompt shares two full blocks
    with the first; the scheduler's prefix cache hits them, only the suffix is computed (bottom-right causal mask over the
    paged prefix), and the logits equal HF's full recompute of the whole prompt."""
    cfg = dict(GOLD["config"], qk_norm=True)
    bs = 16
    oracle = Q.Qwen3Oracle(cfg, _weights(), num_blocks=16, block_size=bs)
    sched = H.Scheduler(H.Config(kvcache_block_size=bs, num_kvcache_blocks=16, max_num_batched_tokens=256, max_num_seqs=8))
    pre, sfx = GOLD["prefix"]["prefix"], GOLD["prefix"]["suffixes"]
    for i in range(2):                      # one prompt per prefill step so that the second one finds the blocks cached
        s = H.Sequence(pre + sfx[i], H.SamplingParams(temperature=0.0, max_tokens=8, ignore_eos=True), block_size=bs)
        sched.add_sequence(s)
        batch, is_prefill = sched.schedule()
        assert is_prefill and len(batch) == 1
        inp = H.prepare_prefill(batch, bs)
        if i == 1:
            assert s.num_cached_tokens == 32 and inp.block_tables is not None and len(inp.input_ids) == len(sfx[1])
        logits = oracle.forward(inp)
        ref = np.asarray(GOLD["prefix"]["logits"][i])
        sig = ref.std()
        tol = LOGIT_SIGMA_PER_SQRT_LAYER * np.sqrt(cfg["num_hidden_layers"]) * sig + LOGIT_RTOL * np.abs(ref)
        assert (np.abs(logits[0] - ref) <= tol).all(), (i, np.abs(logits[0] - ref).max() / sig)
        sched.postprocess(batch, [1])
