"""Regenerates tests/golden/qwen3_hf.json: logits of an INDEPENDENT implementation of the named architecture — Hugging Face
`transformers` Qwen3ForCausalLM (fp32, CPU, eager attention) — on a tiny random Qwen3 (2 layers, GQA 4/2 heads, head_dim 128,
per-head q/k RMSNorm, half-split RoPE theta 1e6, untied head) whose weights (bf16-representable, norm weights != 1) are
stored in the fixture's seed.  tests/test_oracle_model.py checks oracle/qwen3.py (paged KV cache, prefill + decode steps,
bf16 rounding at the tensor boundaries) against it: this pins the model WIRING of SURVEY.md §8 note N1 (q/k-norm placement,
RoPE convention, GQA mapping, residual / norm order) that the reference itself leaves unpinned.  Run from the repo root."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle.bf16 import round_bf16  # noqa: E402

CFG = dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
           num_key_value_heads=2, head_dim=128, max_position_embeddings=4096, rms_norm_eps=1e-6, rope_theta=1e6,
           tie_word_embeddings=False)
# second fixture (qwen3_hf_wide.json): two layers at the REAL Qwen3-0.6B widths (hidden 1024, 16/8 heads of 128, inter 3072),
# vocabulary cut to 2048 to keep the fixture small — arch "qwen3-0.6b-2l" of nvllm_config_init / oracle.qwen3.ARCHS
CFG_WIDE = dict(CFG, vocab_size=2048, hidden_size=1024, intermediate_size=3072, num_attention_heads=16, num_key_value_heads=8)
CFGS = {"tiny": (CFG, "qwen3_hf.json"), "wide": (CFG_WIDE, "qwen3_hf_wide.json")}
SEED, STD = 20261104, 0.05
PROMPTS = [[5, 17, 900, 3, 44], [1023, 0, 7, 7, 512, 64, 33, 2, 81, 300, 11, 640]]
FORCED = [[10, 20, 30], [700, 5, 999]]   # tokens appended at the three decode steps (teacher forcing)
# prefix caching: two prompts sharing a 32-token prefix (= 2 full blocks of 16), different suffixes
PREFIX = [(i * 37 + 11) % 1024 for i in range(32)]
SUFFIXES = [[3, 1, 4, 1, 5], [9, 2, 6, 5, 3, 5, 8, 9, 7]]


def make_weights(which="tiny"):
    """name -> f32 array of bf16-representable values; norm weights 1 + 0.1 N(0,1)."""
    CFG = CFGS[which][0]
    rng = np.random.default_rng(SEED)
    H, Hkv, D, hid, I, V = (CFG[k] for k in ("num_attention_heads", "num_key_value_heads", "head_dim", "hidden_size",
                                             "intermediate_size", "vocab_size"))
    t = lambda *s: round_bf16((rng.standard_normal(s) * STD).astype(np.float32))
    n = lambda k: round_bf16((1 + 0.1 * rng.standard_normal(k)).astype(np.float32))
    w = {"model.embed_tokens.weight": t(V, hid), "model.norm.weight": n(hid), "lm_head.weight": t(V, hid)}
    for l in range(CFG["num_hidden_layers"]):
        p = f"model.layers.{l}."
        w[p + "input_layernorm.weight"] = n(hid); w[p + "post_attention_layernorm.weight"] = n(hid)
        w[p + "self_attn.q_norm.weight"] = n(D); w[p + "self_attn.k_norm.weight"] = n(D)
        w[p + "self_attn.q_proj.weight"] = t(H * D, hid); w[p + "self_attn.k_proj.weight"] = t(Hkv * D, hid)
        w[p + "self_attn.v_proj.weight"] = t(Hkv * D, hid); w[p + "self_attn.o_proj.weight"] = t(hid, H * D)
        w[p + "mlp.gate_proj.weight"] = t(I, hid); w[p + "mlp.up_proj.weight"] = t(I, hid); w[p + "mlp.down_proj.weight"] = t(hid, I)
    return w


def main(which="tiny"):
    import torch
    from transformers import Qwen3Config, Qwen3ForCausalLM
    CFG, fname = CFGS[which]
    cfg = Qwen3Config(attention_bias=False, attn_implementation="eager", **CFG)
    model = Qwen3ForCausalLM(cfg).eval().float()
    w = make_weights(which)
    sd = model.state_dict()
    assert set(sd) == set(w), set(sd) ^ set(w)
    model.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()})
    out = {"config": CFG, "seed": SEED, "std": STD, "prompts": PROMPTS, "forced": FORCED, "steps": []}
    seqs = [list(p) for p in PROMPTS]
    with torch.no_grad():
        for step in range(1 + len(FORCED[0])):        # step 0 = prefill, then 3 decode steps
            rows = [model(torch.tensor([s])).logits[0, -1].double().numpy() for s in seqs]   # full recompute = no cache logic
            out["steps"].append([[float(x) for x in r] for r in rows])
            if step < len(FORCED[0]):
                for s, f in zip(seqs, FORCED):
                    s.append(f[step])
    with torch.no_grad():
        out["prefix"] = {"prefix": PREFIX, "suffixes": SUFFIXES,
                         "logits": [[float(x) for x in model(torch.tensor([PREFIX + sfx])).logits[0, -1].double().numpy()]
                                    for sfx in SUFFIXES]}
    if which != "tiny":   # the wide fixture stores 6 significant digits (tolerances are ~1e-2 of a logit's sigma)
        rnd = lambda rows: [[float(f"{x:.6g}") for x in r] for r in rows]
        out["steps"] = [rnd(st) for st in out["steps"]]
        out["prefix"]["logits"] = rnd(out["prefix"]["logits"])
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), fname), "w"))
    print("wrote", fname, len(out["steps"]), "steps")


if __name__ == "__main__":
    for which in (sys.argv[1:] or ["tiny", "wide"]):
        main(which)
