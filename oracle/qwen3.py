"""numpy restatement of the reference's model wiring with a paged KV cache (TEST INFRASTRUCTURE — oracle/__init__.py).

Qwen3Model::forward (models/qwen3.rs:487-505) -> Qwen3Layer::forward (:372-392) -> Qwen3Attention::forward
(:208-240) -> Attention::forward (layers/attention.rs:65-99) / Qwen3MLP::forward (qwen3.rs:305-314) ->
ParallelLMHead::forward (layers/embed_head.rs:250-306) -> Sampler::greedy_sample (layers/sampler.rs:109-112).
f32 math on bf16-representable values, rounded to bf16 wherever the reference materialises a model-dtype tensor.
Deviations N1 of SURVEY.md §8 (qk-norm, explicit head_dim, f32 RoPE table, context_lens truncation, seeded weights
incl. the LM head) are flag-controlled / stated in oracle/layers.py."""
import numpy as np

from . import layers as L
from .bf16 import bf16_bits_to_f32
from .weights import init_weight_bf16_bits

ARCHS = {
    "qwen3-0.6b": dict(vocab_size=151936, hidden_size=1024, intermediate_size=3072, num_hidden_layers=28,
                       num_attention_heads=16, num_key_value_heads=8, head_dim=128, tie_word_embeddings=True),
    "qwen3-8b": dict(vocab_size=151936, hidden_size=4096, intermediate_size=12288, num_hidden_layers=36,
                     num_attention_heads=32, num_key_value_heads=8, head_dim=128, tie_word_embeddings=False),
    "qwen3-32b": dict(vocab_size=151936, hidden_size=5120, intermediate_size=25600, num_hidden_layers=64,
                      num_attention_heads=64, num_key_value_heads=8, head_dim=128, tie_word_embeddings=False),
    "qwen3-0.6b-2l": dict(vocab_size=2048, hidden_size=1024, intermediate_size=3072, num_hidden_layers=2,
                          num_attention_heads=16, num_key_value_heads=8, head_dim=128, tie_word_embeddings=False),
    "tiny-tp8": dict(vocab_size=2048, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                     num_attention_heads=16, num_key_value_heads=8, head_dim=128, tie_word_embeddings=False),
    "tiny": dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                 num_attention_heads=4, num_key_value_heads=2, head_dim=128, tie_word_embeddings=False),
}


def arch_config(name, **over):
    c = dict(rms_norm_eps=1e-6, rope_theta=1e6, qk_norm=True, max_position_embeddings=4096)
    c.update(ARCHS[name])
    c.update(over)
    return c


def seeded_weights(cfg, seed, std):
    """Logical (unsharded) tensors with the ids csrc/host/model_runner.cu uses: 1 embed, 2 lm_head,
    16*(l+1) + {0 q, 1 k, 2 v, 3 o, 4 gate, 5 up, 6 down}; norm weights = 1.  HF tensor names."""
    H, Hkv, D, hid, I, V = (cfg[k] for k in ("num_attention_heads", "num_key_value_heads", "head_dim", "hidden_size",
                                             "intermediate_size", "vocab_size"))
    t = lambda shape, tid: bf16_bits_to_f32(init_weight_bf16_bits(shape, seed, tid, std))
    w = {"model.embed_tokens.weight": t((V, hid), 1), "model.norm.weight": np.ones(hid, np.float32)}
    if not cfg["tie_word_embeddings"]:
        w["lm_head.weight"] = t((V, hid), 2)
    for l in range(cfg["num_hidden_layers"]):
        b, p = 16 * (l + 1), f"model.layers.{l}."
        w[p + "input_layernorm.weight"] = np.ones(hid, np.float32)
        w[p + "post_attention_layernorm.weight"] = np.ones(hid, np.float32)
        w[p + "self_attn.q_norm.weight"] = np.ones(D, np.float32)
        w[p + "self_attn.k_norm.weight"] = np.ones(D, np.float32)
        w[p + "self_attn.q_proj.weight"] = t((H * D, hid), b + 0)
        w[p + "self_attn.k_proj.weight"] = t((Hkv * D, hid), b + 1)
        w[p + "self_attn.v_proj.weight"] = t((Hkv * D, hid), b + 2)
        w[p + "self_attn.o_proj.weight"] = t((hid, H * D), b + 3)
        w[p + "mlp.gate_proj.weight"] = t((I, hid), b + 4)
        w[p + "mlp.up_proj.weight"] = t((I, hid), b + 5)
        w[p + "mlp.down_proj.weight"] = t((hid, I), b + 6)
    return w


def shard_weights(w, cfg, rank, tp):
    """This rank's tensor-parallel slices, by the reference's sharding arithmetic: q/k/v/gate/up rows and o/down columns
    divided by tp (qwen3.rs:157-159,273-292; linear.rs:154-172,202,249-267,303-307,401), LM head rows = vocab shard
    (embed_head.rs:142-161), embedding and norms replicated.  Returns (sharded cfg, sharded weights)."""
    H, Hkv, D, I, V = (cfg[k] for k in ("num_attention_heads", "num_key_value_heads", "head_dim", "intermediate_size", "vocab_size"))
    assert H % tp == 0 and Hkv % tp == 0 and I % tp == 0 and V % tp == 0
    c = dict(cfg, num_attention_heads=H // tp, num_key_value_heads=Hkv // tp, intermediate_size=I // tp)
    rows = lambda a, n: a[rank * n:(rank + 1) * n]
    cols = lambda a, n: a[:, rank * n:(rank + 1) * n]
    out = {}
    for name, a in w.items():
        if name.endswith(("q_proj.weight",)):
            out[name] = rows(a, H // tp * D)
        elif name.endswith(("k_proj.weight", "v_proj.weight")):
            out[name] = rows(a, Hkv // tp * D)
        elif name.endswith(("gate_proj.weight", "up_proj.weight")):
            out[name] = rows(a, I // tp)
        elif name.endswith("o_proj.weight"):
            out[name] = cols(a, H // tp * D)
        elif name.endswith("down_proj.weight"):
            out[name] = cols(a, I // tp)
        else:
            out[name] = a
    head = w["model.embed_tokens.weight"] if cfg["tie_word_embeddings"] else w["lm_head.weight"]
    out["__lm_head_shard__"] = rows(head, V // tp)
    return c, out


class Qwen3Oracle:
    def __init__(self, cfg, weights, num_blocks, block_size, allreduce=None):
        """allreduce: callable(np.ndarray) -> sum over tensor-parallel ranks (the linear.rs:236 TODO); None = TP 1."""
        self.cfg, self.w, self.bs = cfg, weights, block_size
        self.allreduce = allreduce
        Hkv, D = cfg["num_key_value_heads"], cfg["head_dim"]
        n = cfg["num_hidden_layers"]
        # create_kv_cache, model_runner.rs:364-396: per layer K and V zeros [num_blocks, block_size, Hkv, D]
        self.k_cache = [np.zeros((num_blocks, block_size, Hkv, D), np.float32) for _ in range(n)]
        self.v_cache = [np.zeros((num_blocks, block_size, Hkv, D), np.float32) for _ in range(n)]
        self.rope = L.rope_cos_sin_table(cfg["max_position_embeddings"], D, cfg["rope_theta"])
        self.scale = 1.0 / np.sqrt(np.float32(D))  # attention.rs:45

    def forward(self, inp):
        """inp: oracle.host.StepInputs (or any object with the same fields). Returns f32 logits [num_seqs, vocab]."""
        c, w = self.cfg, self.w
        H, Hkv, D, eps = c["num_attention_heads"], c["num_key_value_heads"], c["head_dim"], c["rms_norm_eps"]
        h = L.embedding(w["model.embed_tokens.weight"], inp.input_ids)  # qwen3.rs:492
        N = h.shape[0]
        mlp_out = None
        for l in range(c["num_hidden_layers"]):
            p = f"model.layers.{l}."
            if l == 0:
                normed = L.rms_norm(h, w[p + "input_layernorm.weight"], eps)  # qwen3.rs:378
            else:
                normed, h = L.add_rms_norm(mlp_out, h, w[p + "input_layernorm.weight"], eps)  # :389 then :378
            wqkv = np.concatenate([w[p + "self_attn.q_proj.weight"], w[p + "self_attn.k_proj.weight"],
                                   w[p + "self_attn.v_proj.weight"]], 0)  # linear.rs:331-340 packing
            qkv = L.linear(normed, wqkv)  # qwen3.rs:216
            q = qkv[:, : H * D].reshape(N, H, D)  # split_qkv + reshape, qwen3.rs:219-226
            k = qkv[:, H * D : (H + Hkv) * D].reshape(N, Hkv, D)
            v = qkv[:, (H + Hkv) * D :].reshape(N, Hkv, D)
            qn = w[p + "self_attn.q_norm.weight"] if c["qk_norm"] else None
            kn = w[p + "self_attn.k_norm.weight"] if c["qk_norm"] else None
            q, k = L.qknorm_rope(q, k, inp.positions, self.rope, qn, kn, eps)  # qwen3.rs:229
            L.store_kv_cache(k, v, self.k_cache[l], self.v_cache[l], inp.slot_mapping)  # attention.rs:78-81
            if not inp.is_prefill:  # attention.rs:92-95
                a = L.flash_attention_decode(q, self.k_cache[l], self.v_cache[l], inp.block_tables, inp.context_lens, self.scale)
            elif inp.block_tables is not None:  # attention.rs:85-87
                a = L.flash_attention_varlen_with_cache(q, self.k_cache[l], self.v_cache[l], inp.block_tables,
                                                        inp.cu_seqlens_q, inp.cu_seqlens_k, self.scale)
            else:  # attention.rs:88-91
                a = L.flash_attention_varlen(q, k, v, inp.cu_seqlens_q, inp.cu_seqlens_k, self.scale)
            o = L.linear(a.reshape(N, H * D), w[p + "self_attn.o_proj.weight"])  # qwen3.rs:236
            if self.allreduce is not None:  # row-parallel partial sums, each a bf16 tensor (linear.rs:228-239)
                o = L.round_bf16(self.allreduce(o))
            normed, h = L.add_rms_norm(o, h, w[p + "post_attention_layernorm.weight"], eps)  # :382-385
            wgu = np.concatenate([w[p + "mlp.gate_proj.weight"], w[p + "mlp.up_proj.weight"]], 0)  # linear.rs:421-433
            act = L.silu_and_mul(L.linear(normed, wgu))  # qwen3.rs:307-310
            mlp_out = L.linear(act, w[p + "mlp.down_proj.weight"])  # :312
            if self.allreduce is not None:
                mlp_out = L.round_bf16(self.allreduce(mlp_out))
        normed, h = L.add_rms_norm(mlp_out, h, w["model.norm.weight"], eps)  # qwen3.rs:389,501
        if inp.is_prefill:
            normed = L.extract_last_tokens(normed, inp.cu_seqlens_q)  # embed_head.rs:250-289
        if "__lm_head_shard__" in w:  # vocab-sharded head: this rank's logits [num_seqs, vocab / tp]
            return L.linear(normed, w["__lm_head_shard__"], out_bf16=False)
        head = w["model.embed_tokens.weight"] if c["tie_word_embeddings"] else w["lm_head.weight"]
        return L.linear(normed, head, out_bf16=False)  # embed_head.rs:292-306 (f32 logits for the sampler)
