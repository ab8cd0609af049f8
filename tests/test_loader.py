"""Safetensors loading (reference src/utils/loader.rs).  CPU: format round-trip, dtype conversion,
packed-module mapping.  GPU: a checkpoint written from the oracle's seeded tensors and loaded through
nvllm_runner_load_weight produces exactly the logits of the engine's own seeded init (same bytes, same slices)."""
import numpy as np
import pytest

from oracle import qwen3 as Q
from oracle.bf16 import f32_to_bf16_bits


def test_packed_mapping_loader_rs_175_198():
    from nano_vllm_rs_b200 import loader as L
    assert L.find_packed_mapping("model.layers.3.self_attn.q_proj.weight") == ("model.layers.3.self_attn.qkv_proj.weight", "q")
    assert L.find_packed_mapping("model.layers.0.mlp.up_proj.weight") == ("model.layers.0.mlp.gate_up_proj.weight", "1")
    assert L.find_packed_mapping("model.norm.weight") is None


def test_safetensors_roundtrip_and_dtype_conversion(tmp_path):
    from nano_vllm_rs_b200 import loader as L
    rng = np.random.default_rng(0)
    f32 = rng.standard_normal((5, 8)).astype(np.float32)
    bf = f32_to_bf16_bits(rng.standard_normal((3, 16)).astype(np.float32))
    f16 = rng.standard_normal(7).astype(np.float16)
    p = tmp_path / "m.safetensors"
    L.write_safetensors(p, {"a": f32, "b": bf, "c": f16})
    got = {n: (dt, sh, L.to_bf16_bits(dt, sh, raw)) for n, dt, sh, raw in L.read_safetensors(p)}
    assert got["a"][0] == "F32" and got["a"][1] == (5, 8) and np.array_equal(got["a"][2], f32_to_bf16_bits(f32))
    assert got["b"][0] == "BF16" and np.array_equal(got["b"][2], bf)
    assert np.array_equal(got["c"][2], f32_to_bf16_bits(f16.astype(np.float32)))
    with pytest.raises(ValueError):
        list(L.read_safetensors(__file__))


@pytest.mark.gpu
def test_loaded_checkpoint_equals_seeded_init(cuda, tmp_path):
    from nano_vllm_rs_b200 import engine as E
    from nano_vllm_rs_b200 import loader as L
    ocfg = Q.arch_config("tiny")
    w = Q.seeded_weights(ocfg, 77, 0.05)
    files = {}
    for i, (name, a) in enumerate(w.items()):   # two shards, mixed dtypes (norm weights as F32)
        files.setdefault(i % 2, {})[name] = a.astype(np.float32) if a.ndim == 1 else f32_to_bf16_bits(a)
    for i, t in files.items():
        L.write_safetensors(tmp_path / f"model-{i:05d}-of-00002.safetensors", t)
    cfg = E.Config("tiny", kvcache_block_size=16, num_kvcache_blocks=32)
    rng = np.random.default_rng(1)
    prompts = [rng.integers(0, 1024, n).tolist() for n in (9, 30)]
    outs = []
    for mode in ("seeded", "loaded"):
        sched, runner = E.Scheduler(cfg), E.ModelRunner(cfg)
        if mode == "seeded":
            runner.init_weights(77, 0.05)
        else:
            assert L.ModelLoader(runner).load_safetensors_dir(str(tmp_path)) == len(w)
        for p in prompts:
            sched.add_sequence(p, E.SamplingParams(temperature=0.0, max_tokens=4, ignore_eos=True))
        trace = []
        while not sched.is_finished():
            b = sched.schedule()
            toks, logits = runner.execute_model(b, return_logits=True)
            trace.append((toks.copy(), logits.copy()))
            sched.postprocess(toks)
        outs.append(trace)
    for (t0, l0), (t1, l1) in zip(*outs):
        assert np.array_equal(t0, t1) and np.array_equal(l0, l1)
