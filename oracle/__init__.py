"""CPU oracle — TEST INFRASTRUCTURE ONLY.

A plain numpy / pure-Python (+ one small C file) restatement of the reference's algorithm for the
paged-attention decode hot path of ssvgopal/nano-vllm-rs.  Each function cites the reference
file:line it follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (nano-vllm-rs_b200/) never does.

Parity pinning: the reference cannot be compiled or run here (no Rust toolchain, invalid
manifest, type errors on the path), and it holds NO numeric golden vectors for attention outputs, logits or
tokens.  What it does pin, and what tests/test_oracle_*.py check this oracle against:
  * block accounting / scheduler / sequence / runner-input known answers of the reference's in-file unit tests
    (block_manager.rs:426-539, sequence.rs:288-303, scheduler.rs:417-528, model_runner.rs:501-520),
  * XXH64 (third-party xxhash-rust 0.8, call site block_manager.rs:122): canonical known-answer vectors,
    cross-checked against the independent `xxhash` Python package and tests/golden/xxh64.json,
  * formula-level facts: scale(64)=0.125, causal mask pattern, RoPE identities, RMS≈1, SiLU·mul values,
    greedy argmax (attention.rs:456-497, rotary_embedding.rs:322-386, layernorm.rs:195-221,
    activation.rs:264-290, sampler.rs:333-342).
Floating-point outputs (attention, logits, tokens) are therefore "parity unpinned" by the reference itself;
they are additionally cross-checked against independent implementations: torch-CPU layer math (tests/golden/layers.json)
and, end to end (prefill + decode steps over the paged cache), Hugging Face transformers' Qwen3ForCausalLM in fp32
(tests/golden/qwen3_hf.json, tests/test_oracle_model.py: within 0.04 sigma = bf16 rounding noise)."""
