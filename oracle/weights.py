"""Seeded synthetic weights shared bit-for-bit with the device (nvllm_init_weight, csrc/fused_ops.cu).

The reference leaves weight init unspecified (linear.rs:37 calls a candle overload that does not exist;
the LM head is re-randomised on every call, embed_head.rs:309-318).  SURVEY.md §8d: seeded N(0, 0.02^2),
identical bytes for oracle and GPU.  To be reproducible without transcendental functions the normal is
approximated by an Irwin-Hall sum: the four 16-bit fields of splitmix64(key + (i+1)*golden), minus their
mean 131070, times std/37837.23 — integer arithmetic and one exact int->f32 multiply."""
import numpy as np

from .bf16 import f32_to_bf16_bits

_GOLD = np.uint64(0x9E3779B97F4A7C15)
_SCALE = np.float32(2.6428998e-05)


def _splitmix64(x):
    x = x ^ (x >> np.uint64(30)); x = x * np.uint64(0xBF58476D1CE4E5B9)
    x = x ^ (x >> np.uint64(27)); x = x * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def init_weight_f32(count, seed, tensor_id, std, chunk=1 << 24):
    """float32 values before the bf16 rounding."""
    key = np.uint64(seed) ^ (np.uint64(tensor_id) << np.uint64(40))
    scale = np.float32(std) * _SCALE
    out = np.empty(count, dtype=np.float32)
    with np.errstate(over="ignore"):
        for s in range(0, count, chunk):
            e = min(count, s + chunk)
            i = np.arange(s + 1, e + 1, dtype=np.uint64)
            r = _splitmix64(key + i * _GOLD)
            m = np.uint64(0xFFFF)
            ssum = ((r & m) + ((r >> np.uint64(16)) & m) + ((r >> np.uint64(32)) & m) + (r >> np.uint64(48))).astype(np.int64)
            out[s:e] = (ssum - 131070).astype(np.float32) * scale
    return out


def init_weight_bf16_bits(shape, seed, tensor_id, std):
    n = int(np.prod(shape))
    return f32_to_bf16_bits(init_weight_f32(n, seed, tensor_id, std)).reshape(shape)
