"""bfloat16 emulation in numpy (round-to-nearest-even), so the oracle can round at the same tensor
boundaries as the reference's bf16 tensors (candle `to_dtype(BF16)`)."""
import numpy as np


def f32_to_bf16_bits(x):
    """float32 array -> uint16 bf16 bit patterns (RNE; NaN kept quiet)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    rounded = (u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) >> np.uint32(16)
    nan = np.isnan(x)
    if nan.any():
        rounded = np.where(nan, (u >> np.uint32(16)) | np.uint32(0x40), rounded)
    return rounded.astype(np.uint16)


def bf16_bits_to_f32(b):
    b = np.ascontiguousarray(b, dtype=np.uint16)
    return (b.astype(np.uint32) << np.uint32(16)).view(np.float32)


def round_bf16(x):
    """float32 -> nearest bf16, returned as float32."""
    return bf16_bits_to_f32(f32_to_bf16_bits(x))
