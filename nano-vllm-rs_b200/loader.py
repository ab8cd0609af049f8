"""Weight loading from HF safetensors checkpoints — mirror of the reference's ModelLoader (src/utils/loader.rs:13-198).

The reference never calls its loader (ModelRunner::load_weights is a stub, model_runner.rs:437-448) and every linear's
load_weight validates shapes and discards the tensor (linear.rs:154-172).  Here the file format is parsed on the host
(safetensors = 8-byte little-endian header length, JSON header {name: {dtype, shape, data_offsets}}, raw little-endian
data) and every tensor goes to nvllm_runner_load_weight, which keeps this rank's tensor-parallel slice and packs
q/k/v -> qkv_proj and gate/up -> gate_up_proj (create_standard_loader's mapping, loader.rs:187-198).
Only host-side file parsing lives here; there is no compute and no CPU fallback of any kernel."""
import json
import os
import struct

import numpy as np

# loader.rs:187-198 — original module name -> (packed parameter, shard id)
PACKED_MODULES_MAPPING = {
    "q_proj": ("qkv_proj", "q"), "k_proj": ("qkv_proj", "k"), "v_proj": ("qkv_proj", "v"),
    "gate_proj": ("gate_up_proj", "0"), "up_proj": ("gate_up_proj", "1"),
}
_DTYPES = {"BF16": (np.uint16, 2), "F16": (np.float16, 2), "F32": (np.float32, 4)}


def find_packed_mapping(tensor_name):
    """ModelLoader::find_packed_mapping (loader.rs:175-183): 'model.layers.0.self_attn.q_proj.weight' ->
    ('model.layers.0.self_attn.qkv_proj.weight', 'q'); None for unpacked tensors."""
    for original, (packed, shard) in PACKED_MODULES_MAPPING.items():
        if original in tensor_name:
            return tensor_name.replace(original, packed), shard
    return None


def read_safetensors(path):
    """Yields (name, dtype string, shape, raw little-endian bytes as a numpy view)."""
    data = np.memmap(path, dtype=np.uint8, mode="r")
    if data.size < 8:
        raise ValueError(f"{path}: not a safetensors file")
    (hlen,) = struct.unpack("<Q", bytes(data[:8]))
    if 8 + hlen > data.size:
        raise ValueError(f"{path}: header length {hlen} exceeds the file size")
    header = json.loads(bytes(data[8 : 8 + hlen]).decode("utf-8"))
    base = 8 + hlen
    for name, meta in header.items():
        if name == "__metadata__":
            continue
        b, e = meta["data_offsets"]
        yield name, meta["dtype"], tuple(meta["shape"]), data[base + b : base + e]


def to_bf16_bits(dtype, shape, raw):
    """Any supported checkpoint dtype -> uint16 bf16 bit patterns (round-to-nearest-even for F32 / F16)."""
    if dtype not in _DTYPES:
        raise ValueError(f"unsupported safetensors dtype {dtype}")
    np_dt, _ = _DTYPES[dtype]
    a = np.frombuffer(raw, dtype=np_dt).reshape(shape)
    if dtype == "BF16":
        return np.ascontiguousarray(a)
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    return ((u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) >> np.uint32(16)).astype(np.uint16)


def write_safetensors(path, tensors):
    """Minimal writer (tests / tooling): tensors = {name: uint16 bf16-bit array | float32 | float16 array}."""
    header, blobs, off = {}, [], 0
    for name, a in tensors.items():
        a = np.ascontiguousarray(a)
        dt = {np.dtype(np.uint16): "BF16", np.dtype(np.float32): "F32", np.dtype(np.float16): "F16"}[a.dtype]
        raw = a.tobytes()
        header[name] = {"dtype": dt, "shape": list(a.shape), "data_offsets": [off, off + len(raw)]}
        blobs.append(raw)
        off += len(raw)
    h = json.dumps(header, separators=(",", ":")).encode()
    h += b" " * ((8 - len(h) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(h)) + h + b"".join(blobs))


class ModelLoader:
    """ModelLoader::load_safetensors / load_safetensors_dir (loader.rs:63-125) targeting a ModelRunner."""

    def __init__(self, runner):
        self.runner = runner

    def load_safetensors(self, path):
        n = 0
        for name, dtype, shape, raw in read_safetensors(path):
            if len(shape) not in (1, 2):
                raise ValueError(f"{name}: unsupported rank {len(shape)}")
            self.runner.load_weight(name, to_bf16_bits(dtype, shape, raw))   # slices + packs on the device side
            n += 1
        return n

    def load_safetensors_dir(self, dir_path):
        if not os.path.isdir(dir_path):
            raise ValueError(f"Path is not a directory: {dir_path}")
        files = sorted(f for f in os.listdir(dir_path) if f.endswith(".safetensors"))   # loader.rs:103-116
        if not files:
            raise ValueError(f"No safetensors files found in directory: {dir_path}")
        return sum(self.load_safetensors(os.path.join(dir_path, f)) for f in files)
