"""XXH64 restated from the published algorithm (xxHash specification, "XXH64 algorithm description").

Third-party arithmetic that is NOT under /root/reference: the reference calls `xxhash_rust::xxh64::xxh64(&data, 0)`
(Cargo.toml:39 `xxhash-rust = "0.8"` feature xxh64, unpinned — no Cargo.lock; call site
src/engine/block_manager.rs:7,122).  xxhash-rust implements canonical XXH64, so parity is anchored on the
canonical known-answer vectors (tests/golden/xxh64.json) and on the reference's own call site semantics:
data = [prefix_hash as u64 LE]? || token_ids as i64 LE, seed 0 (block_manager.rs:109-123)."""
import struct

_M = (1 << 64) - 1
P1, P2, P3, P4, P5 = (0x9E3779B185EBCA87, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9, 0x85EBCA77C2B2AE63,
                      0x27D4EB2F165667C5)


def _rotl(x, r):
    return ((x << r) | (x >> (64 - r))) & _M


def _round(acc, inp):
    acc = (acc + inp * P2) & _M
    return (_rotl(acc, 31) * P1) & _M


def _merge(acc, val):
    acc ^= _round(0, val)
    return (acc * P1 + P4) & _M


def xxh64(data: bytes, seed: int = 0) -> int:
    n = len(data)
    i = 0
    if n >= 32:
        v1, v2, v3, v4 = (seed + P1 + P2) & _M, (seed + P2) & _M, seed & _M, (seed - P1) & _M
        while i + 32 <= n:
            a, b, c, d = struct.unpack_from("<4Q", data, i)
            v1, v2, v3, v4 = _round(v1, a), _round(v2, b), _round(v3, c), _round(v4, d)
            i += 32
        h = (_rotl(v1, 1) + _rotl(v2, 7) + _rotl(v3, 12) + _rotl(v4, 18)) & _M
        h = _merge(h, v1); h = _merge(h, v2); h = _merge(h, v3); h = _merge(h, v4)
    else:
        h = (seed + P5) & _M
    h = (h + n) & _M
    while i + 8 <= n:
        (k,) = struct.unpack_from("<Q", data, i)
        h ^= _round(0, k)
        h = (_rotl(h, 27) * P1 + P4) & _M
        i += 8
    if i + 4 <= n:
        (k,) = struct.unpack_from("<I", data, i)
        h ^= (k * P1) & _M
        h = (_rotl(h, 23) * P2 + P3) & _M
        i += 4
    while i < n:
        h ^= (data[i] * P5) & _M
        h = (_rotl(h, 11) * P1) & _M
        i += 1
    h ^= h >> 33
    h = (h * P2) & _M
    h ^= h >> 29
    h = (h * P3) & _M
    h ^= h >> 32
    return h


def compute_hash(token_ids, prefix_hash=None) -> int:
    """BlockManager::compute_hash, src/engine/block_manager.rs:109-123."""
    data = b""
    if prefix_hash is not None:  # block_manager.rs:113-115
        data += struct.pack("<Q", prefix_hash)
    data += struct.pack(f"<{len(token_ids)}q", *[int(t) for t in token_ids])  # block_manager.rs:118-120
    return xxh64(data, 0)  # block_manager.rs:122
