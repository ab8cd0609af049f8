"""Regenerates the compute_hash digests of tests/golden/xxh64.json with the independent `xxhash` package."""
import json, struct, xxhash

def h(tokens, prefix=None):
    d = (struct.pack("<Q", prefix) if prefix is not None else b"") + struct.pack(f"<{len(tokens)}q", *tokens)
    return xxhash.xxh64(d, seed=0).hexdigest()

if __name__ == "__main__":
    a = h([1, 2, 3, 4])
    print(json.dumps([h([1, 2, 3, 4, 5]), h([1, 2, 3, 4, 6]), h([1, 2, 3, 4, 5], 12345), a,
                      h([5, 6, 7, 8], int(a, 16)), h([9, 10, 11, 12], int(a, 16))], indent=1))
