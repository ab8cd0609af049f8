"""Regenerates tests/golden/sampler.json: known answers of the seeded sampling contract (oracle/layers.py — the same
arithmetic as csrc/sampler.cu and ModelRunner::sample in csrc/host/model_runner.cu).  The splitmix64 finaliser values are
computed here with plain Python integers, independently of the numpy code they pin."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import layers as O  # noqa: E402

M64 = (1 << 64) - 1
GOLDEN = 0x9E3779B97F4A7C15


def mix64(x):
    x &= M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def row_seed(engine_seed, seq_id, n):
    return mix64((engine_seed & M64) ^ mix64(seq_id * GOLDEN + n))


def uniform_bits(seed, j):   # the 23-bit integer behind u[j] = (bits + 0.5) * 2^-23
    return mix64(seed + (j + 1) * GOLDEN) >> 41


def main():
    rng = np.random.default_rng(20260)
    logits = np.round(rng.standard_normal((6, 64)) * 3, 2).astype(np.float32)
    logits[4] = np.round(logits[4])          # ties
    temps = [0.0, 1.0, 0.7, 1.3, 1.0, 0.9]
    top_k = [0, 0, 8, 0, 5, 12]
    top_p = [0.0, 0.0, 0.0, 0.6, 0.8, 0.9]
    seeds = [row_seed(77, i, 10 + i) for i in range(6)]
    rows = []
    for i in range(6):
        tok, filt, _ = O.sample_single(logits[i], temps[i], top_p[i] if top_p[i] > 0 else None, top_k[i] or None, seeds[i])
        rows.append({"temperature": temps[i], "top_k": top_k[i], "top_p": top_p[i], "row_seed": str(seeds[i]), "token": tok,
                     "kept": np.nonzero(np.isfinite(filt))[0].tolist() if temps[i] != 0 else None})
    out = {"mix64": {str(x): str(mix64(x)) for x in (0, 1, 0x9E3779B97F4A7C15, M64, 1234)},
           "row_seed": [{"engine_seed": e, "seq_id": s, "len": n, "value": str(row_seed(e, s, n))}
                        for e, s, n in ((0, 0, 1), (77, 3, 129), (1234, 255, 1024), (-1, 7, 5))],
           "uniform_bits": {"seed": str(seeds[1]), "bits": [uniform_bits(seeds[1], j) for j in range(8)]},
           "logits": [[float(v) for v in r] for r in logits], "rows": rows}
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "sampler.json"), "w"), indent=1)
    print("wrote sampler.json")


if __name__ == "__main__":
    main()
