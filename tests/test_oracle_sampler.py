"""CPU tests: the oracle's sampler (oracle/layers.py, restating sampler.rs:26-254) against every known answer the
reference's own unit tests state (sampler.rs:331-462), and the seeded-uniform contract against tests/golden/sampler.json
(generator: tests/golden/make_sampler_golden.py, plain-integer arithmetic independent of the numpy code)."""
import json
import os

import numpy as np

from oracle import layers as O

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sampler.json")))


def test_reference_unit_test_facts():
    # test_greedy_sampling (sampler.rs:333-342), test_temperature_scaling (:345-356), test_batch_sampling (:423-441)
    assert O.sample_single(np.array([1.0, 2.0, 5.0, 1.5], np.float32), 0.0, None, None, 0)[0] == 2
    assert O.sampler_forward(np.array([[1.0, 2.0, 3.0]], np.float32), [0.0], None, None, [0])[0] == 2
    toks = O.sampler_forward(np.array([[1, 2, 3], [4, 5, 6]], np.float32), [0.0, 1.0], None, None, [1, 2])
    assert toks.shape == (2,) and toks[0] == 2 and 0 <= toks[1] < 3
    # test_top_k_filtering (:359-374): keeps 5, 4, 3 at indices 1, 3, 4
    f = O.apply_top_k(np.array([1.0, 5.0, 2.0, 4.0, 3.0], np.float32), 3)
    assert f[1] == 5.0 and f[3] == 4.0 and f[4] == 3.0 and f[0] == -np.inf and f[2] == -np.inf
    # test_top_p_filtering (:377-389): the highest logit survives; here p(10) = 0.993 >= 0.9 so it is the only survivor
    f = O.apply_top_p(np.array([0.0, 10.0, 5.0, 1.0], np.float32), 0.9)
    assert f[1] == 10.0 and np.isfinite(f).sum() == 1
    # test_gumbel_sampling (:392-404): finite for every representable uniform of the contract, incl. the extremes
    u = np.array([0.5 * 2.0 ** -23, 0.5, 1.0 - 0.5 * 2.0 ** -23], np.float32)
    assert np.isfinite(O.sample_gumbel(u)).all()
    # test_multinomial_sampling_deterministic (:444-462): a 20-nat gap is never overturned by Gumbel noise
    picks = [O.sample_single(np.array([-10.0, 10.0, -10.0], np.float32), 1.0, None, None, O.sampler_row_seed(3, i, 1))[0]
             for i in range(200)]
    assert picks.count(1) == 200


def test_top_k_stable_tie_rule_and_clamps():
    x = np.array([2.0, 3.0, 3.0, 1.0, 3.0, 3.0], np.float32)
    f = O.apply_top_k(x, 3)                      # stable descending sort: the first three 3.0 by index
    assert np.nonzero(np.isfinite(f))[0].tolist() == [1, 2, 4]
    assert np.isfinite(O.apply_top_k(x, 100)).all()                      # k.min(vocab), sampler.rs:117
    # top_k == 0 / top_p outside (0, 1) mean None in the build (batch_sample maps None to 0 / 1.0, sampler.rs:235-248)
    _, filt, _ = O.sample_single(x, 1.0, 1.0, 0, 5)
    assert np.isfinite(filt).all()


def test_top_p_cut_includes_the_crossing_element():
    x = np.log(np.array([0.5, 0.3, 0.15, 0.05], np.float32))
    assert np.isfinite(O.apply_top_p(x, 0.5)).sum() == 1      # cum >= p at the first element (sampler.rs:172-178)
    assert np.isfinite(O.apply_top_p(x, 0.6)).sum() == 2
    assert np.isfinite(O.apply_top_p(x, 0.96)).sum() == 4
    # filters compose as in sample_single (sampler.rs:91-103): top-k first, top-p over the renormalised survivors
    _, filt, _ = O.sample_single(x, 1.0, 0.7, 2, 9)            # survivors {0.5, 0.3} -> renormalised 0.625, 0.375 -> both
    assert np.nonzero(np.isfinite(filt))[0].tolist() == [0, 1]


def test_seeded_contract_golden():
    for x, want in GOLD["mix64"].items():
        assert int(O.mix64(np.uint64(int(x)))) == int(want)
    for r in GOLD["row_seed"]:
        assert O.sampler_row_seed(r["engine_seed"], r["seq_id"], r["len"]) == int(r["value"])
    ub = GOLD["uniform_bits"]
    u = O.sampler_uniform(int(ub["seed"]), len(ub["bits"]))
    want = (np.array(ub["bits"], np.float32) + np.float32(0.5)) * np.float32(2.0 ** -23)
    assert (u == want).all() and (u > 0).all() and (u < 1).all()
    logits = np.array(GOLD["logits"], np.float32)
    for i, r in enumerate(GOLD["rows"]):
        p, k = (r["top_p"] if r["top_p"] > 0 else None), (r["top_k"] or None)
        tok, filt, _ = O.sample_single(logits[i], r["temperature"], p, k, int(r["row_seed"]))
        assert tok == r["token"]
        if r["kept"] is not None:
            assert np.nonzero(np.isfinite(filt))[0].tolist() == r["kept"]


def test_gumbel_max_samples_the_softmax():
    row = np.array([2.0, 1.0, 0.5, 0.0, -0.5, -1.0, 1.5, -3.0], np.float32)
    n = 20000
    toks = [O.sample_single(row, 1.0, None, None, O.sampler_row_seed(11, i, 4))[0] for i in range(n)]
    freq = np.bincount(toks, minlength=8) / n
    prob = np.exp(row - row.max()); prob /= prob.sum()
    assert np.abs(freq - prob).max() < 0.012
