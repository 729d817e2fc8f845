"""CPU: the oracle reproduces the committed golden vectors (tests/golden/, written by tests/golden/make_golden.py), so
a change of the oracle's arithmetic cannot pass unnoticed.  The fixtures are oracle outputs (TensorFlow is not
installable here); parity with TF itself is unpinned."""
import importlib.util
import json
import os

import numpy as np
import pytest

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "make_golden.py"))
mk = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mk)


def _close(a, b, rtol=1e-7, atol=1e-12):
    return abs(a - b) <= atol + rtol * abs(b)


@pytest.mark.parametrize("variant", ["c1", "c1_dropout", "c2", "c4"])
def test_oracle_reproduces_golden_train_steps(variant):
    want = json.load(open(os.path.join(HERE, f"tiny_{variant}.json")))
    got = mk.run(variant)
    for s, (g, w) in enumerate(zip(got["steps"], want["steps"])):
        assert g.keys() == w.keys()
        for k in w:
            assert _close(g[k], w[k]), (variant, s, k, g[k], w[k])
    for k, w in want["grad_norms_step0"].items():
        assert _close(got["grad_norms_step0"][k], w, rtol=1e-6, atol=1e-15), (variant, k)
    for k, w in want["weights_after"].items():
        assert _close(got["weights_after"][k], w, rtol=1e-7, atol=1e-9), (variant, k)
    if "greedy_ids_after" in want:
        assert got["greedy_ids_after"] == want["greedy_ids_after"]


def test_oracle_reproduces_golden_full_config():
    """SURVEY 8(c) known-answer item (2): full C1 shapes, B = 16, three Adam steps (about half a minute of float64 oracle)."""
    want = json.load(open(os.path.join(HERE, "full_c1.json")))
    got = mk.run_full()
    for s, (g, w) in enumerate(zip(got["steps"], want["steps"])):
        assert g.keys() == w.keys()
        for k in w:
            assert _close(g[k], w[k]), (s, k, g[k], w[k])
    for k, w in want["grad_norms_step0"].items():
        assert _close(got["grad_norms_step0"][k], w, rtol=1e-6, atol=1e-15), k
    for k, w in want["weights_after"].items():
        assert _close(got["weights_after"][k], w, rtol=1e-7, atol=1e-9), k
    # ids may only differ where the committed margin says near-tie (thread-count dependent summation order in BLAS)
    m = np.asarray(want["greedy_rel_margin"])
    for b, (g, w) in enumerate(zip(got["greedy_ids_after"], want["greedy_ids_after"])):
        bad = [t for t in range(len(w)) if g[t] != w[t]]
        assert not bad or m[b, bad[0]] < 1e-5, (b, bad[0], float(m[b, bad[0]]))


def test_oracle_reproduces_golden_generator_batches():
    want = json.load(open(os.path.join(HERE, "generator.json")))
    got = mk.generator_case()
    assert got["order_epoch0"] == want["order_epoch0"] and got["order_epoch1"] == want["order_epoch1"]
    for g, w in zip(got["batches"], want["batches"]):
        assert g["beta_rows"] == w["beta_rows"] and g["cap"] == w["cap"] and g["target_ids"] == w["target_ids"]
        assert _close(g["betas_sum"], w["betas_sum"], rtol=1e-12)
