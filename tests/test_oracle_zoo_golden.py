"""The oracles of the two larq-zoo models (oracle/alexnet_oracle.py, oracle/bireal_oracle.py) against the
committed vectors tests/golden/zoo_steps.npz (written by ``python -m oracle.make_golden zoo``): loss,
logits, norm and sum of every gradient, absolute sums of the kept activations of one seeded train step
each.  The vectors pin the oracles against drift; they are the oracles' own output (parity unpinned:
larq-zoo is not in the reference tree and cannot be installed here).  No GPU."""
import os

import numpy as np

from oracle import make_golden as G

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "zoo_steps.npz")


def test_zoo_oracles_reproduce_their_golden_steps():
    gold = np.load(GOLDEN)
    seen = set()
    for name, run in G.zoo_cases():
        got = G.zoo_summary(run())
        for k, v in got.items():
            key = f"{name}/{k}"
            assert key in gold.files, key
            seen.add(key)
            np.testing.assert_allclose(v, gold[key], rtol=1e-9, atol=1e-12, err_msg=key)
    assert seen == set(gold.files)
    # the integer maps are exact
    assert float(gold["alexnet/act_abs_sum_y1"]) == int(gold["alexnet/act_abs_sum_y1"])
    assert float(gold["bireal/act_abs_sum_y0"]) == int(gold["bireal/act_abs_sum_y0"])
