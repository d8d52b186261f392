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


def test_alexnet_oracle_pool_against_torch_and_tie_rule():
    """MaxPool2D(3, 2) 'valid' as the oracle restates it: equal to torch's max_pool2d (values and
    gradient) on tie-free float maps, and on tied maps the FIRST maximum in row-major window order takes
    the whole gradient, overlapping windows adding up."""
    import torch

    from oracle import alexnet_oracle as A

    rng = np.random.default_rng(3)
    x = torch.tensor(rng.normal(size=(2, 9, 8, 5)), requires_grad=True)
    y = A.maxpool3s2_valid(x)
    g = torch.tensor(rng.normal(size=tuple(y.shape)))
    y.backward(g)
    xt = x.detach().clone().requires_grad_(True)
    yt = torch.nn.functional.max_pool2d(xt.permute(0, 3, 1, 2), 3, 2).permute(0, 2, 3, 1)
    yt.backward(g)
    assert torch.equal(y.detach(), yt.detach()) and torch.allclose(x.grad, xt.grad)
    # all-equal 5x5 map: windows at (0,0), (0,2), (2,0), (2,2) each pick their own top-left element
    z = torch.zeros(1, 5, 5, 1, requires_grad=True)
    A.maxpool3s2_valid(z).sum().backward()
    want = torch.zeros(5, 5)
    want[0, 0] = want[0, 2] = want[2, 0] = want[2, 2] = 1.0
    assert torch.equal(z.grad[0, :, :, 0], want)
    # a maximum shared by two overlapping windows collects both gradients
    z2 = torch.zeros(1, 3, 5, 1)
    z2[0, 1, 2, 0] = 7.0
    z2.requires_grad_(True)
    A.maxpool3s2_valid(z2).sum().backward()
    assert float(z2.grad[0, 1, 2, 0]) == 2.0 and float(z2.grad.sum()) == 2.0
