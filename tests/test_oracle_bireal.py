"""CPU checks of oracle/bireal_oracle.py (the restatement the Bi-Real GPU tests compare with): the
two quantizers' forward / backward rules as published, the stem pool against torch's max_pool2d
under TensorFlow's 'same' padding, and the optimizer's clip.  No GPU."""
import numpy as np
import torch

from oracle import bireal_oracle as O


def test_approx_sign_and_magnitude_aware_sign_rules():
    x = torch.tensor([-1.5, -1.0, -0.25, -0.0, 0.0, 0.5, 1.0, 1.0000001], dtype=torch.float64, requires_grad=True)
    y = O.approx_sign(x)
    assert y.tolist() == [-1, -1, -1, 1, 1, 1, 1, 1]                       # sign(0) = sign(-0.0) = +1
    y.backward(torch.full_like(x, 2.0))
    assert np.allclose(x.grad.numpy(), 2.0 * np.array([0, 0, 1.5, 2, 2, 1, 0, 0]))   # dy * (2 - 2|x|) * [|x| <= 1]
    w = torch.tensor(np.random.default_rng(0).normal(0, 0.7, (3, 3, 4, 5)), requires_grad=True)
    q = O.magnitude_aware_sign(w)
    scale = w.detach().abs().mean(dim=(0, 1, 2))
    assert torch.equal(q.detach(), scale * torch.where(w.detach() >= 0, 1.0, -1.0))
    q.sum().backward()
    # the scale's gradient is stopped; the sign passes the gradient where |w| <= 1
    assert torch.allclose(w.grad, scale.expand_as(w) * (w.detach().abs() <= 1))


def test_stem_pool_matches_torch_under_tf_same_padding():
    rng = np.random.default_rng(1)
    for h, w in ((7, 5), (8, 6), (14, 14), (9, 9)):
        x = rng.normal(size=(2, h, w, 3)).astype(np.float32)
        out, slot = O.maxpool3s2_same(x)
        oh, ow = (h + 1) // 2, (w + 1) // 2
        th, tw = max((oh - 1) * 2 + 3 - h, 0), max((ow - 1) * 2 + 3 - w, 0)
        xt = torch.tensor(x).permute(0, 3, 1, 2).requires_grad_(True)
        ref = torch.nn.functional.max_pool2d(
            torch.nn.functional.pad(xt, (tw // 2, tw - tw // 2, th // 2, th - th // 2), value=-np.inf), 3, 2)
        assert np.array_equal(out, ref.detach().permute(0, 2, 3, 1).numpy())
        g = rng.normal(size=out.shape).astype(np.float32)
        ref.backward(torch.tensor(g).permute(0, 3, 1, 2))
        # distinct float values: no ties, so the first-max rule and torch's agree
        np.testing.assert_allclose(O.maxpool3s2_same_backward(g.astype(np.float64), slot, (h, w)),
                                   xt.grad.permute(0, 2, 3, 1).numpy(), rtol=1e-6, atol=1e-7)
    # ties: the first maximum in row-major window order wins
    x = np.zeros((1, 4, 4, 1), dtype=np.float32)
    _, slot = O.maxpool3s2_same(x)
    assert slot[0, 0, 0, 0] == 0 and slot[0, 1, 1, 0] == 0


def test_step_shapes_gradients_and_clip():
    plan = ("plain", "down")
    p = O.init_params(plan, 8, 5, seed=0)
    assert p["kernel1"].shape == (3, 3, 8, 16) and p["sc_kernel1"].shape == (8, 16) and p["dense_kernel"].shape == (16, 5)
    rng = np.random.default_rng(2)
    h0 = rng.normal(0, 0.8, (3, 6, 6, 8)).astype(np.float32)
    lab = rng.integers(0, 5, (3,))
    r = O.forward_backward(p, plan, h0, lab, keep=True)
    assert r["acts"]["h2"].shape == (3, 3, 3, 16) and set(r["grads"]) == set(p) | {"h0"}
    assert all(np.isfinite(v).all() and np.abs(v).max() > 0 for v in r["grads"].values())
    # finite-difference check of one float parameter (the quantizers are piecewise constant, gamma is smooth)
    eps, k = 1e-6, "gamma1"
    p2 = {n: v.astype(np.float64).copy() for n, v in p.items()}
    p2[k][3] += eps
    lp = O.forward_backward(p2, plan, h0, lab)["loss"]
    p2[k][3] -= 2 * eps
    lm = O.forward_backward(p2, plan, h0, lab)["loss"]
    assert abs((lp - lm) / (2 * eps) - r["grads"][k][3]) < 1e-6
    big = {n: (v * 50 if n.startswith("kernel") else v) for n, v in r["grads"].items() if n in p}
    zero = {n: np.zeros_like(v) for n, v in p.items()}
    new, _, _ = O.adam_step(p, big, zero, zero, 1, lr=2.0)
    assert all(np.abs(new[n]).max() <= 1.0 for n in new if n.startswith("kernel"))      # latent binary kernels are clipped
    assert np.abs(new["dense_kernel"]).max() > 1.0                                      # float weights are not


def test_stem_conv_follows_tf_same_padding_and_feeds_the_units():
    """The stem restated two ways: torch conv2d over explicit TensorFlow 'same' padding (what the oracle
    runs) against a direct numpy loop over the definition out[oy,ox] = sum x[oy*s - pad + kh, ...];
    7x7/2 on even and odd sizes, 11x11/4; the pool behind it against the numpy pool; and the whole
    model's gradient dictionary holds the stem's parameters."""
    rng = np.random.default_rng(0)
    for k, s, (h, w) in ((7, 2, (12, 9)), (7, 2, (16, 16)), (11, 4, (21, 18))):
        x = rng.uniform(-1, 1, (2, h, w, 3))
        wk = rng.normal(size=(k, k, 3, 4))
        oh, ow = -(-h // s), -(-w // s)
        ph, pw = max((oh - 1) * s + k - h, 0) // 2, max((ow - 1) * s + k - w, 0) // 2
        ref = np.zeros((2, oh, ow, 4))
        for oy in range(oh):
            for ox in range(ow):
                for kh in range(k):
                    for kw in range(k):
                        iy, ix = oy * s - ph + kh, ox * s - pw + kw
                        if 0 <= iy < h and 0 <= ix < w:
                            ref[:, oy, ox] += x[:, iy, ix] @ wk[kh, kw]
        got = torch.nn.functional.conv2d(O._same_pad(torch.tensor(x).permute(0, 3, 1, 2), k, s),
                                         torch.tensor(wk).permute(3, 2, 0, 1), stride=s).permute(0, 2, 3, 1).numpy()
        np.testing.assert_allclose(got, ref, rtol=1e-12, atol=1e-12)
    plan = ("plain", "down")
    p = O.init_params(plan, 64, 10, seed=1, stem_ksize=7)
    img = rng.uniform(-1, 1, (2, 32, 32, 3)).astype(np.float32)
    r = O.forward_backward(p, plan, img, rng.integers(0, 10, (2,)), torch.float64, keep=True)
    out, _ = O.maxpool3s2_same(r["acts"]["stem_a"])
    assert np.array_equal(out, r["acts"]["h0"]) and r["acts"]["h0"].shape == (2, 8, 8, 64)
    assert r["grads"]["stem_kernel"].shape == (7, 7, 3, 64) and np.abs(r["grads"]["stem_kernel"]).max() > 0
    assert O.bireal18_plan().count("down") == 3 and len(O.bireal18_plan()) == 16
