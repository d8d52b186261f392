"""CPU checks of oracle/sampled.py -- the per-entry float64 recomputation the batch-4096 GPU parity
test relies on -- against torch autograd of the same operators (the way oracle/binarynet_oracle.py
computes whole steps): forward, dgrad and wgrad entries of a 'same' 3x3 binary conv, the dense
forms, BatchNorm backward with the ste mask, and the whole first-layer backward.  No GPU."""
import numpy as np
import torch

from oracle import sampled as S


def _patches(m, pix, flip):
    """numpy twin of the GPU gather in tests/test_gpu_full_size_parity.py."""
    B, H, W, C = m.shape
    b, r = pix // (H * W), pix % (H * W)
    y, x = r // W, r % W
    out = np.zeros((len(pix), 3, 3, C), dtype=m.dtype)
    for kh in range(3):
        for kw in range(3):
            dy_, dx_ = (1 - kh, 1 - kw) if flip else (kh - 1, kw - 1)
            yy, xx = y + dy_, x + dx_
            ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
            v = m[b, np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)]
            out[:, kh, kw] = np.where(ok[:, None], v, 0)
    return out


def _shifted_columns(m, ch):
    sub = m[..., ch]
    B, H, W, n = sub.shape
    pad = np.pad(sub, ((0, 0), (1, 1), (1, 1), (0, 0)))
    return np.stack([pad[:, kh:kh + H, kw:kw + W].reshape(-1, n) for kh in range(3) for kw in range(3)])


def test_conv_entries_match_autograd():
    rng = np.random.default_rng(0)
    B, H, W, Ci, Co = 3, 7, 5, 6, 4
    s = rng.choice([-1, 1], size=(B, H, W, Ci)).astype(np.int8)
    w = rng.normal(size=(3, 3, Ci, Co)).astype(np.float32)
    dy = rng.normal(size=(B, H, W, Co))
    st = torch.tensor(s, dtype=torch.float64).permute(0, 3, 1, 2).requires_grad_(True)
    wq = torch.tensor(np.where(w >= 0, 1.0, -1.0)).permute(3, 2, 0, 1).requires_grad_(True)
    y = torch.nn.functional.conv2d(st, wq, padding=1)
    y.backward(torch.tensor(dy).permute(0, 3, 1, 2))
    y_ref = y.detach().permute(0, 2, 3, 1).numpy().reshape(-1, Co)
    ds_ref = st.grad.permute(0, 2, 3, 1).numpy().reshape(-1, Ci)
    dw_ref = wq.grad.permute(2, 3, 1, 0).numpy().reshape(9, Ci, Co)
    pix = rng.permutation(B * H * W)[:40]
    assert np.array_equal(S.conv_forward_at(_patches(s, pix, False), w), np.rint(y_ref[pix]).astype(np.int64))
    np.testing.assert_allclose(S.conv_dgrad_at(_patches(dy, pix, True), w), ds_ref[pix], rtol=1e-12, atol=1e-12)
    ci, co = np.array([0, 2, 5]), np.array([1, 3])
    got = S.conv_wgrad_entries(_shifted_columns(s, ci), dy.reshape(-1, Co)[:, co])
    np.testing.assert_allclose(got, dw_ref[:, ci][:, :, co], rtol=1e-12, atol=1e-12)


def test_dense_entries_match_matmul():
    rng = np.random.default_rng(1)
    s = rng.choice([-1, 1], size=(9, 12)).astype(np.int8)
    w = rng.normal(size=(12, 5)).astype(np.float32)
    dy = rng.normal(size=(9, 5))
    q = np.where(w >= 0, 1.0, -1.0)
    assert np.array_equal(S.dense_forward(s, w), (s.astype(np.float64) @ q).astype(np.int64))
    np.testing.assert_allclose(S.dense_dgrad(dy, w), dy @ q.T)
    np.testing.assert_allclose(S.dense_wgrad_entries(s[:, [1, 4]], dy[:, [0, 2]]), s[:, [1, 4]].T.astype(np.float64) @ dy[:, [0, 2]])


def test_bn_backward_columns_match_autograd_and_flag_border_elements():
    rng = np.random.default_rng(2)
    M, C = 200, 3
    x = rng.integers(-30, 31, (M, C)).astype(np.float64)
    ds = rng.normal(size=(M, C))
    beta = rng.normal(0, 0.3, C)
    mean, var = x.mean(0), x.var(0)
    rstd = 1.0 / np.sqrt(var + 1e-3)
    xt = torch.tensor(x, requires_grad=True)
    xh = (xt - xt.mean(0)) * torch.rsqrt(xt.var(0, unbiased=False) + 1e-3)
    a = xh + torch.tensor(beta)
    mask = (a.detach().abs() <= 1).double()          # ste_sign backward: dy * [|a| <= 1]
    (a * mask * torch.tensor(ds)).sum().backward()   # d/da of sign(a) under the STE, contracted with ds
    r = S.bn_backward_columns(ds, x, mean, rstd, beta, ste=True)
    np.testing.assert_allclose(r["dy"], xt.grad.numpy(), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(r["sum_da"], (ds * mask.numpy()).sum(0), rtol=1e-12)
    assert not r["border"].any() and np.all(r["slack_da"] == 0)
    # an element placed exactly on the clip boundary is flagged and its gradient becomes the slack
    x2 = x.copy()
    x2[0, 0] = mean[0] + (1.0 - beta[0]) / rstd[0]
    r2 = S.bn_backward_columns(ds, x2, mean, rstd, beta, ste=True)
    assert r2["border"][0, 0] and r2["border"].sum() == 1
    assert np.isclose(r2["slack_da"][0], abs(ds[0, 0]))
    assert S.excess(np.array([0.5, 0.0]), np.array([0.4, 0.0])) == np.float64(0.5 - 0.4)


def test_first_layer_backward_columns_match_autograd():
    rng = np.random.default_rng(3)
    B, H, W, C = 2, 6, 6, 5
    img = rng.uniform(-1, 1, (B, H, W)).astype(np.float64)
    w = np.where(rng.normal(size=(9, C)) >= 0, 1.0, -1.0)
    ds0 = rng.normal(size=(B * 16, C))
    beta = rng.normal(0, 0.3, C)
    taps = np.stack([img[:, kh:kh + 4, kw:kw + 4].reshape(-1) for kh in range(3) for kw in range(3)], axis=1)
    wt = torch.tensor(w, requires_grad=True)
    bt = torch.tensor(beta, requires_grad=True)
    y = torch.tensor(taps) @ wt
    xh = (y - y.mean(0)) * torch.rsqrt(y.var(0, unbiased=False) + 1e-3)
    a = xh + bt
    mask = (a.detach().abs() <= 1).double()
    (a * mask * torch.tensor(ds0)).sum().backward()
    r = S.first_layer_backward_columns(taps, w, ds0, beta)
    np.testing.assert_allclose(r["dw"], wt.grad.numpy(), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(r["dbeta"], bt.grad.numpy(), rtol=1e-9, atol=1e-12)
    assert np.all(r["slack_dw"] == 0)
