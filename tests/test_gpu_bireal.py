"""Bi-Real-Net style binary residual units: kernel-level parity of the new
operators and one full train step of residual units + head against oracle/bireal_oracle.py.
Bit-exact: the integer binary conv output (the +-1 part of magnitude_aware_sign), pool slots.
Tolerance: loss, every gradient tensor (incl. the gradient flowing out over the float shortcut) and
the updated weights within 1e-4 relative (Frobenius norm per tensor), as for BinaryNet."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def test_quantizer_scale_backward_and_head_kernels(native_lib):
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(0)
    # magnitude_aware_sign scale
    w = rng.normal(0, 0.3, (3, 3, 128, 256)).astype(np.float32)
    dw = torch.as_tensor(w).cuda()
    alpha = torch.zeros(256, device="cuda")
    zn.call("zkb_weight_abs_mean", zn.ptr(dw), 9 * 128, 256, zn.ptr(alpha), 0)
    assert _rel(alpha.cpu().numpy(), np.abs(w.astype(np.float64)).mean(axis=(0, 1, 2))) < 1e-6
    # approx_sign backward + shortcut (values at and around the clip boundary included)
    h = rng.normal(0, 0.8, (4, 5, 5, 128)).astype(np.float32)
    h.reshape(-1)[:6] = [1.0, -1.0, 0.0, -0.0, 1.0000001, 0.5]
    ds = rng.normal(size=h.shape).astype(np.float32)
    dsh = rng.normal(size=h.shape).astype(np.float32)
    out = torch.zeros(h.shape, device="cuda")
    d_ds, d_h, d_dsh = (torch.as_tensor(t).cuda() for t in (ds, h, dsh))     # (keep the device copies alive)
    zn.call("zkb_approx_sign_bwd_add", zn.ptr(d_ds), zn.ptr(d_h), zn.ptr(d_dsh), zn.ptr(out), h.size, 0)
    torch.cuda.synchronize()
    a = np.abs(h.astype(np.float64))
    ref = dsh + ds * (2 - 2 * a) * (a <= 1)
    assert _rel(out.cpu().numpy(), ref) < 1e-6
    # global average pool and its backward
    x = torch.as_tensor(rng.normal(size=(6, 7 * 7, 256)).astype(np.float32)).cuda()
    pooled = torch.zeros(6, 256, device="cuda")
    zn.call("zkb_global_avgpool_fwd", zn.ptr(x), 6, 49, 256, zn.ptr(pooled), 0)
    assert _rel(pooled.cpu().numpy(), x.double().mean(1).cpu().numpy()) < 1e-6
    dp = torch.as_tensor(rng.normal(size=(6, 256)).astype(np.float32)).cuda()
    dx = torch.zeros_like(x)
    zn.call("zkb_global_avgpool_bwd", zn.ptr(dp), 6, 49, 256, zn.ptr(dx), 0)
    assert _rel(dx.cpu().numpy(), (dp.double() / 49)[:, None, :].expand(6, 49, 256).cpu().numpy()) < 1e-6
    # float classifier
    M, K, N = 9, 256, 1000
    xs = torch.as_tensor(rng.normal(size=(M, K)).astype(np.float32)).cuda()
    wk = torch.as_tensor(rng.normal(0, 0.05, (K, N)).astype(np.float32)).cuda()
    bk = torch.as_tensor(rng.normal(size=(N,)).astype(np.float32)).cuda()
    y = torch.zeros(M, N, device="cuda")
    zn.call("zkb_dense_f32_fwd", zn.ptr(xs), zn.ptr(wk), zn.ptr(bk), zn.ptr(y), M, K, N, N, 0)
    assert _rel(y.cpu().numpy(), (xs.double() @ wk.double() + bk.double()).cpu().numpy()) < 1e-6
    dy = torch.as_tensor(rng.normal(size=(M, N)).astype(np.float32)).cuda()
    dxs, dwk, dbk = torch.zeros_like(xs), torch.zeros_like(wk), torch.zeros_like(bk)
    zn.call("zkb_dense_f32_bwd", zn.ptr(xs), zn.ptr(wk), zn.ptr(dy), N, zn.ptr(dxs), zn.ptr(dwk), zn.ptr(dbk), M, K, N, 0)
    torch.cuda.synchronize()
    assert _rel(dxs.cpu().numpy(), (dy.double() @ wk.double().T).cpu().numpy()) < 1e-6
    assert _rel(dwk.cpu().numpy(), (xs.double().T @ dy.double()).cpu().numpy()) < 1e-6
    assert _rel(dbk.cpu().numpy(), dy.double().sum(0).cpu().numpy()) < 1e-6


@pytest.mark.parametrize("M,K,N,ld", [(5000, 64, 128, 128), (777, 70, 10, 12), (130, 512, 1000, 1000), (1, 3, 5, 5),
                                      (20000, 128, 256, 256)])
def test_float_gemm_all_directions(native_lib, M, K, N, ld):
    """zkb_dense_f32_fwd / _bwd (classifier and 1x1 shortcut conv): ragged and tile-aligned shapes, a
    row pitch wider than N, outputs written (not accumulated) whatever they held before."""
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(M + K)
    x = torch.randn(M, K, generator=g).cuda()
    w = (torch.randn(K, N, generator=g) * 0.1).cuda()
    b = torch.randn(N, generator=g).cuda()
    y = torch.full((M, ld), 7.0, device="cuda")
    zn.call("zkb_dense_f32_fwd", zn.ptr(x), zn.ptr(w), zn.ptr(b), zn.ptr(y), M, K, N, ld, 0)
    torch.cuda.synchronize()
    ref = x.double() @ w.double() + b.double()
    assert _rel(y[:, :N].cpu().numpy(), ref.cpu().numpy()) < 1e-6
    assert bool((y[:, N:] == 7.0).all()), "columns past N were written"
    dy = torch.zeros(M, ld, device="cuda")
    dy[:, :N] = torch.randn(M, N, generator=g).cuda()
    dy[:, N:] = 1e6          # padding columns must not be read
    dx, dw, db = torch.full((M, K), 3.0, device="cuda"), torch.full((K, N), 3.0, device="cuda"), torch.full((N,), 3.0, device="cuda")
    zn.call("zkb_dense_f32_bwd", zn.ptr(x), zn.ptr(w), zn.ptr(dy), ld, zn.ptr(dx), zn.ptr(dw), zn.ptr(db), M, K, N, 0)
    torch.cuda.synchronize()
    d64 = dy[:, :N].double()
    assert _rel(dx.cpu().numpy(), (d64 @ w.double().T).cpu().numpy()) < 1e-6
    assert _rel(dw.cpu().numpy(), (x.double().T @ d64).cpu().numpy()) < 2e-6
    assert _rel(db.cpu().numpy(), d64.sum(0).cpu().numpy()) < 2e-6
    # optional outputs
    dw2 = torch.full((K, N), 3.0, device="cuda")
    zn.call("zkb_dense_f32_bwd", zn.ptr(x), zn.ptr(w), zn.ptr(dy), ld, 0, zn.ptr(dw2), 0, M, K, N, 0)
    torch.cuda.synchronize()
    assert _rel(dw2.cpu().numpy(), (x.double().T @ d64).cpu().numpy()) < 2e-6


@pytest.mark.parametrize("H,W", [(14, 14), (7, 5), (8, 6)])
def test_stem_maxpool_3x3_stride2_same(native_lib, H, W):
    from oracle import bireal_oracle as O
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(H * 31 + W)
    # small integers: ties in most windows, so the first-max rule is exercised
    x = rng.integers(-3, 4, (3, H, W, 64)).astype(np.float32)
    ref, slot = O.maxpool3s2_same(x)
    OH, OW = ref.shape[1:3]
    out = torch.zeros(3, OH, OW, 64, device="cuda")
    arg = torch.zeros(3, OH, OW, 64, dtype=torch.uint8, device="cuda")
    dx_ = torch.as_tensor(x).cuda()
    zn.call("zkb_maxpool3s2_f32_fwd", zn.ptr(dx_), 3, H, W, 64, zn.ptr(out), zn.ptr(arg), 0)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(arg.cpu().numpy(), slot)
    dout = rng.normal(size=ref.shape).astype(np.float32)
    dx = torch.zeros(3, H, W, 64, device="cuda")
    d_dout = torch.as_tensor(dout).cuda()
    zn.call("zkb_maxpool3s2_f32_bwd", zn.ptr(d_dout), zn.ptr(arg), 3, H, W, 64, zn.ptr(dx), 0)
    torch.cuda.synchronize()
    assert _rel(dx.cpu().numpy(), O.maxpool3s2_same_backward(dout.astype(np.float64), slot, (H, W))) < 1e-6


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(3, 14, 14, 128, 256), (2, 7, 9, 128, 128), (5, 28, 28, 128, 256), (2, 13, 8, 256, 512)])
def test_stride2_binary_conv_forward_wgrad_dgrad(native_lib, B, H, W, Cin, Cout):
    """Down-sampling binary conv (3x3, stride 2, TensorFlow 'same' padding: asymmetric on even sizes):
    forward integers exact, weight gradient and input gradient (stride-1 dgrad over the zero-inserted
    map) against the float64 strided convolution, even and odd sizes."""
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(H * 100 + W)
    x = (torch.randint(0, 2, (B, H, W, Cin), generator=g, dtype=torch.int8) * 2 - 1).cuda()
    w = (torch.randint(0, 2, (3, 3, Cin, Cout), generator=g, dtype=torch.int8) * 2 - 1).cuda()
    wf = torch.empty(Cout, 9 * Cin, dtype=torch.bfloat16, device="cuda")
    wd = torch.empty(Cin, 9 * Cout, dtype=torch.bfloat16, device="cuda")
    wf8 = torch.empty(Cout, 9 * Cin, dtype=torch.int8, device="cuda")
    w32 = w.float().contiguous()
    zn.call("zkb_prep_conv_weights", zn.ptr(w32), 3, Cin, Cout, zn.ptr(wf), zn.ptr(wd), zn.ptr(wf8), 0)
    OH, OW = (H + 1) // 2, (W + 1) // 2
    th, tw = max((OH - 1) * 2 + 3 - H, 0), max((OW - 1) * 2 + 3 - W, 0)
    xr = x.permute(0, 3, 1, 2).double().requires_grad_(True)
    wr = w.permute(3, 2, 0, 1).double().requires_grad_(True)
    xp = torch.nn.functional.pad(xr, (tw // 2, tw - tw // 2, th // 2, th - th // 2))
    ref = torch.nn.functional.conv2d(xp, wr, stride=2)
    assert ref.shape[2:] == (OH, OW)
    y = torch.full((B, OH, OW, Cout), -7, dtype=torch.int16, device="cuda")
    xc = x.contiguous()
    zn.call("zkb_bconv2d_fwd_i8_s2", zn.ptr(xc), zn.ptr(wf8), zn.ptr(y), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    assert torch.equal(y.double(), ref.detach().permute(0, 2, 3, 1)), "stride-2 forward integers"
    dy = torch.randn(B, OH, OW, Cout, generator=g).cuda()
    ref.backward(dy.permute(0, 3, 1, 2).double())
    hi = dy.to(torch.bfloat16).contiguous()
    lo = (dy - hi.float()).to(torch.bfloat16).contiguous()
    xb = x.to(torch.bfloat16).contiguous()
    dw = torch.zeros(3, 3, Cin, Cout, dtype=torch.float32, device="cuda")
    zn.call("zkb_bconv2d_wgrad_s2", zn.ptr(xb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), B, H, W, Cin, Cout, 3, 0)
    Dh = torch.full((B, H, W, Cout), 3.0, dtype=torch.bfloat16, device="cuda")
    Dl = torch.full((B, H, W, Cout), 3.0, dtype=torch.bfloat16, device="cuda")
    zn.call("zkb_dilate2_hi_lo", zn.ptr(hi), zn.ptr(lo), B, H, W, Cout, 3, zn.ptr(Dh), zn.ptr(Dl), 0)
    dx = torch.empty(B, H, W, Cin, dtype=torch.float32, device="cuda")
    zn.call("zkb_bconv2d_dgrad", zn.ptr(Dh), zn.ptr(Dl), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    assert int((Dh.float() != 0).sum()) <= B * OH * OW * Cout          # zeros everywhere else
    for got, rf, what in ((dw.double(), wr.grad.permute(2, 3, 1, 0), "wgrad"), (dx.double(), xr.grad.permute(0, 2, 3, 1), "dgrad")):
        rel = ((got - rf).norm() / rf.norm()).item()
        assert rel < 1e-5, f"stride-2 {what}: {rel:.3g}"


@pytest.mark.parametrize("units,C,hw,B,NC", [(2, 128, (14, 14), 8, 10), (1, 256, (7, 7), 5, 1000),
                                             (("plain", "down", "plain"), 128, (14, 14), 4, 10),
                                             (("down", "down"), 128, (12, 8), 3, 10),
                                             (("plain", "plain"), 64, (14, 14), 6, 10),           # Bi-Real-Net-18 stage 1
                                             (("plain", "down", "plain"), 64, (12, 12), 4, 10)])  # ... into stage 2
def test_residual_units_train_step_matches_oracle(native_lib, units, C, hw, B, NC):
    """a Bi-Real-Net-18 residual unit trains one step within 1e-4 of the oracle --
    plain units, and stacks with down-sampling units (stride-2 binary conv; average-pool + 1x1 float
    conv + BatchNorm shortcut)."""
    from oracle import bireal_oracle as O
    from zookeeper_b200.bireal import BiRealUnitsEngine

    p = O.init_params(units, C, NC, seed=3)
    rng = np.random.default_rng(7)
    h0 = rng.normal(0, 0.8, (B,) + hw + (C,)).astype(np.float32)
    lab = rng.integers(0, NC, (B,)).astype(np.int64)
    ref = O.forward_backward(p, units, h0, lab, torch.float64, keep=True)

    eng = BiRealUnitsEngine(units, C, hw, B, NC)
    eng.load_params(p)
    eng.train_step(torch.as_tensor(h0).cuda(), torch.as_tensor(lab).cuda(), apply_update=False)
    torch.cuda.synchronize()
    for u in range(eng.U):
        # the +-1 contraction is exact; the magnitude scale only enters through the BatchNorm gain
        assert np.array_equal(eng.units[u].y.cpu().numpy().astype(np.int64), ref["acts"][f"y{u}"]), f"unit {u}: integers"
        assert _rel(eng.units[u].alpha.cpu().numpy(), ref["acts"][f"alpha{u}"]) < 1e-6
        assert _rel(eng.h[u + 1].cpu().numpy(), ref["acts"][f"h{u + 1}"]) < 1e-5, f"unit {u}: output map"
    assert abs(eng.last_loss() - ref["loss"]) <= REL_TOL * abs(ref["loss"])
    assert _rel(eng.logits.cpu().numpy(), ref["logits"]) < REL_TOL
    assert eng.last_accuracy() == pytest.approx(ref["accuracy"])
    g = eng.get_grads()
    for k, v in ref["grads"].items():
        assert _rel(g[k], v) < REL_TOL, f"grad {k}: rel {_rel(g[k], v):.3g}"
    # optimizer: Adam everywhere, clip on the latent binary kernels only
    zero = {k: np.zeros_like(v) for k, v in p.items()}
    exp64, _, _ = O.adam_step(p, {k: ref["grads"][k] for k in p}, zero, zero, 1)
    exp32, _, _ = O.adam_step(p, {k: g[k] for k in p}, zero, zero, 1, dtype=np.float32)
    eng._stream = torch.cuda.current_stream().cuda_stream
    eng.optimizer_step()
    torch.cuda.synchronize()
    got = eng.get_params()
    for k in p:
        assert _rel(got[k], exp32[k]) < 2e-6, f"adam kernel {k}"
        if k.startswith("kernel"):
            assert _rel(got[k], exp64[k]) < REL_TOL and np.abs(got[k]).max() <= 1.0, k


def test_residual_units_learn(native_lib):
    """A few hundred steps on a learnable task: the loss falls and the batch accuracy rises (the
    quantizer gradients, the gain BatchNorm and the shortcut work together across steps)."""
    from oracle import bireal_oracle as O
    from zookeeper_b200.bireal import BiRealUnitsEngine

    units, C, hw, B, NC = 2, 128, (7, 7), 64, 10
    rng = np.random.default_rng(11)
    protos = rng.normal(0, 0.7, (NC,) + hw + (C,)).astype(np.float32)
    eng = BiRealUnitsEngine(units, C, hw, B, NC)
    eng.load_params(O.init_params(units, C, NC, seed=5))
    first, last = [], []
    for s in range(120):
        lab = rng.integers(0, NC, (B,))
        h0 = (protos[lab] * 0.5 + rng.normal(0, 0.6, (B,) + hw + (C,))).astype(np.float32)
        eng.train_step(torch.as_tensor(h0).cuda(), torch.as_tensor(lab).cuda())
        (first if s < 10 else last).append((eng.last_loss(), eng.last_accuracy()))
    assert np.mean([a for _, a in last[-10:]]) > 0.9 > np.mean([a for _, a in first])
    assert np.mean([l for l, _ in last[-10:]]) < 0.5 * np.mean([l for l, _ in first])


@pytest.mark.parametrize("plan,hw,B,NC,seed", [(("plain", "down", "plain"), (32, 32), 4, 10, 4),
                                               ("bireal18", (64, 64), 2, 10, 4),
                                               (("plain",), (30, 22), 3, 1000, 5)])
def test_whole_net_with_stem_train_step_matches_oracle(native_lib, plan, hw, B, NC, seed):
    """Stem (float 7x7/2 conv -> BatchNorm -> 3x3/2 max-pool) + residual units + head: one train step
    of the whole model -- Bi-Real-Net-18's own unit plan in one case -- within 1e-4 of the oracle."""
    from oracle import bireal_oracle as O
    from zookeeper_b200.bireal import BiRealNetEngine, bireal18_plan

    if plan == "bireal18":
        plan = O.bireal18_plan()
        assert plan == bireal18_plan()
    p = O.init_params(plan, 64, NC, seed=seed, stem_ksize=7)
    rng = np.random.default_rng(seed + 100)
    img = rng.uniform(-1, 1, (B,) + hw + (3,)).astype(np.float32)
    lab = rng.integers(0, NC, (B,)).astype(np.int64)
    ref = O.forward_backward(p, plan, img, lab, torch.float64, keep=True)

    eng = BiRealNetEngine(plan, hw, B, NC)
    eng.load_params(p)
    eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), apply_update=False)
    torch.cuda.synchronize()
    assert _rel(eng.z.cpu().numpy(), ref["acts"]["stem_z"]) < 1e-6, "stem conv output"
    assert _rel(eng.a.cpu().numpy(), ref["acts"]["stem_a"]) < 1e-5, "stem BatchNorm output"
    assert _rel(eng.body.h[0].cpu().numpy(), ref["acts"]["h0"]) < 1e-5, "stem pool output"
    for u in range(eng.body.U):
        assert np.array_equal(eng.body.units[u].y.cpu().numpy().astype(np.int64), ref["acts"][f"y{u}"]), f"unit {u}: integers"
    assert abs(eng.last_loss() - ref["loss"]) <= REL_TOL * abs(ref["loss"])
    assert _rel(eng.body.logits.cpu().numpy(), ref["logits"]) < REL_TOL
    g = eng.get_grads()
    worst = {k: _rel(g[k], v) for k, v in ref["grads"].items() if k != "h0"}
    bad = {k: f"{r:.3g}" for k, r in worst.items() if not r < REL_TOL}
    assert not bad, f"gradients outside 1e-4: {bad}"
    # optimizer: the stem kernel is a float Conv2D -- Adam without the clip
    # (step 1 of Adam moves every weight by ~lr * sign(g): the expectation is built from the gradients of
    # the SAME pass -- the weight-gradient sums are atomic, so a gradient within rounding of zero may
    # change sign between two passes over the same batch)
    eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), apply_update=True)
    torch.cuda.synchronize()
    g2 = eng.get_grads()
    for k in p:
        assert _rel(g2[k], g[k]) < 1e-5, f"second pass over the same batch: grad {k}"
    zero = {k: np.zeros_like(v) for k, v in p.items()}
    exp32, _, _ = O.adam_step(p, {k: g2[k] for k in p}, zero, zero, 1, dtype=np.float32)
    got = eng.get_params()
    for k in p:
        assert _rel(got[k], exp32[k]) < 2e-6, f"adam kernel {k}"
    assert eng.launches_per_step > 0


def test_whole_net_learns(native_lib):
    """A hundred steps of the stem + units + head on a learnable task: the loss falls."""
    from oracle import bireal_oracle as O
    from zookeeper_b200.bireal import BiRealNetEngine

    plan, hw, B, NC = ("plain", "down", "plain"), (32, 32), 32, 10
    rng = np.random.default_rng(21)
    protos = rng.uniform(-1, 1, (NC,) + hw + (3,)).astype(np.float32)
    eng = BiRealNetEngine(plan, hw, B, NC)
    eng.load_params(O.init_params(plan, 64, NC, seed=6, stem_ksize=7))
    first, last = [], []
    for s in range(100):
        lab = rng.integers(0, NC, (B,))
        img = np.clip(protos[lab] * 0.6 + rng.normal(0, 0.3, (B,) + hw + (3,)), -1, 1).astype(np.float32)
        eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda())
        (first if s < 10 else last).append(eng.last_loss())
    assert np.isfinite(last).all()
    assert np.mean(last[-10:]) < 0.5 * np.mean(first), (np.mean(first), np.mean(last[-10:]))
