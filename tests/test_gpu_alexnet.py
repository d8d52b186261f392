"""BinaryAlexNet (BASELINE.json configs[3]): the overlapping 3x3/2 pool on
integer and float maps, the per-channel BatchNorm kernels on channel counts that do not divide a
256-thread block (192, 384) and on maps wider than one block (4096), and one full train step of the
model against oracle/alexnet_oracle.py.  Bit-exact: pooled values, pool slots, integer conv / dense
outputs.  Tolerance: loss, every gradient tensor and the updated weights within 1e-4 relative."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


@pytest.mark.parametrize("dtype", ["i16", "f32"])
@pytest.mark.parametrize("B,H,W,C,same", [(3, 27, 27, 192, 0), (2, 13, 14, 256, 0), (2, 56, 56, 64, 0), (2, 3, 3, 4, 0),
                                          (2, 14, 9, 64, 1)])
def test_maxpool_3x3_stride2(native_lib, dtype, B, H, W, C, same):
    from oracle import alexnet_oracle as A
    from oracle import bireal_oracle as R
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(H * 31 + W + C)
    if dtype == "i16":
        xn = rng.integers(-6, 7, (B, H, W, C)).astype(np.int16)        # few distinct values: many ties
        x = torch.as_tensor(xn).cuda()
    else:
        xn = rng.normal(size=(B, H, W, C)).astype(np.float32)
        x = torch.as_tensor(xn).cuda()
    OH, OW = ((H + 1) // 2, (W + 1) // 2) if same else ((H - 3) // 2 + 1, (W - 3) // 2 + 1)
    out = torch.zeros(B, OH, OW, C, dtype=x.dtype, device="cuda")
    arg = torch.full((B, OH, OW, C), 77, dtype=torch.uint8, device="cuda")
    zn.call("zkb_maxpool3s2_fwd", zn.ptr(x), int(dtype == "i16"), B, H, W, C, same, zn.ptr(out), zn.ptr(arg), 0)
    torch.cuda.synchronize()
    xt = torch.tensor(xn.astype(np.float64), requires_grad=True)
    if same:
        ref, slot = R.maxpool3s2_same(xn.astype(np.float64))
        assert np.array_equal(out.cpu().numpy().astype(np.float64), ref)
        assert np.array_equal(arg.cpu().numpy(), slot)
    else:
        ref = A.maxpool3s2_valid(xt)
        assert np.array_equal(out.cpu().numpy().astype(np.float64), ref.detach().numpy())
    dout = rng.normal(size=(B, OH, OW, C)).astype(np.float32)
    d = torch.as_tensor(dout).cuda()
    dx = torch.full((B, H, W, C), 5.0, device="cuda")
    zn.call("zkb_maxpool3s2_bwd", zn.ptr(d), zn.ptr(arg), B, H, W, C, same, zn.ptr(dx), 0, 0, 0)
    hi = torch.zeros(B, H, W, C, dtype=torch.bfloat16, device="cuda")
    lo = torch.zeros(B, H, W, C, dtype=torch.bfloat16, device="cuda")
    zn.call("zkb_maxpool3s2_bwd", zn.ptr(d), zn.ptr(arg), B, H, W, C, same, 0, zn.ptr(hi), zn.ptr(lo), 0)
    torch.cuda.synchronize()
    if same:
        ref_dx = R.maxpool3s2_same_backward(dout.astype(np.float64), slot, (H, W))
    else:
        ref.backward(torch.tensor(dout.astype(np.float64)))
        ref_dx = xt.grad.numpy()
    assert _rel(dx.cpu().numpy(), ref_dx) < 1e-6
    assert _rel((hi.float() + lo.float()).cpu().numpy(), ref_dx) < 2e-5          # bf16 pair: ~16 mantissa bits
    with pytest.raises(RuntimeError, match="dx_f32 or"):
        zn.call("zkb_maxpool3s2_bwd", zn.ptr(d), zn.ptr(arg), B, H, W, C, same, zn.ptr(dx), zn.ptr(hi), zn.ptr(lo), 0)


@pytest.mark.parametrize("rows,C", [(700, 192), (333, 384), (64, 4096), (50, 6144), (37, 1000 + 24)])
def test_batchnorm_kernels_on_odd_and_wide_channel_counts(native_lib, rows, C):
    """stats / apply+sign / backward reduce / backward apply against float64 torch, int16 and f32 input."""
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(C)
    for is_i16 in (1, 0):
        xn = rng.integers(-300, 301, (rows, C)).astype(np.int16) if is_i16 else rng.normal(0, 30, (rows, C)).astype(np.float32)
        x = torch.as_tensor(xn).cuda()
        acc = torch.zeros(4, C, dtype=torch.float64, device="cuda")
        zn.call("zkb_bn_stats", zn.ptr(x), is_i16, rows, C, zn.ptr(acc[0]), zn.ptr(acc[1]), 0)
        mean, rstd = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
        zn.call("zkb_bn_finalize", zn.ptr(acc[0]), zn.ptr(acc[1]), float(rows), 1e-3, 0.9, 0, C, zn.ptr(mean), zn.ptr(rstd), 0, 0, 0)
        beta = torch.as_tensor(rng.normal(0, 0.3, C).astype(np.float32)).cuda()
        s = torch.zeros(rows, C, dtype=torch.bfloat16, device="cuda")
        s8 = torch.zeros(rows, C, dtype=torch.int8, device="cuda")
        a = torch.zeros(rows, C, device="cuda")
        zn.call("zkb_bn_apply_sign", zn.ptr(x), is_i16, rows, C, zn.ptr(mean), zn.ptr(rstd), zn.ptr(beta), zn.ptr(s), zn.ptr(a), zn.ptr(s8), 0)
        torch.cuda.synchronize()
        x64 = torch.tensor(xn.astype(np.float64), requires_grad=True)
        xh = (x64 - x64.mean(0)) * torch.rsqrt(x64.var(0, unbiased=False) + 1e-3)
        a64 = xh + beta.double().cpu()
        assert _rel(acc[0].cpu().numpy(), xn.astype(np.float64).sum(0)) < 1e-6
        assert _rel(a.cpu().numpy(), a64.detach().numpy()) < 1e-5
        far = a64.detach().abs() > 1e-4            # away from the sign / clip boundaries
        assert torch.equal((s.float().cpu() > 0)[far], (a64.detach() >= 0)[far])
        assert torch.equal(s8.float().cpu(), s.float().cpu())
        # backward with the ste mask
        # (no gradient arrives at elements within 1e-3 of the clip boundary |a| = 1: which side of it the
        # float32 kernels see is then immaterial to every output)
        gn = rng.normal(size=(rows, C)).astype(np.float32)
        gn[np.abs(np.abs(a64.detach().numpy()) - 1.0) < 1e-3] = 0.0
        g = torch.as_tensor(gn).cuda()
        mda, mdx = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
        dbeta = torch.zeros(C, device="cuda")
        zn.call("zkb_bn_bwd_reduce", zn.ptr(g), zn.ptr(x), is_i16, rows, C, zn.ptr(mean), zn.ptr(rstd), zn.ptr(beta), 1,
                zn.ptr(acc[2]), zn.ptr(acc[3]), 0)
        zn.call("zkb_bn_bwd_finalize", zn.ptr(acc[2]), zn.ptr(acc[3]), float(rows), 1.0, C, zn.ptr(dbeta), zn.ptr(mda), zn.ptr(mdx), 0)
        dxf = torch.zeros(rows, C, device="cuda")
        hi = torch.zeros(rows, C, dtype=torch.bfloat16, device="cuda")
        lo = torch.zeros(rows, C, dtype=torch.bfloat16, device="cuda")
        args = (zn.ptr(g), zn.ptr(x), is_i16, rows, C, zn.ptr(mean), zn.ptr(rstd), zn.ptr(beta), zn.ptr(mda), zn.ptr(mdx), 1)
        zn.call("zkb_bn_bwd_apply", *args, 0, zn.ptr(dxf), 0, 0, 0, 0, 0, 0)
        zn.call("zkb_bn_bwd_apply", *args, 1, 0, zn.ptr(hi), zn.ptr(lo), 0, 0, 0, 0)
        torch.cuda.synchronize()
        mask = (a64.detach().abs() <= 1).double()
        (a64 * mask * g.double().cpu()).sum().backward()
        assert _rel(dbeta.cpu().numpy(), (g.double().cpu() * mask).sum(0).numpy()) < 1e-5
        assert _rel(dxf.cpu().numpy(), x64.grad.numpy()) < 1e-4
        assert _rel((hi.float() + lo.float()).cpu().numpy(), dxf.cpu().numpy()) < 2e-5


@pytest.mark.parametrize("hw,B,NC,dense,seed", [((96, 96), 4, 10, 256, 1), ((100, 80), 3, 1000, 384, 2)])
def test_binary_alexnet_train_step_matches_oracle(native_lib, hw, B, NC, dense, seed):
    from oracle import alexnet_oracle as A
    from zookeeper_b200.alexnet import BinaryAlexNetEngine

    p = A.init_params(NC, hw, dense_units=dense, seed=seed)
    rng = np.random.default_rng(seed + 50)
    img = rng.uniform(-1, 1, (B,) + hw + (3,)).astype(np.float32)
    lab = rng.integers(0, NC, (B,)).astype(np.int64)
    ref = A.forward_backward(p, img, lab, torch.float64, keep=True)

    eng = BinaryAlexNetEngine(hw, B, NC, dense_units=dense)
    eng.load_params(p)
    eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), apply_update=False)
    torch.cuda.synchronize()
    assert _rel(eng.L[0].y.cpu().numpy(), ref["acts"]["y0"]) < 1e-6, "stem output"
    for L in eng.L[1:]:
        got = L.y.cpu().numpy().astype(np.int64)
        got = got[:, :L.cout] if L.kind == "dense" else got
        assert np.array_equal(got, np.rint(ref["acts"][f"y{L.idx}"]).astype(np.int64)), f"layer {L.idx}: integers"
        if L.kind == "conv" and L.pool:
            assert np.array_equal(L.p.cpu().numpy().astype(np.int64), np.rint(ref["acts"][f"p{L.idx}"]).astype(np.int64)), f"layer {L.idx}: pooled"
    assert abs(eng.last_loss() - ref["loss"]) <= REL_TOL * abs(ref["loss"])
    assert _rel(eng.logits(), ref["logits"]) < REL_TOL
    assert eng.last_accuracy() == pytest.approx(ref["accuracy"])
    g = eng.get_grads()
    bad = {k: f"{_rel(g[k], v):.3g}" for k, v in ref["grads"].items() if not _rel(g[k], v) < REL_TOL}
    assert not bad, f"gradients outside 1e-4: {bad}"
    eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), apply_update=True)
    torch.cuda.synchronize()
    g2 = eng.get_grads()
    zero = {k: np.zeros_like(v) for k, v in p.items()}
    exp32, _, _ = A.adam_clip_step(p, {k: g2[k] for k in p}, zero, zero, 1, dtype=np.float32)
    exp64, _, _ = A.adam_clip_step(p, {k: ref["grads"][k] for k in p}, zero, zero, 1)
    got = eng.get_params()
    for k in p:
        assert _rel(got[k], exp32[k]) < 2e-6, f"adam kernel {k}"
        if k.startswith("kernel"):
            assert np.abs(got[k]).max() <= 1.0


def test_binary_alexnet_learns(native_lib):
    from oracle import alexnet_oracle as A
    from zookeeper_b200.alexnet import BinaryAlexNetEngine

    hw, B, NC = (96, 96), 32, 10
    rng = np.random.default_rng(31)
    protos = rng.uniform(-1, 1, (NC,) + hw + (3,)).astype(np.float32)
    eng = BinaryAlexNetEngine(hw, B, NC, dense_units=256)
    eng.load_params(A.init_params(NC, hw, dense_units=256, seed=7))
    first, last = [], []
    for s in range(100):
        lab = rng.integers(0, NC, (B,))
        img = np.clip(protos[lab] * 0.6 + rng.normal(0, 0.3, (B,) + hw + (3,)), -1, 1).astype(np.float32)
        eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda())
        (first if s < 10 else last).append(eng.last_loss())
    assert np.isfinite(last).all()
    assert np.mean(last[-10:]) < 0.6 * np.mean(first), (np.mean(first), np.mean(last[-10:]))


@pytest.mark.parametrize("B,NC,ld", [(300, 1000, 1024), (33, 65, 128), (256, 10, 64), (5, 1000, 1000)])
def test_softmax_cross_entropy_narrow_and_wide_heads(native_lib, B, NC, ld):
    """Thread-per-row (<= 64 classes) and warp-per-row forms: loss sum, accuracy with the first-maximum
    argmax on tied rows, gradient incl. zeros in the padding columns."""
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(B + NC)
    z = rng.normal(0, 2, (B, ld)).astype(np.float32)
    lab = rng.integers(0, NC, (B,)).astype(np.int32)
    z[0, :NC] = 0.5                                   # all tied: argmax is class 0
    lab[0] = 0
    z[1, :NC] = -1.0
    z[1, NC - 1] = z[1, NC // 2] = 3.0                # two maxima: the first one counts
    lab[1] = NC - 1
    zt, lt = torch.as_tensor(z).cuda(), torch.as_tensor(lab).cuda()
    d = torch.full((B, ld), 9.0, device="cuda")
    loss = torch.zeros(1, dtype=torch.float64, device="cuda")
    correct = torch.zeros(1, dtype=torch.int32, device="cuda")
    zn.call("zkb_softmax_ce", zn.ptr(zt), ld, zn.ptr(lt), B, NC, 1.0 / B, zn.ptr(d), zn.ptr(loss), zn.ptr(correct), 0)
    torch.cuda.synchronize()
    z64 = torch.tensor(z[:, :NC].astype(np.float64), requires_grad=True)
    ref = torch.nn.functional.cross_entropy(z64, torch.as_tensor(lab.astype(np.int64)), reduction="sum")
    (ref / B).backward()
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    first_max = np.argmax(z[:, :NC], axis=1)          # numpy returns the first maximum
    assert correct.item() == int((first_max == lab).sum())
    assert _rel(d[:, :NC].cpu().numpy(), z64.grad.numpy()) < 1e-5
    assert bool((d[:, NC:] == 0).all())
