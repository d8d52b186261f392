"""SURVEY.md section 8f.2, the pieces Bi-Real-Net-18 / BinaryAlexNet need besides the residual units:
binary convs whose channel counts are multiples of 64 but not of 128 (the 64-channel first stage),
5x5 binary kernels, and the float 7x7/2 and 11x11/4 stem convolutions.  References: float64
torch convolutions on the same device with TensorFlow's 'same' padding written out explicitly.
Bit-exact: integer forward outputs.  Tolerance: gradients 1e-5 relative (Frobenius), stem forward 1e-6."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _same_pad(x_nchw, k, s, value=0.0):
    h, w = x_nchw.shape[2:]
    oh, ow = -(-h // s), -(-w // s)
    th, tw = max((oh - 1) * s + k - h, 0), max((ow - 1) * s + k - w, 0)
    return torch.nn.functional.pad(x_nchw, (tw // 2, tw - tw // 2, th // 2, th - th // 2), value=value), oh, ow


def _rel(got, ref):
    return ((got.double() - ref.double()).norm() / ref.double().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("B,H,W,Cin,Cout,ks,stride", [
    (4, 14, 14, 64, 64, 3, 1),       # Bi-Real-Net-18 stage 1
    (3, 9, 7, 64, 64, 3, 1),         # odd map
    (3, 14, 10, 64, 128, 3, 2),      # first down-sampling unit: 64 -> 128, stride 2
    (2, 9, 7, 64, 128, 3, 2),
    (2, 12, 12, 128, 64, 3, 1),      # narrow output
    (2, 12, 12, 128, 256, 5, 1),     # 5x5 kernel on the int8 path
    (2, 10, 9, 64, 128, 5, 1),       # 5x5 kernel, 64 input channels
    (2, 27, 27, 64, 192, 5, 1),      # BinaryAlexNet block 2 (output channels not a multiple of the MMA tile)
    (2, 13, 13, 192, 384, 3, 1),     # ... block 3
    (2, 13, 13, 384, 384, 3, 1),     # ... block 4
    (3, 13, 13, 384, 256, 3, 1),     # ... block 5
], ids=lambda v: str(v))
def test_binary_conv_narrow_and_5x5(native_lib, B, H, W, Cin, Cout, ks, stride):
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(1000 * Cin + 10 * H + ks)
    x = (torch.randint(0, 2, (B, H, W, Cin), generator=g, dtype=torch.int8) * 2 - 1).cuda().contiguous()
    w = (torch.randint(0, 2, (ks, ks, Cin, Cout), generator=g, dtype=torch.int8) * 2 - 1).cuda()
    K = ks * ks
    wf = torch.empty(Cout, K * Cin, dtype=torch.bfloat16, device="cuda")
    wd = torch.empty(Cin, K * Cout, dtype=torch.bfloat16, device="cuda")
    wf8 = torch.empty(Cout, K * Cin, dtype=torch.int8, device="cuda")
    w32 = w.float().contiguous()
    zn.call("zkb_prep_conv_weights", zn.ptr(w32), ks, Cin, Cout, zn.ptr(wf), zn.ptr(wd), zn.ptr(wf8), 0)

    xr = x.permute(0, 3, 1, 2).double().requires_grad_(True)
    wr = w.permute(3, 2, 0, 1).double().requires_grad_(True)
    xp, OH, OW = _same_pad(xr, ks, stride)
    ref = torch.nn.functional.conv2d(xp, wr, stride=stride)
    assert ref.shape[2:] == (OH, OW)

    xb = x.to(torch.bfloat16).contiguous()
    y = torch.full((B, OH, OW, Cout), -7, dtype=torch.int16, device="cuda")
    name = "zkb_bconv2d_fwd" if stride == 1 else "zkb_bconv2d_fwd_s2"
    zn.call(name, zn.ptr(xb), zn.ptr(wf), zn.ptr(y), B, H, W, Cin, Cout, ks, 0)
    torch.cuda.synchronize()
    ref_y = ref.detach().permute(0, 2, 3, 1)
    bad = (y.double() != ref_y)
    assert not bool(bad.any()), f"bf16 forward: {int(bad.sum())} of {bad.numel()} integers differ"
    if Cin % 128 == 0:
        y8 = torch.full((B, OH, OW, Cout), -7, dtype=torch.int16, device="cuda")
        name8 = "zkb_bconv2d_fwd_i8" if stride == 1 else "zkb_bconv2d_fwd_i8_s2"
        zn.call(name8, zn.ptr(x), zn.ptr(wf8), zn.ptr(y8), B, H, W, Cin, Cout, ks, 0)
        torch.cuda.synchronize()
        assert torch.equal(y8, y), "int8 forward"

    dy = torch.randn(B, OH, OW, Cout, generator=g).cuda()
    ref.backward(dy.permute(0, 3, 1, 2).double())
    hi = dy.to(torch.bfloat16).contiguous()
    lo = (dy - hi.float()).to(torch.bfloat16).contiguous()
    dw = torch.zeros(ks, ks, Cin, Cout, dtype=torch.float32, device="cuda")
    dx = torch.full((B, H, W, Cin), 5.0, dtype=torch.float32, device="cuda")
    if stride == 1:
        zn.call("zkb_bconv2d_wgrad", zn.ptr(xb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), B, H, W, Cin, Cout, ks, 0)
        zn.call("zkb_bconv2d_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, ks, 0)
    else:
        zn.call("zkb_bconv2d_wgrad_s2", zn.ptr(xb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), B, H, W, Cin, Cout, ks, 0)
        Dh = torch.empty(B, H, W, Cout, dtype=torch.bfloat16, device="cuda")
        Dl = torch.empty(B, H, W, Cout, dtype=torch.bfloat16, device="cuda")
        zn.call("zkb_dilate2_hi_lo", zn.ptr(hi), zn.ptr(lo), B, H, W, Cout, ks, zn.ptr(Dh), zn.ptr(Dl), 0)
        zn.call("zkb_bconv2d_dgrad", zn.ptr(Dh), zn.ptr(Dl), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, ks, 0)
    torch.cuda.synchronize()
    ref_dw = wr.grad.permute(2, 3, 1, 0)
    ref_dx = xr.grad.permute(0, 2, 3, 1)
    # per-tap report: a wrong tap / channel-half mapping shows up as whole blocks
    if _rel(dw, ref_dw) >= 1e-5:
        per_tap = [[round(_rel(dw[a, b2], ref_dw[a, b2]), 6) for b2 in range(ks)] for a in range(ks)]
        halves = [round(_rel(dw[:, :, c:c + 64], ref_dw[:, :, c:c + 64]), 6) for c in range(0, Cin, 64)]
        raise AssertionError(f"wgrad rel {_rel(dw, ref_dw):.3g}; per tap {per_tap}; per 64 input channels {halves}")
    assert _rel(dx, ref_dx) < 1e-5, f"dgrad rel {_rel(dx, ref_dx):.3g}"


@pytest.mark.parametrize("K", [64, 192, 320])
def test_dense_wgrad_odd_multiple_of_64(native_lib, K):
    """zkb_bdense_wgrad with a reduced operand whose width is an odd multiple of 64."""
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(K)
    rows, N = 300, 128
    s = (torch.randint(0, 2, (rows, K), generator=g, dtype=torch.int8) * 2 - 1).cuda()
    dy = torch.randn(rows, N, generator=g).cuda()
    hi = dy.to(torch.bfloat16).contiguous()
    lo = (dy - hi.float()).to(torch.bfloat16).contiguous()
    sb = s.to(torch.bfloat16).contiguous()
    dw = torch.zeros(K, N, dtype=torch.float32, device="cuda")
    zn.call("zkb_bdense_wgrad", zn.ptr(sb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), N, rows, K, N, N, 0)
    torch.cuda.synchronize()
    ref = s.double().t() @ dy.double()
    assert _rel(dw, ref) < 1e-5, f"dense wgrad K={K}: {_rel(dw, ref):.3g}"


@pytest.mark.parametrize("ks,stride,B,H,W,Cout", [
    (7, 2, 3, 32, 32, 64),
    (7, 2, 2, 30, 23, 128),      # odd sizes, ragged tiles, two channel groups
    (7, 2, 2, 224, 224, 64),     # ImageNet-shaped: every persistent block visits several tiles
    (11, 4, 2, 35, 40, 64),
    (11, 4, 1, 224, 224, 64),
], ids=lambda v: str(v))
def test_stem_conv_forward_and_wgrad(native_lib, ks, stride, B, H, W, Cout):
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(ks * 1000 + H)
    x = (torch.rand(B, H, W, 3, generator=g) * 2 - 1).cuda().contiguous()
    w = (torch.randn(ks, ks, 3, Cout, generator=g) * 0.1).cuda().contiguous()
    xr = x.permute(0, 3, 1, 2).double()
    wr = w.permute(3, 2, 0, 1).double().requires_grad_(True)
    xp, OH, OW = _same_pad(xr, ks, stride)
    ref = torch.nn.functional.conv2d(xp, wr, stride=stride)
    assert ref.shape[2:] == (OH, OW)
    y = torch.full((B, OH, OW, Cout), 9.0, dtype=torch.float32, device="cuda")
    zn.call("zkb_conv_stem_f32_fwd", zn.ptr(x), zn.ptr(w), zn.ptr(y), B, H, W, 3, Cout, ks, stride, 0)
    torch.cuda.synchronize()
    assert _rel(y, ref.detach().permute(0, 2, 3, 1)) < 1e-6, f"stem forward: {_rel(y, ref.detach().permute(0, 2, 3, 1)):.3g}"
    dy = torch.randn(B, OH, OW, Cout, generator=g).cuda().contiguous()
    ref.backward(dy.permute(0, 3, 1, 2).double())
    dw = torch.zeros(ks, ks, 3, Cout, dtype=torch.float32, device="cuda")
    zn.call("zkb_conv_stem_f32_wgrad", zn.ptr(x), zn.ptr(dy), zn.ptr(dw), B, H, W, 3, Cout, ks, stride, 0)
    torch.cuda.synchronize()
    ref_dw = wr.grad.permute(2, 3, 1, 0)
    assert _rel(dw, ref_dw) < 1e-5, f"stem wgrad: {_rel(dw, ref_dw):.3g}"
    # accumulates: a second call doubles it
    zn.call("zkb_conv_stem_f32_wgrad", zn.ptr(x), zn.ptr(dy), zn.ptr(dw), B, H, W, 3, Cout, ks, stride, 0)
    torch.cuda.synchronize()
    assert _rel(dw, 2 * ref_dw) < 1e-5


def test_stem_conv_rejects_unsupported_shapes(native_lib):
    from zookeeper_b200 import native as zn

    x = torch.zeros(1, 8, 8, 3, device="cuda")
    w = torch.zeros(3, 3, 3, 64, device="cuda")
    y = torch.zeros(1, 8, 8, 64, device="cuda")
    with pytest.raises(RuntimeError, match="7x7"):
        zn.call("zkb_conv_stem_f32_fwd", zn.ptr(x), zn.ptr(w), zn.ptr(y), 1, 8, 8, 3, 64, 3, 1, 0)
    with pytest.raises(RuntimeError, match="multiple of 64"):
        zn.call("zkb_conv_stem_f32_fwd", zn.ptr(x), zn.ptr(w), zn.ptr(y), 1, 8, 8, 3, 32, 7, 2, 0)
