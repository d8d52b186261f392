"""Kernel-level parity through the C ABI: preprocessing variants, sign + bit-pack, XNOR-popcount
GEMM, and the tcgen05 GEMM engines -- against the CPU oracle / golden vectors at small sizes and
through size-independent properties at BASELINE.json's full sizes (batch 4096)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _pm1(shape, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randint(0, 2, shape, generator=g, dtype=torch.int8) * 2 - 1).cuda()


def _split(x):
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    return hi.contiguous(), lo.contiguous()


def _preprocess(zn, img, out_hw, training, pad=0, aug=None, labels=None, binarize=False, onehot_classes=0):
    B, IH, IW, C = img.shape
    d_img = torch.as_tensor(img).cuda()
    out = torch.full((B, out_hw[0], out_hw[1], C), 7.0, dtype=torch.float32, device="cuda")
    d_aug = torch.as_tensor(aug).cuda() if aug is not None else None
    d_lab = torch.as_tensor(labels).cuda() if labels is not None else None
    lab_out = torch.zeros(B, dtype=torch.int32, device="cuda") if labels is not None else None
    oh = torch.zeros(B, onehot_classes, dtype=torch.float32, device="cuda") if onehot_classes else None
    zn.call("zkb_preprocess_u8", zn.ptr(d_img), zn.ptr(d_lab), 0, zn.ptr(d_aug), zn.ptr(out), zn.ptr(lab_out),
            zn.ptr(oh), B, IH, IW, C, out_hw[0], out_hw[1], pad, int(training), int(binarize), max(onehot_classes, 1), 0, 0, 0, 0)
    torch.cuda.synchronize()
    return out.cpu().numpy(), (lab_out.cpu().numpy() if lab_out is not None else None), (oh.cpu().numpy() if oh is not None else None)


def test_preprocess_golden_variants(native_lib):
    from oracle import binarynet_oracle as O
    from zookeeper_b200 import native as zn

    g = np.load(os.path.join(GOLD, "preprocess.npz"))
    img, aug = g["img"], g["aug"]
    labels = np.array([3, 0, 9], dtype=np.int64)
    out, lab, onehot = _preprocess(zn, img, (32, 32), True, 40, aug, labels, onehot_classes=10)
    assert np.array_equal(out, g["train"])                     # bit-exact (true fp32 division)
    assert np.array_equal(lab, labels.astype(np.int32))
    assert np.array_equal(onehot, np.eye(10, dtype=np.float32)[labels])   # optional one-hot output
    assert np.array_equal(_preprocess(zn, img, (28, 28), False)[0], g["eval_crop"])
    assert np.array_equal(_preprocess(zn, img, (36, 36), False)[0], g["eval_pad"])
    b = _preprocess(zn, img, (32, 32), True, 40, aug, binarize=True)[0]    # optional sign-binarise
    assert np.array_equal(b, O.preprocess(img, (32, 32), True, 40, aug, binarize=True))
    assert set(np.unique(b).tolist()) <= {-1.0, 1.0}
    # empty batch is a no-op
    zn.call("zkb_preprocess_u8", 16, 0, 0, 0, 16, 0, 0, 0, 32, 32, 3, 32, 32, 40, 1, 0, 1, 0, 0, 0, 0)


def test_preprocess_gathers_by_index_and_uses_counter_rng(native_lib):
    from oracle import binarynet_oracle as O
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(3)
    pool = rng.integers(0, 256, (50, 28, 28, 1), dtype=np.uint8)
    labels = rng.integers(0, 10, (50,)).astype(np.int64)
    index = rng.permutation(50)[:16].astype(np.int32)
    d_pool, d_lab, d_idx = torch.as_tensor(pool).cuda(), torch.as_tensor(labels).cuda(), torch.as_tensor(index).cuda()
    out = torch.empty(16, 28, 28, 1, dtype=torch.float32, device="cuda")
    lab = torch.empty(16, dtype=torch.int32, device="cuda")
    zn.call("zkb_preprocess_u8", zn.ptr(d_pool), zn.ptr(d_lab), zn.ptr(d_idx), 0, zn.ptr(out), zn.ptr(lab), 0,
            16, 28, 28, 1, 28, 28, 32, 1, 0, 10, 77, 5, 0, 0)
    torch.cuda.synchronize()
    aug = O.augmentation_params(77, 5, 16, 32, 28, 28)
    assert np.array_equal(out.cpu().numpy(), O.preprocess(pool[index], (28, 28), True, 32, aug))
    assert np.array_equal(lab.cpu().numpy(), labels[index].astype(np.int32))


def test_sign_pack_and_popcount_gemm_golden(native_lib):
    from zookeeper_b200 import native as zn

    g = np.load(os.path.join(GOLD, "bitops.npz"))
    a, b = torch.as_tensor(g["a"]).cuda(), torch.as_tensor(g["b"]).cuda()
    pa = torch.zeros(8, 4, dtype=torch.int32, device="cuda")
    pb = torch.zeros(6, 4, dtype=torch.int32, device="cuda")
    zn.call("zkb_sign_pack", zn.ptr(a), zn.ptr(pa), a.numel(), 0, 0)
    zn.call("zkb_sign_pack", zn.ptr(b), zn.ptr(pb), b.numel(), 0, 0)
    y = torch.zeros(8, 6, dtype=torch.int32, device="cuda")
    zn.call("zkb_bgemm_popc", zn.ptr(pa), zn.ptr(pb), zn.ptr(y), 6, 8, 6, 4, 0)
    torch.cuda.synchronize()
    assert np.array_equal(pa.cpu().numpy().view(np.uint32), g["pa"])     # includes sign(+-0) = +1
    assert np.array_equal(pb.cpu().numpy().view(np.uint32), g["pb"])
    assert np.array_equal(y.cpu().numpy(), g["y"])
    # bf16 +-1 input packs to the same words as its float source
    pa2 = torch.zeros_like(pa)
    s = torch.where(a >= 0, 1.0, -1.0).to(torch.bfloat16)
    zn.call("zkb_sign_pack", zn.ptr(s), zn.ptr(pa2), s.numel(), 1, 0)
    torch.cuda.synchronize()
    assert torch.equal(pa, pa2)


def test_three_formulations_of_the_binary_dense_forward_agree_at_full_size(native_lib):
    """Batch 4096 x K 4608 x N 1024 (dense1 of BinaryNet at configs[1]): tcgen05 (bf16 operands, f32
    accumulate), tcgen05 kind::i8 (int32 accumulate), CUDA-core XNOR-popcount over packed words, and a float matmul must give the same
    integers -- the exactness claim of DESIGN.md section 3.1."""
    from zookeeper_b200 import native as zn

    M, K, N = 4096, 4608, 1024
    a, w = _pm1((M, K), 1), _pm1((N, K), 2)
    ab, wb = a.to(torch.bfloat16).contiguous(), w.to(torch.bfloat16).contiguous()
    y_tc = torch.empty(M, N, dtype=torch.int16, device="cuda")
    zn.call("zkb_bdense_fwd", zn.ptr(ab), zn.ptr(wb), zn.ptr(y_tc), N, M, K, N, 0)
    pa = torch.empty(M, K // 32, dtype=torch.int32, device="cuda")
    pw = torch.empty(N, K // 32, dtype=torch.int32, device="cuda")
    zn.call("zkb_sign_pack", zn.ptr(ab), zn.ptr(pa), ab.numel(), 1, 0)
    zn.call("zkb_sign_pack", zn.ptr(wb), zn.ptr(pw), wb.numel(), 1, 0)
    y_pc = torch.empty(M, N, dtype=torch.int32, device="cuda")
    zn.call("zkb_bgemm_popc", zn.ptr(pa), zn.ptr(pw), zn.ptr(y_pc), N, M, N, K // 32, 0)
    torch.cuda.synchronize()
    y_i8 = torch.empty(M, N, dtype=torch.int16, device="cuda")
    zn.call("zkb_bdense_fwd_i8", zn.ptr(a), zn.ptr(w), zn.ptr(y_i8), N, M, K, N, 0)
    torch.cuda.synchronize()
    ref = (a.float() @ w.float().t()).to(torch.int32)
    assert torch.equal(y_pc, ref)
    assert torch.equal(y_tc.to(torch.int32), ref)
    assert torch.equal(y_i8.to(torch.int32), ref)
    assert int(ref.abs().max()) <= K and bool(((ref + K) % 2 == 0).all())    # parity of a +-1 sum


def _conv_ref(x_nhwc, w_hwio):
    y = torch.nn.functional.conv2d(x_nhwc.permute(0, 3, 1, 2).double(), w_hwio.permute(3, 2, 0, 1).double(),
                                   padding=w_hwio.shape[0] // 2)
    return y.permute(0, 2, 3, 1).contiguous()


def _prep_conv(zn, w_hwio):
    ks, _, cin, cout = w_hwio.shape
    wf = torch.empty(cout, ks * ks * cin, dtype=torch.bfloat16, device="cuda")
    wd = torch.empty(cin, ks * ks * cout, dtype=torch.bfloat16, device="cuda")
    wf8 = torch.empty(cout, ks * ks * cin, dtype=torch.int8, device="cuda")
    zn.call("zkb_prep_conv_weights", zn.ptr(w_hwio.float().contiguous()), ks, cin, cout, zn.ptr(wf), zn.ptr(wd),
            zn.ptr(wf8), 0)
    _prep_conv.last_i8 = wf8
    return wf, wd


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(1, 6, 6, 128, 128), (3, 26, 26, 128, 128), (5, 13, 13, 256, 256),
                                             (7, 6, 6, 512, 512), (2, 7, 5, 128, 256)])
def test_binary_conv_forward_is_bit_exact_with_zero_same_padding(native_lib, B, H, W, Cin, Cout):
    from zookeeper_b200 import native as zn

    x, w = _pm1((B, H, W, Cin), 3), _pm1((3, 3, Cin, Cout), 4)
    wf, _ = _prep_conv(zn, w)
    xb = x.to(torch.bfloat16).contiguous()
    y = torch.full((B, H, W, Cout), -7, dtype=torch.int16, device="cuda")
    zn.call("zkb_bconv2d_fwd", zn.ptr(xb), zn.ptr(wf), zn.ptr(y), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    ref = _conv_ref(x, w)
    assert torch.equal(y.double(), ref)
    # the same layer on the int8 pipe (kind::i8, int32 accumulators): identical integers
    y8 = torch.full((B, H, W, Cout), -7, dtype=torch.int16, device="cuda")
    x8 = x.contiguous()
    zn.call("zkb_bconv2d_fwd_i8", zn.ptr(x8), zn.ptr(_prep_conv.last_i8), zn.ptr(y8), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    assert torch.equal(y8, y)
    # border outputs sum fewer taps: corner has 4 of 9 -> |y| <= 4*Cin there
    assert float(ref[:, 0, 0].abs().max()) <= 4 * Cin


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(5, 13, 13, 128, 256), (3, 26, 26, 128, 128), (64, 6, 6, 512, 512),
                                             (1, 6, 6, 256, 256), (3, 7, 5, 256, 128), (9, 6, 6, 512, 256)])
def test_binary_conv_backward_gemms(native_lib, B, H, W, Cin, Cout):
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(5)
    dy = (torch.randn(B, H, W, Cout, generator=g) * torch.rand(B, H, W, Cout, generator=g)).cuda()
    x, w = _pm1((B, H, W, Cin), 8), _pm1((3, 3, Cin, Cout), 6)
    _, wd = _prep_conv(zn, w)
    hi, lo = _split(dy)
    xb = x.to(torch.bfloat16).contiguous()
    dx = torch.empty(B, H, W, Cin, dtype=torch.float32, device="cuda")
    dw = torch.zeros(3, 3, Cin, Cout, dtype=torch.float32, device="cuda")
    zn.call("zkb_bconv2d_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wd), zn.ptr(dx), B, H, W, Cin, Cout, 3, 0)
    zn.call("zkb_bconv2d_wgrad", zn.ptr(xb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    xr = x.permute(0, 3, 1, 2).double().requires_grad_(True)
    wr = w.permute(3, 2, 0, 1).double().requires_grad_(True)
    torch.nn.functional.conv2d(xr, wr, padding=1).backward(dy.permute(0, 3, 1, 2).double())
    for got, ref in ((dx.double(), xr.grad.permute(0, 2, 3, 1)), (dw.double(), wr.grad.permute(2, 3, 1, 0))):
        rel = (got - ref).norm() / ref.norm()
        assert rel.item() < 1e-5, rel.item()          # hi+lo split: ~2^-17; budget of the path is 1e-4


@pytest.mark.parametrize("nb_i8,nb_bf16", [(28, 14), (29, 15)])
def test_im2col_maps_on_either_side_of_128_kib(native_lib, nb_i8, nb_bf16):
    """zkb_tmap_im2col_nhwc adjusts the encoded im2col tensor map for activations smaller than
    128 KiB (a driver quirk, see csrc/zkb_host.cu and DESIGN.md): a 6x6x128 int8 map is 4608 B per
    image (28 images: 129 024 B, 29 images: 133 632 B), a bf16 one 9216 B (14 / 15 images).  Forward,
    dgrad and wgrad on both sides of the threshold against the float64 convolution."""
    from zookeeper_b200 import native as zn

    H = W = 6
    C = 128
    w = _pm1((3, 3, C, C), 21)
    wf, wd = _prep_conv(zn, w)
    x = _pm1((nb_i8, H, W, C), 22)
    assert (x.numel() < 131072) == (nb_i8 == 28)
    y = torch.full((nb_i8, H, W, C), -7, dtype=torch.int16, device="cuda")
    zn.call("zkb_bconv2d_fwd_i8", zn.ptr(x.contiguous()), zn.ptr(_prep_conv.last_i8), zn.ptr(y), nb_i8, H, W, C, C, 3, 0)
    torch.cuda.synchronize()
    assert torch.equal(y.double(), _conv_ref(x, w))
    g = torch.Generator(device="cpu").manual_seed(23)
    dy = torch.randn(nb_bf16, H, W, C, generator=g).cuda()
    hi, lo = _split(dy)
    assert (hi.numel() * 2 < 131072) == (nb_bf16 == 14)
    xs = _pm1((nb_bf16, H, W, C), 24)
    dx = torch.empty(nb_bf16, H, W, C, dtype=torch.float32, device="cuda")
    dw = torch.zeros(3, 3, C, C, dtype=torch.float32, device="cuda")
    zn.call("zkb_bconv2d_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wd), zn.ptr(dx), nb_bf16, H, W, C, C, 3, 0)
    zn.call("zkb_bconv2d_wgrad", zn.ptr(xs.to(torch.bfloat16).contiguous()), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw),
            nb_bf16, H, W, C, C, 3, 0)
    torch.cuda.synchronize()
    xr = xs.permute(0, 3, 1, 2).double().requires_grad_(True)
    wr = w.permute(3, 2, 0, 1).double().requires_grad_(True)
    torch.nn.functional.conv2d(xr, wr, padding=1).backward(dy.permute(0, 3, 1, 2).double())
    for got, ref in ((dx.double(), xr.grad.permute(0, 2, 3, 1)), (dw.double(), wr.grad.permute(2, 3, 1, 0))):
        assert ((got - ref).norm() / ref.norm()).item() < 1e-5


def test_dense_backward_gemms_and_padded_head(native_lib):
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(9)
    for M, K, N, n_out in ((256, 4608, 1024, 1024), (512, 1024, 64, 10), (37, 1024, 1024, 1024)):
        dy = torch.randn(M, N, generator=g).cuda()
        if n_out < N:
            dy[:, n_out:] = 0
        x, w = _pm1((M, K), 12), _pm1((K, N), 10)
        hi, lo = _split(dy)
        xb, wb = x.to(torch.bfloat16).contiguous(), w.to(torch.bfloat16).contiguous()
        dx = torch.empty(M, K, dtype=torch.float32, device="cuda")
        dw = torch.zeros(K, n_out, dtype=torch.float32, device="cuda")
        zn.call("zkb_bdense_dgrad", zn.ptr(hi), zn.ptr(lo), zn.ptr(wb), zn.ptr(dx), M, K, N, 0)
        zn.call("zkb_bdense_wgrad", zn.ptr(xb), zn.ptr(hi), zn.ptr(lo), zn.ptr(dw), n_out, M, K, N, n_out, 0)
        torch.cuda.synchronize()
        ref_dx = dy.double() @ w.double().t()
        ref_dw = (x.double().t() @ dy.double())[:, :n_out]
        assert ((dx.double() - ref_dx).norm() / ref_dx.norm()).item() < 1e-5
        assert ((dw.double() - ref_dw).norm() / ref_dw.norm()).item() < 1e-5


@pytest.mark.parametrize("B,H,W", [(3, 28, 28), (2, 5, 40), (64, 28, 28)])
def test_first_conv_gram_gives_the_batch_statistics(native_lib, B, H, W):
    """zkb_conv_first_gram + zkb_conv_first_stats_from_gram against the direct sums of the conv
    output (float64), and against zkb_conv_first_bn_stats (the kernel they stand in for)."""
    from zookeeper_b200 import native as zn

    rng = np.random.default_rng(B)
    x = torch.tensor(rng.uniform(-1, 1, (B, H, W, 1)).astype(np.float32)).cuda()
    wq = torch.tensor(np.where(rng.standard_normal((9, 128)) >= 0, 1.0, -1.0).astype(np.float32)).cuda()
    gram = torch.zeros(90, dtype=torch.float64, device="cuda")
    zn.call("zkb_conv_first_gram", zn.ptr(x), B, H, W, zn.ptr(gram), 0)
    taps = torch.stack([x[:, kh:kh + H - 2, kw:kw + W - 2, 0].double().reshape(-1)
                        for kh in range(3) for kw in range(3)])                 # [9, pixels]
    # (the kernel sums up to 32 products per thread in float before going to float64)
    tol = 2e-8 * taps.shape[1] + 1e-6
    np.testing.assert_allclose(gram[:81].reshape(9, 9).cpu().numpy(), (taps @ taps.T).cpu().numpy(), rtol=0, atol=tol)
    np.testing.assert_allclose(gram[81:].cpu().numpy(), taps.sum(1).cpu().numpy(), rtol=0, atol=tol)

    s1, q1, s2, q2 = (torch.zeros(128, dtype=torch.float64, device="cuda") for _ in range(4))
    zn.call("zkb_conv_first_stats_from_gram", zn.ptr(gram), zn.ptr(wq), 128, zn.ptr(s1), zn.ptr(q1), 0)
    zn.call("zkb_conv_first_bn_stats", zn.ptr(x), zn.ptr(wq), B, H, W, 1, 128, 3, zn.ptr(s2), zn.ptr(q2), 0)
    y = taps.T @ wq.double()                                                     # [pixels, 128]
    torch.cuda.synchronize()
    n = y.shape[0]
    np.testing.assert_allclose(s1.cpu().numpy(), y.sum(0).cpu().numpy(), rtol=0, atol=1e-6 * n)
    np.testing.assert_allclose(q1.cpu().numpy(), (y * y).sum(0).cpu().numpy(), rtol=1e-6)
    np.testing.assert_allclose(s1.cpu().numpy(), s2.cpu().numpy(), rtol=1e-5, atol=1e-5 * n)
    np.testing.assert_allclose(q1.cpu().numpy(), q2.cpu().numpy(), rtol=1e-5)


@pytest.mark.parametrize("rows,K,N", [(300, 1024, 1024), (96, 256, 64), (4100, 512, 256)])
def test_forward_gemm_leaves_exact_batch_statistics(native_lib, rows, K, N):
    """zkb_bdense_fwd_i8_stats == zkb_bdense_fwd_i8 + zkb_bn_stats, bit for bit (integer sums)."""
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cuda").manual_seed(rows)
    a = (torch.randint(0, 2, (rows, K), device="cuda", generator=g, dtype=torch.int8) * 2 - 1)
    w = (torch.randint(0, 2, (N, K), device="cuda", generator=g, dtype=torch.int8) * 2 - 1)
    y1 = torch.zeros(rows, N, dtype=torch.int16, device="cuda")
    y2 = torch.zeros_like(y1)
    s1, q1, s2, q2 = (torch.zeros(N, dtype=torch.float64, device="cuda") for _ in range(4))
    zn.call("zkb_bdense_fwd_i8_stats", zn.ptr(a), zn.ptr(w), zn.ptr(y1), N, rows, K, N, zn.ptr(s1), zn.ptr(q1), 0)
    zn.call("zkb_bdense_fwd_i8", zn.ptr(a), zn.ptr(w), zn.ptr(y2), N, rows, K, N, 0)
    zn.call("zkb_bn_stats", zn.ptr(y2), 1, rows, N, zn.ptr(s2), zn.ptr(q2), 0)
    torch.cuda.synchronize()
    ref = a.double() @ w.double().T
    assert torch.equal(y1, y2) and torch.equal(y1.double(), ref)
    assert torch.equal(s1, ref.sum(0)) and torch.equal(q1, (ref * ref).sum(0))
    assert torch.equal(s1, s2) and torch.equal(q1, q2)


def test_conv_forward_gemm_leaves_exact_batch_statistics(native_lib):
    from zookeeper_b200 import native as zn

    B, H, W, Cin, Cout = 5, 13, 13, 128, 256
    s8 = (torch.randint(0, 2, (B, H, W, Cin), device="cuda", dtype=torch.int8) * 2 - 1)
    w = torch.randn(3, 3, Cin, Cout, device="cuda")
    wf = torch.empty(Cout, 9 * Cin, dtype=torch.bfloat16, device="cuda")
    wd = torch.empty(Cin, 9 * Cout, dtype=torch.bfloat16, device="cuda")
    wf8 = torch.empty(Cout, 9 * Cin, dtype=torch.int8, device="cuda")
    zn.call("zkb_prep_conv_weights", zn.ptr(w), 3, Cin, Cout, zn.ptr(wf), zn.ptr(wd), zn.ptr(wf8), 0)
    y1 = torch.zeros(B, H, W, Cout, dtype=torch.int16, device="cuda")
    y2 = torch.zeros_like(y1)
    s1, q1 = (torch.zeros(Cout, dtype=torch.float64, device="cuda") for _ in range(2))
    zn.call("zkb_bconv2d_fwd_i8_stats", zn.ptr(s8), zn.ptr(wf8), zn.ptr(y1), B, H, W, Cin, Cout, 3, zn.ptr(s1), zn.ptr(q1), 0)
    zn.call("zkb_bconv2d_fwd_i8", zn.ptr(s8), zn.ptr(wf8), zn.ptr(y2), B, H, W, Cin, Cout, 3, 0)
    torch.cuda.synchronize()
    assert torch.equal(y1, y2)
    yd = y2.double().reshape(-1, Cout)
    assert torch.equal(s1, yd.sum(0)) and torch.equal(q1, (yd * yd).sum(0))


def test_sign_thresholds_reproduce_the_float_batchnorm_sign(native_lib):
    """thr[c] from zkb_bn_sign_thresholds: y >= thr[c]  <=>  zkb_bn_apply_sign gives +1, for every
    int16 y (edge channels included: never / always +1, huge and tiny rstd)."""
    from zookeeper_b200 import native as zn

    C = 128
    rng = np.random.default_rng(0)
    mean = rng.normal(0, 50, C).astype(np.float32)
    rstd = (1.0 / np.sqrt(rng.uniform(1.0, 4000.0, C))).astype(np.float32)
    beta = rng.normal(0, 1, C).astype(np.float32)
    mean[0], beta[0] = 1e9, 0.0            # never +1
    mean[1], beta[1] = -1e9, 0.0           # always +1
    rstd[2], beta[2] = 1e-6, 1e-3          # beta decides alone
    rstd[3], mean[3], beta[3] = 30.0, 0.4999, 0.0
    d = lambda a: torch.tensor(a).cuda()
    dm, dr, db = d(mean), d(rstd), d(beta)
    thr = torch.zeros(C, dtype=torch.int32, device="cuda")
    zn.call("zkb_bn_sign_thresholds", zn.ptr(dm), zn.ptr(dr), zn.ptr(db), C, zn.ptr(thr), 0)
    ys = torch.arange(-32768, 32768, dtype=torch.int16, device="cuda")
    y = ys[:, None].expand(-1, C).contiguous()
    s8 = torch.zeros(y.shape, dtype=torch.int8, device="cuda")
    zn.call("zkb_bn_apply_sign", zn.ptr(y), 1, y.shape[0], C, zn.ptr(dm), zn.ptr(dr), zn.ptr(db), 0, 0, zn.ptr(s8), 0)
    torch.cuda.synchronize()
    want = s8 > 0
    got = y.int() >= thr[None, :]
    assert torch.equal(want, got)
    assert int(thr[0]) == 2**31 - 1 and int(thr[1]) == -2**31


def test_sign_pool_and_threshold_gemm(native_lib):
    """zkb_bdense_fwd_i8_sign == (zkb_bdense_fwd_i8 >= thr); zkb_maxpool2_sign_i8 == max of 4."""
    from zookeeper_b200 import native as zn

    rows, K, N = 333, 512, 256
    g = torch.Generator(device="cuda").manual_seed(3)
    a = (torch.randint(0, 2, (rows, K), device="cuda", generator=g, dtype=torch.int8) * 2 - 1)
    w = (torch.randint(0, 2, (N, K), device="cuda", generator=g, dtype=torch.int8) * 2 - 1)
    thr = torch.randint(-40, 40, (N,), device="cuda", generator=g, dtype=torch.int32)
    thr[0], thr[1] = 2**31 - 1, -2**31
    s = torch.zeros(rows, N, dtype=torch.int8, device="cuda")
    zn.call("zkb_bdense_fwd_i8_sign", zn.ptr(a), zn.ptr(w), zn.ptr(thr), zn.ptr(s), N, rows, K, N, 0)
    torch.cuda.synchronize()
    y = (a.float() @ w.float().T).int()
    assert torch.equal(s, torch.where(y >= thr[None, :], 1, -1).to(torch.int8))

    B, H, W, C = 3, 13, 26, 64
    t = (torch.randint(0, 2, (B, H, W, C), device="cuda", generator=g, dtype=torch.int8) * 2 - 1)
    out = torch.zeros(B, H // 2, W // 2, C, dtype=torch.int8, device="cuda")
    zn.call("zkb_maxpool2_sign_i8", zn.ptr(t), B, H, W, C, zn.ptr(out), 0)
    torch.cuda.synchronize()
    ref = torch.nn.functional.max_pool2d(t.permute(0, 3, 1, 2).float(), 2).permute(0, 2, 3, 1).to(torch.int8)
    assert torch.equal(out, ref)


def test_full_size_train_step_properties(native_lib):
    """configs[1] (batch 4096): too large for the CPU oracle, so check properties that do not depend
    on size: integer layer outputs respect |y| <= K and the +-1 parity, pooled = max of its window,
    BatchNorm output has zero mean / unit variance per channel, the loss starts near ln(10), gradients
    are finite and weights stay clipped after the step."""
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    B = 4096
    eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    eng.init_glorot(0)
    g = torch.Generator(device="cpu").manual_seed(0)
    img = torch.randint(0, 256, (B, 28, 28, 1), generator=g, dtype=torch.uint8).cuda()
    lab = torch.randint(0, 10, (B,), generator=g, dtype=torch.int64).cuda()
    eng.train_step(img, lab, None, pad_size=32, seed=1)
    torch.cuda.synchronize()
    assert abs(eng.last_loss() - np.log(10)) < 0.5
    for L in eng.L[1:]:
        l = L.spec
        K = l.cin * (l.ksize ** 2 if l.kind == "conv" else 1)
        fused_pool = l.kind == "conv" and l.pool and eng._pool_in_epilogue(L)
        y = (L.p if fused_pool else L.y)[..., : l.cout].to(torch.int32)     # (fused conv+pool layers store only p)
        assert int(y.abs().max()) <= K
        if l.kind == "dense":
            assert bool(((y + K) % 2 == 0).all())
        else:
            inner = y[:, 1:-1, 1:-1]                          # all 9 taps valid there
            assert bool(((inner + K) % 2 == 0).all())
        if l.pool and not eng._pool_in_epilogue(L):
            oh, ow = L.out_hw
            win = L.y[:, : oh // 2 * 2, : ow // 2 * 2].reshape(B, oh // 2, 2, ow // 2, 2, L.cp)
            assert torch.equal(win.amax(dim=(2, 4)), L.p)
    # BN statistics of layer 1: recompute from the pooled map
    L = eng.L[1]
    p = L.p.double().reshape(-1, L.cp)
    assert torch.allclose(L.mean.double(), p.mean(0), rtol=1e-6, atol=1e-6)
    assert torch.allclose(L.rstd.double(), torch.rsqrt(p.var(0, unbiased=False) + 1e-3), rtol=1e-5)
    assert bool(torch.isfinite(eng.grads).all()) and float(eng.grads.abs().max()) > 0
    assert float(eng.params[eng.kbase:].abs().max()) <= 1.0
