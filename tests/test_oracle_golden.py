"""The CPU oracle against the committed golden vectors (tests/golden, made by
oracle/make_golden.py) and against properties that must hold regardless of implementation.
Parity is unpinned against TensorFlow (see oracle/binarynet_oracle.py); these tests pin the oracle
against drift and check its internal consistency."""
import os

import numpy as np
import torch

from oracle import binarynet_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_train_step_golden():
    g = np.load(os.path.join(GOLD, "binarynet_b4_step.npz"))
    layers = O.binarynet_layers((28, 28, 1), 128, 1024, 10)
    params = O.glorot_uniform_params(layers, 0)
    aug = O.augmentation_params(2024, 0, 4, 32, 28, 28)
    assert np.array_equal(aug, g["aug"])
    x = O.preprocess(g["images"], (28, 28), True, 32, aug)
    assert np.array_equal(x[0, 0, :, 0], g["x_first_row"])
    assert float(x.astype(np.float64).sum()) == float(g["x_sum"])
    r = O.forward_backward(params, layers, x, g["labels"], torch.float64, keep_acts=True)
    assert abs(r.loss - float(g["loss"])) < 1e-12
    np.testing.assert_allclose(r.logits, g["logits"], rtol=1e-10, atol=1e-12)
    for k in ("kernel0", "kernel8", "beta8"):
        np.testing.assert_allclose(r.grads[k], g[f"grad_{k}"], rtol=1e-9, atol=1e-14)
    for i in range(len(layers)):
        assert abs(np.abs(r.acts[f"y{i}"]).sum() - float(g[f"y{i}_abs_sum"])) <= 1e-9 * float(g[f"y{i}_abs_sum"])
        assert abs(np.linalg.norm(r.grads[f"kernel{i}"]) - float(g[f"grad_kernel{i}_norm"])) <= 1e-9 * float(g[f"grad_kernel{i}_norm"])
        np.testing.assert_allclose(r.batch_stats[f"mean{i}"], g[f"mean{i}"], rtol=1e-10, atol=1e-12)
    m0 = {k: np.zeros_like(v) for k, v in params.items()}
    p1, _, _ = O.adam_clip_step(params, r.grads, m0, m0, 1, dtype=np.float32)
    np.testing.assert_array_equal(p1["kernel0"], g["new_kernel0"])
    np.testing.assert_array_equal(p1["beta8"], g["new_beta8"])


def test_binary_layers_produce_exact_integers_and_f32_tracks_f64():
    layers = O.binarynet_layers((28, 28, 1), 128, 1024, 10)
    params = O.glorot_uniform_params(layers, 0)
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (8, 28, 28, 1), dtype=np.uint8)
    lab = rng.integers(0, 10, (8,)).astype(np.int64)
    x = O.preprocess(img, (28, 28), False)
    r64 = O.forward_backward(params, layers, x, lab, torch.float64, keep_acts=True)
    r32 = O.forward_backward(params, layers, x, lab, torch.float32, keep_acts=True)
    for i in range(1, len(layers)):
        y = r64.acts[f"y{i}"]
        assert np.array_equal(y, np.rint(y)), f"layer {i} output is not integer valued"
        k = layers[i].cin * (layers[i].ksize ** 2 if layers[i].kind == "conv" else 1)
        assert np.abs(y).max() <= k
    for k, v in r64.grads.items():
        rel = np.linalg.norm(r32.grads[k] - v) / max(np.linalg.norm(v), 1e-30)
        assert rel < 1e-4, (k, rel)


def test_bitops_golden_and_identity_with_float_gemm():
    g = np.load(os.path.join(GOLD, "bitops.npz"))
    pa, pb = O.sign_pack(g["a"]), O.sign_pack(g["b"])
    assert np.array_equal(pa, g["pa"]) and np.array_equal(pb, g["pb"])
    # sign(0) = sign(-0.0) = +1 (larq): first four entries of row 0 are 0, -0, +tiny, -tiny
    assert (int(pa[0, 0]) & 0xF) == 0b0111
    y = O.xnor_popcount_gemm(pa, pb, 128)
    assert np.array_equal(y, g["y"])
    sa = np.where(g["a"] >= 0, 1.0, -1.0)
    sb = np.where(g["b"] >= 0, 1.0, -1.0)
    assert np.array_equal(y, (sa @ sb.T).astype(np.int32))


def test_preprocess_golden_and_semantics():
    g = np.load(os.path.join(GOLD, "preprocess.npz"))
    img, aug = g["img"], g["aug"]
    assert np.array_equal(O.preprocess(img, (32, 32), True, 40, aug), g["train"])
    assert np.array_equal(O.preprocess(img, (28, 28), False), g["eval_crop"])
    assert np.array_equal(O.preprocess(img, (36, 36), False), g["eval_pad"])
    # eval crop is the centre crop; eval pad puts -1.0 (u8 zero) around the image
    ref = img[:, 2:30, 2:30].astype(np.float32) / np.float32(127.5) - np.float32(1.0)
    assert np.array_equal(g["eval_crop"], ref)
    assert np.all(g["eval_pad"][:, :2] == -1.0) and np.all(g["eval_pad"][:, :, -2:] == -1.0)
    # training: sample 0 has ox=0, oy=8 (pad 4 each side of 32->40), flipped
    canvas = np.zeros((40, 40, 3), np.uint8)
    canvas[4:36, 4:36] = img[0]
    want = canvas[8:40, 0:32][:, ::-1].astype(np.float32) / np.float32(127.5) - np.float32(1.0)
    assert np.array_equal(g["train"][0], want)
    # empty batch
    assert O.preprocess(img[:0], (28, 28), False).shape == (0, 28, 28, 3)


def test_augmentation_rng_is_uniform_enough():
    aug = O.augmentation_params(7, 3, 20000, 32, 28, 28)
    assert aug[:, 0].min() == 0 and aug[:, 0].max() == 4 and aug[:, 1].max() == 4
    for col, n in ((0, 5), (1, 5), (2, 2)):
        counts = np.bincount(aug[:, col], minlength=n)
        assert np.all(np.abs(counts / 20000 - 1 / n) < 0.02)
    assert not np.array_equal(aug, O.augmentation_params(7, 4, 20000, 32, 28, 28))


def test_pool_tie_rule_and_ste_mask():
    x = torch.tensor([[[[1.0], [1.0]], [[1.0], [1.0]]]], dtype=torch.float64, requires_grad=True)
    O.maxpool2(x).sum().backward()
    assert x.grad.flatten().tolist() == [1.0, 0.0, 0.0, 0.0]       # first maximum wins
    x = torch.tensor([[[[0.0], [2.0]], [[2.0], [1.0]]]], dtype=torch.float64, requires_grad=True)
    O.maxpool2(x).sum().backward()
    assert x.grad.flatten().tolist() == [0.0, 1.0, 0.0, 0.0]
    # odd trailing row/column is dropped ('valid')
    assert O.maxpool2(torch.zeros(1, 13, 13, 2)).shape == (1, 6, 6, 2)
    z = torch.tensor([-2.0, -1.0, -0.0, 0.0, 0.5, 1.0, 1.5], dtype=torch.float64, requires_grad=True)
    s = O.ste_sign(z)
    assert s.tolist() == [-1, -1, 1, 1, 1, 1, 1]
    s.sum().backward()
    assert z.grad.tolist() == [0, 1, 1, 1, 1, 1, 0]               # |x| <= 1 inclusive
