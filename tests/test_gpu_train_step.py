"""Parity of the CUDA train step against the CPU oracle.

Bit-exact: preprocessing output, integer conv / dense outputs (through the sign chain).
Tolerance (north_star): loss, gradients and updated weights within 1e-4 relative; the norm used is
the Frobenius norm per tensor (a per-element ratio is meaningless where g ~ 0).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _setup(B, seed=1234, input_shape=(28, 28, 1), filters=128, dense=1024, nc=10):
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    layers_o = O.binarynet_layers(input_shape, filters, dense, nc)
    params = O.glorot_uniform_params(layers_o, 0)
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (B,) + tuple(input_shape), dtype=np.uint8)
    lab = rng.integers(0, nc, (B,)).astype(np.int64)
    eng = BinaryNetEngine(binarynet_layers(input_shape, filters, dense, nc), input_shape, B, nc)
    eng.load_params(params)
    return O, layers_o, params, img, lab, eng


@pytest.mark.parametrize("B", [8, 37])
def test_train_step_matches_oracle(native_lib, B):
    O, layers_o, params, img, lab, eng = _setup(B)
    aug = O.augmentation_params(1234, 0, B, 32, 28, 28)
    x = O.preprocess(img, (28, 28), True, 32, aug)
    ref = O.forward_backward(params, layers_o, x, lab, torch.float64, keep_acts=True)

    d_img = torch.as_tensor(img).cuda()
    d_lab = torch.as_tensor(lab).cuda()
    eng.train_step(d_img, d_lab, None, pad_size=32, aug=torch.as_tensor(aug).cuda(), apply_update=False)
    torch.cuda.synchronize()

    # preprocessing: bit-exact
    assert np.array_equal(eng.x0.cpu().numpy(), x)
    assert np.array_equal(eng.labels.cpu().numpy(), lab.astype(np.int32))
    # first conv (float): close; binary layers: exact integers given identical sign inputs
    # (the engine recomputes this map inside its fused kernels and never stores it; the stand-alone
    # entry point computes the same thing from the same prepared sign weights)
    from zookeeper_b200 import native as zn
    y0 = torch.empty(B, 26, 26, 128, dtype=torch.float32, device="cuda")
    zn.call("zkb_conv_first_fwd", zn.ptr(eng.x0), zn.ptr(eng.L[0].wq), zn.ptr(y0), B, 28, 28, 1, 128, 3, 0)
    torch.cuda.synchronize()
    assert _rel(y0.cpu().numpy(), ref.acts["y0"]) < 1e-6
    s0 = eng.L[0].s.float().cpu().numpy()
    a0 = ref.acts["a0"]
    agree = (s0 == np.where(a0 >= 0, 1.0, -1.0)) | (np.abs(a0) < 1e-5)
    assert agree.all(), "first-layer sign output differs from the oracle"
    for i in range(1, 9):
        L = eng.L[i]
        if not (L.spec.pool and eng._pool_in_epilogue(L)):     # (fused conv+pool layers never store y)
            y = L.y.cpu().numpy().astype(np.int64)[..., : L.spec.cout]
            yr = np.rint(ref.acts[f"y{i}"]).astype(np.int64)
            assert np.array_equal(y, yr), f"layer {i}: integer conv/dense output differs"
        if L.spec.pool:
            assert np.array_equal(L.p.cpu().numpy().astype(np.int64), np.rint(ref.acts[f"p{i}"]).astype(np.int64))
    assert abs(eng.last_loss() - ref.loss) <= REL_TOL * abs(ref.loss)
    assert eng.last_accuracy() == pytest.approx(ref.accuracy)
    assert _rel(eng.logits(), ref.logits) < REL_TOL
    g = eng.get_grads()
    for k, v in ref.grads.items():
        assert _rel(g[k], v) < REL_TOL, f"grad {k}: rel {_rel(g[k], v):.3g}"

    # moving statistics
    mv = O.update_moving_stats(O.init_moving_stats(layers_o), layers_o, ref.batch_stats)
    got = eng.get_moving()
    for k, v in mv.items():
        assert _rel(got[k], v) < 1e-5, k

    # optimizer kernel in isolation: oracle Adam (fp32 operation order of TF's ApplyAdam) fed with
    # the ENGINE's gradients must reproduce the engine's weights to fp32 rounding
    m0 = {k: np.zeros_like(v) for k, v in params.items()}
    eng_grads = eng.get_grads()
    exp32, _, _ = O.adam_clip_step(params, eng_grads, m0, m0, 1, dtype=np.float32)
    exp64, _, _ = O.adam_clip_step(params, ref.grads, m0, m0, 1, dtype=np.float64)
    eng._begin()
    eng.optimizer_step()
    torch.cuda.synchronize()
    gp = eng.get_params()
    for k in params:
        assert _rel(gp[k], exp32[k]) < 2e-6, f"adam kernel {k}: {_rel(gp[k], exp32[k]):.3g}"
        if k.startswith("kernel"):
            # end to end (oracle gradients, exact arithmetic): updated latent weights within 1e-4
            assert _rel(gp[k], exp64[k]) < REL_TOL, f"param {k}: {_rel(gp[k], exp64[k]):.3g}"
            assert np.abs(gp[k]).max() <= 1.0
        else:
            # beta starts at 0, so the new value IS the step lr*g/(|g|+eps'): elementwise it is as
            # ill-conditioned as g/|g| where g ~ 0; bound the median and the bulk instead
            err = np.abs(gp[k] - exp64[k]) / 5e-3
            assert np.median(err) < 1e-4 and (err > 1e-2).mean() < 0.05, f"beta step {k}"


@pytest.mark.parametrize("B", [5, 37, 64])
def test_fused_first_layer_backward_equals_separate_kernels(native_lib, B, monkeypatch):
    """zkb_bconv2d_dgrad_first_bwd (+ finalize) against zkb_bconv2d_dgrad -> conv_first_bwd_reduce ->
    bn_bwd_finalize -> conv_first_bwd_wgrad on the same state: all gradients equal up to float
    summation order."""
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    params = O.glorot_uniform_params(O.binarynet_layers((28, 28, 1), 128, 1024, 10), 0)
    rng = np.random.default_rng(7)
    img = torch.as_tensor(rng.integers(0, 256, (B, 28, 28, 1), dtype=np.uint8)).cuda()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int64)).cuda()
    grads = []
    for fused in ("1", "0"):
        monkeypatch.setenv("ZKB_FUSE_FIRST_BWD", fused)
        eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
        assert eng.fuse_first_backward == (fused == "1")
        assert (eng.L[0].ds is None) == (fused == "1")
        eng.load_params(params)
        eng.train_step(img, lab, None, pad_size=32, seed=3, apply_update=False)
        torch.cuda.synchronize()
        grads.append(eng.get_grads())
    a, b = grads
    for k in a:
        if k in ("kernel0", "beta0"):
            assert _rel(a[k], b[k]) < 5e-6, f"{k}: {_rel(a[k], b[k]):.3g}"
        else:       # same kernels on both sides (float atomics: summation order may differ)
            assert _rel(a[k], b[k]) < 1e-6, k


@pytest.mark.parametrize("B,shape", [(5, (28, 28, 1)), (37, (28, 28, 1)), (6, (32, 32, 3))])
def test_pool_in_forward_epilogue_equals_separate_pool_kernel(native_lib, monkeypatch, B, shape):
    """zkb_bconv2d_fwd_i8_pool against zkb_bconv2d_fwd_i8 + zkb_maxpool2_i16 inside a full step:
    pooled maps, winning slots, batch statistics and sign maps bit-identical (integers summed
    exactly), gradients up to float summation order.  MNIST shape: layers 1 (26x26) and 5 (6x6) are
    fused, layer 3 (13x13) is not; CIFAR shape: layer 1 (30x30) only."""
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    params = O.glorot_uniform_params(O.binarynet_layers(shape, 128, 1024, 10), 0)
    rng = np.random.default_rng(B)
    img = torch.as_tensor(rng.integers(0, 256, (B,) + shape, dtype=np.uint8)).cuda()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int64)).cuda()
    out = []
    for fused in ("1", "0"):
        monkeypatch.setenv("ZKB_FUSE_POOL", fused)
        eng = BinaryNetEngine(binarynet_layers(shape), shape, B, 10)
        eng.load_params(params)
        which = [i for i, L in enumerate(eng.L) if L.spec.kind == "conv" and L.spec.pool and eng._pool_in_epilogue(L)]
        assert which == (([1, 5] if shape[0] == 28 else [1]) if fused == "1" else [])
        for _ in range(2):       # twice: the staging slabs and accumulator stages are reused across tiles / steps
            eng.train_step(img, lab, None, pad_size=shape[0] + 4, seed=3, apply_update=False)
        torch.cuda.synchronize()
        pooled = [(L.p.cpu().numpy().copy(), L.arg.cpu().numpy().copy(), L.s8.cpu().numpy().copy())
                  for L in eng.L if L.spec.kind == "conv" and L.spec.pool]
        stats = torch.cat([torch.cat([L.mean, L.rstd]) for L in eng.L]).cpu().numpy()
        out.append((pooled, stats, eng.last_loss(), eng.get_grads()))
    (pa, sa, la, ga), (pb, sb, lb, gb) = out
    for k, (a, b) in enumerate(zip(pa, pb)):
        for x, y, what in zip(a, b, ("pooled map", "slot", "sign map")):
            assert np.array_equal(x, y), f"pooling layer {k}: {what} differs"
    assert np.array_equal(sa, sb) and la == lb
    for k in ga:
        assert _rel(ga[k], gb[k]) < 1e-5, k


def test_bn_backward_sums_from_dgrad_epilogue_equal_the_reduction_kernel(native_lib, monkeypatch):
    """zkb_bconv2d_dgrad_bn against zkb_bconv2d_dgrad + zkb_bn_bwd_reduce inside a full step."""
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    B = 48
    params = O.glorot_uniform_params(O.binarynet_layers((28, 28, 1), 128, 1024, 10), 0)
    rng = np.random.default_rng(11)
    img = torch.as_tensor(rng.integers(0, 256, (B, 28, 28, 1), dtype=np.uint8)).cuda()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int64)).cuda()
    out = []
    for fused in ("1", "0"):
        monkeypatch.setenv("ZKB_FUSE_BN_BWD", fused)
        eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
        assert eng.fuse_bn_backward_sums == (fused == "1")
        assert [eng._sums_come_from_dgrad(i) for i in range(9)] == [False] + [fused == "1"] * 4 + [False] * 4
        eng.load_params(params)
        eng.train_step(img, lab, None, pad_size=32, seed=3, apply_update=False)
        torch.cuda.synchronize()
        sums = torch.cat([torch.cat([L.sum_da, L.sum_dax]) for L in eng.L[1:5]]).cpu().numpy()
        out.append((sums, eng.get_grads()))
    (sa, ga), (sb, gb) = out
    assert _rel(sa, sb) < 1e-5          # float partial sums in a different order
    for k in ga:
        assert _rel(ga[k], gb[k]) < 1e-5, k


def test_second_step_from_engine_state(native_lib):
    """Step 2 (Adam t = 2, non-zero moments, re-prepared sign weights, in-kernel augmentation RNG).

    Binarised training is chaotic across steps: a weight with |w| < lr whose gradient is ~0 may
    flip sign after step 1 in one fp implementation and not in the other, and the loss then
    differs at the 1e-2 level.  Parity per step is therefore checked by re-seeding the oracle with
    the ENGINE's state after step 1 and comparing everything step 2 produces."""
    B = 16
    O, layers_o, params, img, lab, eng = _setup(B, seed=7)
    d_img, d_lab = torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda()
    eng.train_step(d_img, d_lab, None, pad_size=32, seed=99)   # step 1, in-kernel counter RNG
    torch.cuda.synchronize()
    aug0 = O.augmentation_params(99, 0, B, 32, 28, 28)
    assert np.array_equal(eng.x0.cpu().numpy(), O.preprocess(img, (28, 28), True, 32, aug0))
    p1 = eng.get_params()
    m1, v1 = eng.get_adam_state()

    aug1 = O.augmentation_params(99, 1, B, 32, 28, 28)
    x1 = O.preprocess(img, (28, 28), True, 32, aug1)
    ref = O.forward_backward(p1, layers_o, x1, lab, torch.float64)
    eng.train_step(d_img, d_lab, None, pad_size=32, seed=99)   # step 2
    torch.cuda.synchronize()
    assert np.array_equal(eng.x0.cpu().numpy(), x1), "in-kernel augmentation differs from oracle RNG"
    assert abs(eng.last_loss() - ref.loss) <= REL_TOL * abs(ref.loss)
    g = eng.get_grads()
    for k, v in ref.grads.items():
        assert _rel(g[k], v) < REL_TOL, f"grad {k}: rel {_rel(g[k], v):.3g}"
    exp32, _, _ = O.adam_clip_step(p1, g, m1, v1, 2, dtype=np.float32)
    exp64, _, _ = O.adam_clip_step(p1, ref.grads, m1, v1, 2, dtype=np.float64)
    gp = eng.get_params()
    for k in p1:
        assert _rel(gp[k], exp32[k]) < 2e-6, f"adam kernel {k}"
        if k.startswith("kernel"):
            assert _rel(gp[k], exp64[k]) < REL_TOL, f"param {k}"


def test_eval_step_matches_oracle(native_lib):
    B = 16
    O, layers_o, params, img, lab, eng = _setup(B, seed=11)
    rng = np.random.default_rng(5)
    moving = {}
    for i, l in enumerate(layers_o):
        moving[f"moving_mean{i}"] = rng.normal(0, 2.0, l.cout).astype(np.float32)
        moving[f"moving_var{i}"] = rng.uniform(50.0, 500.0, l.cout).astype(np.float32)
    eng.load_params(params, moving)
    x = O.preprocess(img, (28, 28), False)
    ref = O.predict(params, moving, layers_o, x)
    eng.eval_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(eng.x0.cpu().numpy(), x)
    assert _rel(eng.logits(), ref) < 1e-4


@pytest.mark.parametrize("B", [16, 200])
def test_folded_inference_is_bit_identical_to_the_batchnorm_path(native_lib, monkeypatch, B):
    """eval_step with BatchNorm + sign folded into integer thresholds inside the GEMM epilogues
    (the default) against eval_step through the int16 maps and zkb_bn_apply_sign: same logits, bit
    for bit, and both match the oracle's predict()."""
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    layers_o = O.binarynet_layers((28, 28, 1), 128, 1024, 10)
    params = O.glorot_uniform_params(layers_o, 0)
    rng = np.random.default_rng(B)
    moving = {}
    for i, l in enumerate(layers_o):
        # means near the typical |y| scale so that thresholds cut through the populated range
        moving[f"moving_mean{i}"] = rng.normal(0, 3.0, l.cout).astype(np.float32)
        moving[f"moving_var{i}"] = rng.uniform(20.0, 400.0, l.cout).astype(np.float32)
    for k in list(params):
        if k.startswith("beta"):
            params[k] = rng.normal(0, 0.5, params[k].shape).astype(np.float32)
    img = rng.integers(0, 256, (B, 28, 28, 1), dtype=np.uint8)
    lab = rng.integers(0, 10, (B,)).astype(np.int64)
    outs = []
    for fold in ("1", "0"):
        monkeypatch.setenv("ZKB_FOLD_EVAL", fold)
        eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
        assert eng.fold_eval == (fold == "1")
        eng.load_params(params, moving)
        eng.eval_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda())
        torch.cuda.synchronize()
        outs.append((eng.L[-1].logits.cpu().numpy().copy(), [L.s8.cpu().numpy().copy() for L in eng.L[:-1]],
                     eng.last_loss(), eng.last_accuracy()))
    for i, (a, b) in enumerate(zip(outs[0][1], outs[1][1])):
        assert np.array_equal(a, b), f"layer {i}: folded signs differ"
    assert np.array_equal(outs[0][0], outs[1][0])
    assert outs[0][2:] == outs[1][2:]
    ref = O.predict(params, moving, layers_o, O.preprocess(img, (28, 28), False))
    assert _rel(outs[0][0][:, :10], ref) < 1e-4


def test_cifar_shaped_config_matches_oracle(native_lib):
    """BASELINE.json configs[2]: the same BinaryNet factory on 32x32x3 inputs (first conv K = 27,
    odd 15x15 -> 7x7 pooling drops the trailing row/column, flatten 3*3*512)."""
    B, shape = 6, (32, 32, 3)
    O, layers_o, params, img, lab, eng = _setup(B, seed=21, input_shape=shape)
    aug = O.augmentation_params(5, 0, B, 40, 32, 32)
    x = O.preprocess(img, (32, 32), True, 40, aug)
    ref = O.forward_backward(params, layers_o, x, lab, torch.float64, keep_acts=True)
    eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), None, pad_size=40,
                   aug=torch.as_tensor(aug).cuda(), apply_update=False)
    torch.cuda.synchronize()
    assert np.array_equal(eng.x0.cpu().numpy(), x)
    for i in range(1, 9):
        L = eng.L[i]
        if L.spec.pool and eng._pool_in_epilogue(L):
            assert np.array_equal(L.p.cpu().numpy().astype(np.int64), np.rint(ref.acts[f"p{i}"]).astype(np.int64)), f"layer {i}"
            continue
        y = L.y.cpu().numpy().astype(np.int64)[..., : L.spec.cout]
        assert np.array_equal(y, np.rint(ref.acts[f"y{i}"]).astype(np.int64)), f"layer {i}"
    assert abs(eng.last_loss() - ref.loss) <= REL_TOL * abs(ref.loss)
    g = eng.get_grads()
    for k, v in ref.grads.items():
        assert _rel(g[k], v) < REL_TOL, f"grad {k}: rel {_rel(g[k], v):.3g}"


def test_training_converges_like_the_oracle_and_checkpoint_roundtrip(native_lib, tmp_path):
    """north_star: "final MNIST-shaped accuracy must agree within 0.5 pt".  MNIST itself is not
    available offline, so both sides train TO CONVERGENCE on the same learnable MNIST-shaped task
    (10 fixed prototypes + noise): 320 steps of batch 64 with the same batches and the same
    augmentation stream, then inference with moving statistics on the same 2048 held-out samples.
    Both must reach >= 97 % and agree within 0.5 pt.  (The two trajectories do not hold the same
    weights -- float summation order flips near-zero latent weights differently, see the module
    docstring -- which is why the criterion is the converged accuracy, as north_star states it.)
    Also checks checkpoint save -> load -> identical eval, and that the oracle's inference on the
    ENGINE's trained state predicts what the engine's validation pass predicts."""
    from oracle import binarynet_oracle as O
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    rng = np.random.default_rng(3)
    protos = rng.integers(0, 256, (10, 28, 28, 1)).astype(np.float32)

    def make(n):
        lab = rng.integers(0, 10, (n,)).astype(np.int64)
        img = np.clip(protos[lab] * 0.6 + rng.normal(60, 40, (n, 28, 28, 1)), 0, 255).astype(np.uint8)
        return img, lab

    B, steps, n_test = 64, 320, 2048
    test_img, test_lab = make(n_test)
    layers_o = O.binarynet_layers()
    p = O.glorot_uniform_params(layers_o, 0)
    m = {k: np.zeros_like(v) for k, v in p.items()}
    v = {k: np.zeros_like(v_) for k, v_ in p.items()}
    mov = O.init_moving_stats(layers_o)
    eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    eng.load_params(p)
    for s in range(steps):
        img, lab = make(B)
        aug = O.augmentation_params(11, s, B, 32, 28, 28)
        r = O.forward_backward(p, layers_o, O.preprocess(img, (28, 28), True, 32, aug), lab, torch.float32)
        p, m, v = O.adam_clip_step(p, r.grads, m, v, s + 1, dtype=np.float32)
        mov = O.update_moving_stats(mov, layers_o, r.batch_stats)
        eng.train_step(torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda(), None, pad_size=32, seed=11)
        if s >= steps - 4:   # batch-statistics accuracy of the last steps: both have learnt the task
            assert r.accuracy > 0.9 and eng.last_accuracy() > 0.9, (s, r.accuracy, eng.last_accuracy())
    torch.cuda.synchronize()
    xt = O.preprocess(test_img, (28, 28), False)
    acc_oracle = float((O.predict(p, mov, layers_o, xt, dtype=torch.float32).argmax(1) == test_lab).mean())

    ev = BinaryNetEngine(binarynet_layers(), (28, 28, 1), n_test, 10)
    ck = str(tmp_path / "ck.npz")
    eng.save_checkpoint(ck)
    ev.load_checkpoint(ck)
    assert ev.step_count == steps
    ev.eval_step(torch.as_tensor(test_img).cuda(), torch.as_tensor(test_lab).cuda())
    torch.cuda.synchronize()
    acc_engine = ev.last_accuracy()
    assert acc_oracle >= 0.97 and acc_engine >= 0.97, (acc_oracle, acc_engine)
    assert abs(acc_engine - acc_oracle) <= 0.005, (acc_oracle, acc_engine)      # 0.5 pt
    # same state, two implementations of inference
    same_state = O.predict(eng.get_params(), eng.get_moving(), layers_o, xt, dtype=torch.float32)
    got = ev.logits()
    agree = (same_state.argmax(1) == got.argmax(1)).mean()
    assert agree >= 0.995, agree                     # ties of near-equal logits aside
    assert abs(float((same_state.argmax(1) == test_lab).mean()) - acc_engine) <= 0.005
    # the checkpoint holds exactly what the engine holds
    for k, a in eng.get_params().items():
        assert np.array_equal(a, ev.get_params()[k])


def test_keras_shaped_fit_runs_on_the_engine(native_lib):
    """The example's run(): dataset plan -> preprocessing lowering -> Sequential.fit/evaluate."""
    import importlib.util
    import os

    from zookeeper_b200.core import configure

    spec = importlib.util.spec_from_file_location(
        "larq_experiment_example", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                "examples", "larq_experiment.py"))
    mod = importlib.util.module_from_spec(spec)
    from zookeeper_b200.core import cli
    saved = dict(cli.commands)
    try:
        spec.loader.exec_module(mod)
        task = mod.BinaryNetMnist()
        configure(task, {"epochs": 2, "batch_size": 64, "dataset.num_train_examples": 256,
                         "dataset.num_validation_examples": 128})
        task.run()
        h = task.model.history
        assert len(h["loss"]) == 2 and all(np.isfinite(h["loss"])) and len(h["val_accuracy"]) == 2
        assert task.model.engine.step_count == 2 * (256 // 64)
    finally:
        cli.commands = saved


def test_validation_at_another_batch_size_sees_the_weights_being_trained(native_lib, tmp_path):
    """Engines of different batch sizes that alias one state (what Sequential.fit/evaluate build
    when validation data is batched differently from training data): the validation engine sees
    every training step, its cached inference preparation is refreshed, the step counter and the
    Adam moments are shared, and a checkpoint written through either holds the trained weights."""
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    rng = np.random.default_rng(9)
    protos = rng.integers(0, 256, (10, 28, 28, 1)).astype(np.float32)

    def make(n):
        lab = rng.integers(0, 10, (n,)).astype(np.int64)
        img = np.clip(protos[lab] * 0.6 + rng.normal(60, 40, (n, 28, 28, 1)), 0, 255).astype(np.uint8)
        return torch.as_tensor(img).cuda(), torch.as_tensor(lab).cuda()

    train = BinaryNetEngine(binarynet_layers(), (28, 28, 1), 64, 10)
    train.init_glorot(0)
    val = BinaryNetEngine(binarynet_layers(), (28, 28, 1), 96, 10, share_state_with=train)
    assert val.params.data_ptr() == train.params.data_ptr() and val.moving_var.data_ptr() == train.moving_var.data_ptr()
    vimg, vlab = make(96)
    losses = []
    for epoch in range(3):
        for _ in range(8):
            train.train_step(*make(64), None, pad_size=32, seed=1)
        val.eval_step(vimg, vlab)
        torch.cuda.synchronize()
        losses.append(val.last_loss())
    assert val.step_count == train.step_count == 24
    assert len({round(v, 6) for v in losses}) == 3, losses          # validation follows the training
    # the same inference on a fresh engine loaded from a checkpoint written through the VAL engine
    ck = str(tmp_path / "shared.npz")
    val.save_checkpoint(ck)
    fresh = BinaryNetEngine(binarynet_layers(), (28, 28, 1), 96, 10)
    fresh.load_checkpoint(ck)
    fresh.eval_step(vimg, vlab)
    torch.cuda.synchronize()
    assert fresh.last_loss() == losses[-1] and fresh.step_count == 24
    for k, a in train.get_params().items():
        assert np.array_equal(a, fresh.get_params()[k])
    m_t, _ = train.get_adam_state()
    m_f, _ = fresh.get_adam_state()
    assert all(np.array_equal(m_t[k], m_f[k]) for k in m_t)


def test_cuda_graph_replay_matches_per_kernel_launches(native_lib):
    """capture_train_step / train_step_graph: same trajectory as ordinary steps (same sample
    indices, same counter-based augmentation, same Adam step sizes)."""
    from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers

    B, n = 32, 96
    rng = np.random.default_rng(2)
    pool = torch.as_tensor(rng.integers(0, 256, (n, 28, 28, 1), dtype=np.uint8)).cuda()
    labs = torch.as_tensor(rng.integers(0, 10, (n,)).astype(np.int64)).cuda()
    idx = [torch.as_tensor(rng.permutation(n)[:B].astype(np.int32)).cuda() for _ in range(4)]
    a = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    b = BinaryNetEngine(binarynet_layers(), (28, 28, 1), B, 10)
    a.init_glorot(0)
    b.init_glorot(0)
    ref = []
    for k in range(4):
        a.train_step(pool, labs, idx[k], pad_size=32, seed=5)
        ref.append(a.last_loss())
    b.train_step(pool, labs, idx[0], pad_size=32, seed=5)
    b.capture_train_step(pool, labs, 32, 5)
    losses = []
    for k in range(1, 4):
        b.train_step_graph(idx[k])
        losses.append(b.last_loss())
    torch.cuda.synchronize()
    assert a.step_count == b.step_count == 4
    assert np.array_equal(a.x0.cpu().numpy(), b.x0.cpu().numpy())          # same augmentation stream
    # the first replay follows one identical update: tight; later steps have amplified the float
    # summation-order differences between the two schedules (sign flips of near-zero weights)
    assert abs(ref[1] - losses[0]) <= 1e-4 * abs(ref[1]), (ref, losses)
    assert abs(ref[3] - losses[2]) <= 3e-2 * abs(ref[3]), (ref, losses)
    pa, pb = a.get_params(), b.get_params()
    for k in pa:
        # float atomics in wgrad make runs differ in the last bits; near-zero gradients then move a
        # few weights by +-lr in the other direction -- bound the fraction, as between any two runs
        assert (np.abs(pa[k] - pb[k]) > 1e-3).mean() < 2e-2, k
