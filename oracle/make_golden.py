"""Generate tests/golden/*.npz from the CPU oracle (run from the repo root:
``python -m oracle.make_golden`` for the BinaryNet / bit-level / preprocessing vectors,
``python -m oracle.make_golden zoo`` for the two larq-zoo models).  The vectors pin the oracle against
drift -- they are NOT TensorFlow outputs (parity unpinned, see oracle/binarynet_oracle.py)."""
import os
import sys

import numpy as np
import torch

from oracle import binarynet_oracle as O

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def small_net():
    # a reduced BinaryNet (filters 128 is the smallest width the CUDA engine accepts; keep the batch
    # tiny instead) -- same topology as examples/larq_experiment.py:59-103
    return O.binarynet_layers((28, 28, 1), 128, 1024, 10)


def main():
    os.makedirs(OUT, exist_ok=True)
    B = 4
    layers = small_net()
    params = O.glorot_uniform_params(layers, 0)
    rng = np.random.default_rng(2024)
    img = rng.integers(0, 256, (B, 28, 28, 1), dtype=np.uint8)
    lab = rng.integers(0, 10, (B,)).astype(np.int64)
    aug = O.augmentation_params(2024, 0, B, 32, 28, 28)
    x = O.preprocess(img, (28, 28), True, 32, aug)
    r = O.forward_backward(params, layers, x, lab, torch.float64, keep_acts=True)
    m0 = {k: np.zeros_like(v) for k, v in params.items()}
    p1, m1, v1 = O.adam_clip_step(params, r.grads, m0, m0, 1, dtype=np.float32)
    # keep the fixture small: checksums of big tensors, full values of small ones
    out = {
        "images": img, "labels": lab, "aug": aug, "x_sum": np.float64(x.astype(np.float64).sum()),
        "x_first_row": x[0, 0, :, 0], "loss": np.float64(r.loss), "logits": r.logits,
        "grad_beta8": r.grads["beta8"], "grad_kernel8": r.grads["kernel8"],
        "grad_kernel0": r.grads["kernel0"], "new_kernel0": p1["kernel0"], "new_beta8": p1["beta8"],
    }
    for i in range(len(layers)):
        out[f"y{i}_abs_sum"] = np.float64(np.abs(r.acts[f"y{i}"]).sum())
        out[f"grad_kernel{i}_norm"] = np.float64(np.linalg.norm(r.grads[f"kernel{i}"]))
        out[f"mean{i}"] = r.batch_stats[f"mean{i}"]
    np.savez_compressed(os.path.join(OUT, "binarynet_b4_step.npz"), **out)

    # bit-level: sign + pack + xnor-popcount
    a = rng.standard_normal((8, 128)).astype(np.float32)
    a[0, :4] = [0.0, -0.0, 1e-30, -1e-30]
    b = rng.standard_normal((6, 128)).astype(np.float32)
    pa, pb = O.sign_pack(a), O.sign_pack(b)
    np.savez_compressed(os.path.join(OUT, "bitops.npz"), a=a, b=b, pa=pa, pb=pb, y=O.xnor_popcount_gemm(pa, pb, 128))

    # preprocessing (eval mode with crop, training with explicit aug on a 3-channel image)
    img3 = rng.integers(0, 256, (3, 32, 32, 3), dtype=np.uint8)
    aug3 = np.array([[0, 8, 1], [8, 0, 0], [3, 5, 1]], dtype=np.int8)
    np.savez_compressed(os.path.join(OUT, "preprocess.npz"), img=img3, aug=aug3,
                        train=O.preprocess(img3, (32, 32), True, 40, aug3),
                        eval_crop=O.preprocess(img3, (28, 28), False),
                        eval_pad=O.preprocess(img3, (36, 36), False))
    print("wrote", sorted(os.listdir(OUT)))


def zoo_cases():
    """(name, run) of the zoo-model steps the fixture holds; tests/test_oracle_zoo_golden.py replays them."""
    from oracle import alexnet_oracle as A
    from oracle import bireal_oracle as R

    def alexnet():
        hw, B, NC, dense, seed = (96, 96), 4, 10, 256, 1
        p = A.init_params(NC, hw, dense_units=dense, seed=seed)
        rng = np.random.default_rng(seed + 50)
        img = rng.uniform(-1, 1, (B,) + hw + (3,)).astype(np.float32)
        lab = rng.integers(0, NC, (B,)).astype(np.int64)
        return A.forward_backward(p, img, lab, torch.float64, keep=True)

    def bireal():
        plan, hw, B, NC, seed = ("plain", "down", "plain"), (32, 32), 4, 10, 4
        p = R.init_params(plan, 64, NC, seed=seed, stem_ksize=7)
        rng = np.random.default_rng(seed + 100)
        img = rng.uniform(-1, 1, (B,) + hw + (3,)).astype(np.float32)
        lab = rng.integers(0, NC, (B,)).astype(np.int64)
        return R.forward_backward(p, plan, img, lab, torch.float64, keep=True)

    return (("alexnet", alexnet), ("bireal", bireal))


def zoo_summary(r):
    """What the fixture keeps of one step: loss, logits, every gradient's norm and sum, and the sum of
    absolute values of every kept activation (exact for the integer maps)."""
    out = {"loss": np.float64(r["loss"]), "logits": r["logits"].astype(np.float64)}
    for k, g in r["grads"].items():
        out[f"grad_norm_{k}"] = np.float64(np.linalg.norm(g))
        out[f"grad_sum_{k}"] = np.float64(g.astype(np.float64).sum())
    for k, a in r["acts"].items():
        if np.ndim(a) > 1:
            out[f"act_abs_sum_{k}"] = np.float64(np.abs(a.astype(np.float64)).sum())
    return out


def main_zoo():
    os.makedirs(OUT, exist_ok=True)
    out = {}
    for name, run in zoo_cases():
        for k, v in zoo_summary(run()).items():
            out[f"{name}/{k}"] = v
    np.savez_compressed(os.path.join(OUT, "zoo_steps.npz"), **out)
    print("wrote zoo_steps.npz with", len(out), "entries")


if __name__ == "__main__":
    main_zoo() if sys.argv[1:] == ["zoo"] else main()
