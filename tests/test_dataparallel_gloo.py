"""N>1 host logic on CPU: two gloo ranks shard a global batch, all-reduce BatchNorm sums and
gradients through zookeeper_b200.dataparallel, and must reproduce the single-process result
("N-GPU result == 1-GPU result at the same global batch")."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from zookeeper_b200.dataparallel import DataParallel

        dp = DataParallel()
        assert (dp.rank, dp.world) == (rank, world)
        rng = np.random.default_rng(0)                    # same stream on both ranks
        B, C, NC = 16, 8, 10
        y = torch.tensor(rng.integers(-50, 50, (B, C)).astype(np.float64))     # pre-BN integers
        w = torch.tensor(rng.standard_normal((C, NC)))
        labels = torch.tensor(rng.integers(0, NC, (B,)))
        idx = dp.shard(np.arange(B))
        assert len(idx) == B // world and set(idx.tolist()) == set(range(rank, B, world))
        yl, ll = y[idx], labels[idx]

        # SyncBN forward: local sums -> all-reduce -> global mean / biased variance
        s, ss = yl.sum(0), (yl * yl).sum(0)
        dp.reduce_bn_sums_(s, ss)
        n = dp.global_count(len(idx), sync=True)
        mean, var = s / n, ss / n - (s / n) ** 2
        a = ((yl - mean) * torch.rsqrt(var + 1e-3)).requires_grad_(True)

        # loss gradient scaled by 1/(local batch * world), gradients summed across ranks
        wl = w.clone().requires_grad_(True)
        loss = torch.nn.functional.cross_entropy(a @ wl, ll, reduction="sum") / len(idx) * dp.loss_scale
        loss.backward()
        flat = wl.grad.reshape(-1).clone()
        dp.reduce_gradients_(flat)
        total_loss = loss.detach().clone().reshape(1)
        dp.all_reduce_sum_(total_loss)
        if rank == 0:
            np.savez(os.path.join(out_dir, "dp.npz"), mean=mean.numpy(), var=var.numpy(), grad=flat.numpy(),
                     loss=total_loss.numpy())
        with pytest.raises(ValueError, match="not divisible"):
            dp.shard(np.arange(B + 1))

        # bucketed gradient all-reduce as the engine issues it: the buckets of the flat arena are
        # started one by one (async handles) and joined at the end -- same sums as one collective
        from zookeeper_b200.engine import gradient_buckets

        k_off = np.array([0, 40, 1000, 3000, 3016])
        arena = torch.arange(3016 + 24, dtype=torch.float64) * (rank + 1)
        whole = arena.clone()
        dp.all_reduce_sum_(whole)
        works = [dp.all_reduce_sum_async_(arena[lo:hi]) for _, lo, hi in gradient_buckets(k_off, 24, min_bytes=4000)]
        for wk in works:
            wk.wait()
        assert torch.equal(arena, whole)
    finally:
        dist.destroy_process_group()


def test_two_rank_shards_reproduce_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "dp.npz")

    rng = np.random.default_rng(0)
    B, C, NC = 16, 8, 10
    y = torch.tensor(rng.integers(-50, 50, (B, C)).astype(np.float64))
    w = torch.tensor(rng.standard_normal((C, NC))).requires_grad_(True)
    labels = torch.tensor(rng.integers(0, NC, (B,)))
    mean, var = y.mean(0), y.var(0, unbiased=False)
    a = (y - mean) * torch.rsqrt(var + 1e-3)
    loss = torch.nn.functional.cross_entropy(a @ w, labels)
    loss.backward()
    np.testing.assert_allclose(got["mean"], mean.numpy(), rtol=1e-12)
    np.testing.assert_allclose(got["var"], var.numpy(), rtol=1e-10)
    np.testing.assert_allclose(got["loss"][0], float(loss), rtol=1e-12)
    np.testing.assert_allclose(got["grad"], w.grad.reshape(-1).numpy(), rtol=1e-10, atol=1e-14)


def test_gradient_buckets_tile_the_arena_in_backward_order():
    """Reverse-layer-order buckets: disjoint slices that cover [betas | kernels] exactly, dense
    layers first, the beta prefix and the first layers' kernels last."""
    sys.path.insert(0, ROOT)
    from zookeeper_b200.engine import binarynet_layers, gradient_buckets

    ls = binarynet_layers()
    ks = [l.ksize * l.ksize * l.cin * l.cout if l.kind == "conv" else l.cin * l.cout for l in ls]
    k_off = np.concatenate([[0], np.cumsum([(s + 3) // 4 * 4 for s in ks])])
    kbase = 3904
    b = gradient_buckets(k_off, kbase)
    assert [x[0] for x in b] == [6, 5, 0]                       # dense1..3 | conv6 | the rest + betas
    assert b[0][2] == kbase + k_off[-1] and b[-1][1] == 0
    assert all(b[i][1] == b[i + 1][2] for i in range(len(b) - 1))
    assert all((hi - lo) * 4 >= 8 << 20 for _, lo, hi in b[:-1])
    assert b[0][1] == kbase + k_off[6] and b[1][1] == kbase + k_off[5]
    # a small model is one bucket
    assert gradient_buckets(np.array([0, 100, 200]), 8) == [(0, 0, 208)]


def test_single_process_is_a_no_op():
    sys.path.insert(0, ROOT)
    from zookeeper_b200.dataparallel import DataParallel

    dp = DataParallel()
    assert (dp.rank, dp.world, dp.loss_scale) == (0, 1, 1.0)
    t = torch.arange(4.0)
    assert torch.equal(dp.reduce_gradients_(t.clone()), t)
    assert np.array_equal(dp.shard(np.arange(6)), np.arange(6))
    assert dp.global_count(10, sync=True) == 10.0
    assert dp.all_reduce_sum_async_(t) is None
