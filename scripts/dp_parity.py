"""N-GPU check (N = WORLD_SIZE ranks): a SyncBN data-parallel step at global batch Bg reproduces the
single-GPU step at batch Bg.  Rank 0 prints the gradient deviation and the
worst tensor, and appends the same lines to gpurun_out/dp_parity_n<N>.log when that directory exists."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from zookeeper_b200.engine import BinaryNetEngine, binarynet_layers  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
Bg = 64
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (Bg, 28, 28, 1), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (Bg,)).astype(np.int64)).cuda()
aug = torch.as_tensor(rng.integers(0, 2, (Bg, 3)).astype(np.int8)).cuda()

eng = BinaryNetEngine(binarynet_layers(), (28, 28, 1), Bg // world, 10, sync_bn=True)
eng.init_glorot(0)
sl = slice(rank, Bg, world)
eng.train_step(img[sl].contiguous(), lab[sl].contiguous(), None, pad_size=32, aug=aug[sl].contiguous(), apply_update=False)
torch.cuda.synchronize()
g_dp = eng.grads.clone()
loss_dp = torch.tensor([eng.last_loss()], device="cuda", dtype=torch.float64)
dist.all_reduce(loss_dp)
if rank == 0:
    single = BinaryNetEngine(binarynet_layers(), (28, 28, 1), Bg, 10, data_parallel=False)
    single.init_glorot(0)
    single.train_step(img, lab, None, pad_size=32, aug=aug, apply_update=False)
    torch.cuda.synchronize()
    rel = ((g_dp - single.grads).norm() / single.grads.norm()).item()
    lines = [f"DP({world}, SyncBN) vs single GPU at global batch {Bg}: grad rel err {rel:.3e}, "
             f"loss {loss_dp.item() / world:.6f} vs {single.last_loss():.6f}"]
    print(lines[0])
    gd = {k: v for k, v in zip(["all"], [None])}
    eng.grads.copy_(g_dp)
    a, b = eng.get_grads(), single.get_grads()
    worst = max((float(np.linalg.norm(a[k] - b[k]) / max(np.linalg.norm(b[k]), 1e-30)), k) for k in b)
    lines.append(f"worst tensor: {worst[1]} rel err {worst[0]:.3e}")
    # moving statistics: identical on every rank under SyncBN and equal to the single-GPU ones
    mv_a, mv_b = eng.get_moving(), single.get_moving()
    mv_err = max(float(np.abs(mv_a[k] - mv_b[k]).max() / max(np.abs(mv_b[k]).max(), 1e-30)) for k in mv_b)
    lines.append(f"moving statistics vs single GPU: max rel err {mv_err:.3e}")
    print("\n".join(lines[1:]))
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, f"dp_parity_n{world}.log"), "w") as f:
            f.write("\n".join(lines) + "\n")
    assert rel < 1e-4 and worst[0] < 1e-4 and mv_err < 1e-5
dist.barrier()
dist.destroy_process_group()
