"""Data-parallel training check, run as `torchrun --nproc-per-node N tests/tools/train_dp_check.py`: every rank
computes gradients on its own batch, the flat device gradient array is averaged with one NCCL all-reduce
(through a zero-copy torch view of the library's buffer), every rank applies the same Adam step.  Rank 0 checks
that all ranks hold bit-identical parameters afterwards and that they match the CPU oracle's step on the
averaged gradient."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import train_oracle as to                                  # noqa: E402
from stereo_matching_b200 import _lib, train                           # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
B, D = 32, 41
ctx = _lib.Context(local)
tr = _lib.Trainer(ctx, _lib.NET_WIN9, B, D)
params, stats = train.init_state("win9_dep9", 11)
tr.set_state(params, stats)
gview = train.DeviceGradView(tr, local).tensor
rng = np.random.default_rng(1000 + rank)
steps, ok = 3, True
p64, m, v = params.astype(np.float64), np.zeros(params.size), np.zeros(params.size)
solid = np.ones(params.size, bool)      # entries whose averaged gradient is well above fp32 round-off at every step
for step in range(1, steps + 1):
    left = rng.standard_normal((B, 9, 9, 3)).astype(np.float32)
    right = rng.standard_normal((B, 9, 9 + D - 1, 3)).astype(np.float32)
    targets = train.soft_labels(B, D)
    tr.forward_backward(left, right, targets)
    train.average_gradients(gview, world)
    torch.cuda.synchronize()
    tr.apply(train.learning_rate(step))
    # oracle: gradients of every rank's batch on rank 0
    batches = [None] * world
    dist.all_gather_object(batches, (left, right))
    if rank == 0:
        g = np.mean([to.forward_backward(p64, stats.copy(), l, r, targets)[2] for l, r in batches], axis=0)
        solid &= np.abs(g) > 1e-2 * np.abs(g).max()
        p64, m, v = to.adam_step(p64, g, m, v, step, train.learning_rate(step))
mine = torch.tensor(tr.get_state()["params"], device=f"cuda:{local}")
allp = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(allp, mine)
if rank == 0:
    same = all(torch.equal(allp[0], q) for q in allp[1:])
    errs = np.abs(allp[0].cpu().numpy() - p64)[solid] / np.abs(p64).max()
    err = float(errs.max())
    print(json.dumps({"world": world, "steps": steps, "ranks_bit_identical": same, "max_rel_err_vs_oracle": err, "p999_rel_err_vs_oracle": float(np.quantile(errs, 0.999)),
                      "ok": bool(same and err < 5e-4)}))
dist.barrier()
dist.destroy_process_group()
