"""Training step time on one GPU (config 5 of SURVEY §8d: B = 128 patches, 9x9 / 9x209, 201-way soft labels)
against the CPU restatement (torch autograd, float32, all host cores) on the same batch."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import train_oracle as to                                  # noqa: E402
from stereo_matching_b200 import _lib, train                           # noqa: E402

B, D = 128, 201
ctx = _lib.Context(0)
tr = _lib.Trainer(ctx, _lib.NET_WIN9, B, D)
params, stats = train.init_state("win9_dep9", 0)
tr.set_state(params, stats)
data = train.SyntheticPatches(2, 9, D, B, seed=1)
batches = [data.next_batch() for _ in range(4)]
for i in range(3):
    tr.forward_backward(*batches[i % 4])
    tr.apply(1e-3)
n = 20
ctx.profile_enable(False)
l0 = ctx.launch_count
t0 = time.perf_counter()
for i in range(n):
    loss, correct, _ = tr.forward_backward(*batches[i % 4])
    tr.apply(train.learning_rate(i + 4))
gpu_s = (time.perf_counter() - t0) / n
launches = (ctx.launch_count - l0) / n
import torch                                                           # noqa: E402
torch.set_num_threads(os.cpu_count() or 1)
st = tr.get_state()
t0 = time.perf_counter()
to.forward_backward(st["params"].astype(np.float32), st["stats"].copy(), *batches[0], dtype=torch.float32)
cpu_s = time.perf_counter() - t0
print(json.dumps({"workload": f"patch training step, base_model_ws_9, B={B}, D={D} (host patches in, loss out)",
                  "gpu_ms_per_step": gpu_s * 1e3, "gpu_steps_per_s": 1 / gpu_s, "patches_per_s": B / gpu_s,
                  "kernel_launches_per_step": launches, "last_loss": loss,
                  "cpu_oracle_ms_per_step": cpu_s * 1e3, "cpu_cores": os.cpu_count(),
                  "algorithmic_gflop_per_step": 3 * 2 * B * 64 * ((49 + 1449) * 27 + (25 + 9 + 1 + 1025 + 609 + 201) * 576) / 1e9}))
