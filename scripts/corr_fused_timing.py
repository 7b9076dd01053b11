"""Fused correlation + WTA on full-size feature maps: kernel-class time per pair, batch of 1 and of 16 pairs.
`python scripts/corr_fused_timing.py [D]`; STEREO_CORR_SWEEPS=2 times one sweep per volume (corr_sweep.cu)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import _lib          # noqa: E402

H, W, D = 375, 1242, int(sys.argv[1]) if len(sys.argv) > 1 else 228
ctx = _lib.Context(0)
dev = torch.device("cuda", 0)
out = {"D": D, "sweeps": 2 if os.environ.get("STEREO_CORR_SWEEPS") == "2" else 1}
for n in (1, 16):
    torch.manual_seed(n)
    fl = torch.randn((n, H, W, 64), dtype=torch.float32, device=dev)
    fr = torch.randn((n, H, W, 64), dtype=torch.float32, device=dev)
    ld = torch.empty((n, H, W), dtype=torch.uint8, device=dev)
    rd = torch.empty_like(ld)
    torch.cuda.synchronize()
    fn = lambda: ctx.corr_wta_dev(fl.data_ptr(), fr.data_ptr(), n, H, W, 64, D, ld.data_ptr(), rd.data_ptr(), _lib.PREC_BF16X3)
    fn()
    ctx.sync()
    ctx.profile_enable(True)
    reps = 10
    for _ in range(reps):
        fn()
    ctx.sync()
    out[f"corr_us_per_pair_n{n}"] = ctx.profile_read()["corr"][0] / reps / n * 1e3
    ctx.profile_enable(False)
print(json.dumps(out))
