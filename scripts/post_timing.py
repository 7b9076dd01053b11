"""Per-kernel-class CUDA-event timing of the post chain on one full-size pair."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import _lib, checkpoint, synth          # noqa: E402
from stereo_matching_b200.models.model import base_model_ws_9     # noqa: E402
from stereo_matching_b200.metric import three_pixel_error         # noqa: E402

H, W, D = 375, 1242, int(sys.argv[1]) if len(sys.argv) > 1 else 228
m = base_model_ws_9((None, None, 3))
m.set_layers(checkpoint.load_net_weights(os.path.join(ROOT, "tests", "golden", "new_checkpoint")))
ctx = m.context
l, r, gt = synth.make_pair(1000, H, W, D)
out = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=True)
res = {}
for rep in range(2):
    ctx.profile_enable(True)
    st = ctx.post(out["left_cost"][0], out["right_cost"][0], out["left_disp"][0], out["right_disp"][0])
    res = ctx.profile_read()
    ctx.profile_enable(False)
vol_bytes = H * W * D * 4
line = {"workload": f"post chain, one {H}x{W} pair, D={D}", "kernel_ms": {k: v[0] for k, v in res.items() if v[1]},
        "launches": {k: v[1] for k, v in res.items() if v[1]},
        "err_wta": three_pixel_error(out["left_disp"][0].astype(np.int64), gt),
        "err_post": three_pixel_error(st["nyu_post"].astype(np.int64), gt)}
cb = res["cbca"][0] / 4 / 1e3
line["cbca_gbs_algorithmic"] = 2 * vol_bytes / cb / 1e9 if cb else None
sg = res["sgm"][0] / 2 / 1e3
line["sgm_gbs_algorithmic"] = 2 * vol_bytes / sg / 1e9 if sg else None
print(json.dumps(line))
