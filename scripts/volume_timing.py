"""Materialising correlation on one full-size pair: kernel-class time of the correlation kernel with and without the
fused WTA (`stereo_pairs_dev` with volumes / `stereo_cost_volume_dev`), CUDA events on the launching stream.
`python scripts/volume_timing.py [D]`; STEREO_CORR_STORE=1 times the per-row stores."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import _lib, checkpoint, synth          # noqa: E402
from stereo_matching_b200.models.model import base_model_ws_9     # noqa: E402

H, W, D = 375, 1242, int(sys.argv[1]) if len(sys.argv) > 1 else 228
m = base_model_ws_9((None, None, 3), precision=_lib.PREC_BF16X3)
m.set_layers(checkpoint.load_net_weights(os.path.join(ROOT, "tests", "golden", "new_checkpoint")))
ctx = m.context
l, r, gt = synth.make_pair(1000, H, W, D)
dev = torch.device("cuda", 0)
d_l, d_r = torch.from_numpy(l).to(dev), torch.from_numpy(r).to(dev)
d_lv = torch.empty((H, W, D), dtype=torch.float32, device=dev)
d_rv = torch.empty_like(d_lv)
d_ld = torch.empty((H, W), dtype=torch.uint8, device=dev)
d_rd = torch.empty_like(d_ld)
d_feat = torch.empty((2, H, W, 64), dtype=torch.float32, device=dev)
d_pad = torch.empty((2, H + 8, W + 8, 3), dtype=torch.float32, device=dev)
prec = _lib.PREC_BF16X3
ctx.normalize_pad_dev(d_l.data_ptr(), 1, H, W, 9, d_pad[0].data_ptr())
ctx.normalize_pad_dev(d_r.data_ptr(), 1, H, W, 9, d_pad[1].data_ptr())
ctx.trunk_dev(m.net, d_pad.data_ptr(), 2, H + 8, W + 8, d_feat.data_ptr(), prec)


def with_wta():
    ctx.pairs_dev(m.net, d_l.data_ptr(), d_r.data_ptr(), 1, H, W, D, prec, d_ld.data_ptr(), d_rd.data_ptr(), None,
                  d_lv.data_ptr(), d_rv.data_ptr())


def volume_only():
    ctx.cost_volume_dev(d_feat[0].data_ptr(), d_feat[1].data_ptr(), 1, H, W, 64, D, d_lv.data_ptr(), d_rv.data_ptr(), prec)


out = {"D": D, "store": "per-row" if os.environ.get("STEREO_CORR_STORE") == "1" else "bulk (aligned D) / per-row"}
reps = 10
for name, fn in (("corr_ms_volumes_and_wta", with_wta), ("corr_ms_volumes_only", volume_only)):
    fn()
    ctx.sync()
    ctx.profile_enable(True)
    for _ in range(reps):
        fn()
    ctx.sync()
    out[name] = ctx.profile_read()["corr"][0] / reps
    ctx.profile_enable(False)
bytes_ = 2 * 4.0 * H * W * 64 + 2 * 4.0 * H * W * D
out["volumes_only_gbs"] = bytes_ / (out["corr_ms_volumes_only"] / 1e3) / 1e9
out["volumes_and_wta_gbs"] = bytes_ / (out["corr_ms_volumes_and_wta"] / 1e3) / 1e9
print(json.dumps(out))
