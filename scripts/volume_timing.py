"""Materialising correlation + standalone WTA on one full-size pair (for ncu captures)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import _lib, checkpoint, synth          # noqa: E402
from stereo_matching_b200.models.model import base_model_ws_9     # noqa: E402

H, W, D = 375, 1242, 228
m = base_model_ws_9((None, None, 3), precision=_lib.PREC_BF16X3)
m.set_layers(checkpoint.load_net_weights(os.path.join(ROOT, "tests", "golden", "new_checkpoint")))
ctx = m.context
l, r, gt = synth.make_pair(1000, H, W, D)
dev = torch.device("cuda", 0)
d_l = torch.from_numpy(l).to(dev)
d_r = torch.from_numpy(r).to(dev)
d_lv = torch.empty((H, W, D), dtype=torch.float32, device=dev)
d_rv = torch.empty((H, W, D), dtype=torch.float32, device=dev)
d_ld = torch.empty((H, W), dtype=torch.uint8, device=dev)
d_rd = torch.empty((H, W), dtype=torch.uint8, device=dev)
for rep in range(3):
    ctx.pairs_dev(_lib.NET_WIN9, d_l.data_ptr(), d_r.data_ptr(), 1, H, W, D, _lib.PREC_BF16X3,
                  d_ld.data_ptr(), d_rd.data_ptr(), None, d_lv.data_ptr(), d_rv.data_ptr())
    ctx.wta_dev(d_lv.data_ptr(), 1, H, W, D, False, d_ld.data_ptr())
ctx.sync()
print("ok", float(d_lv.abs().max()))
