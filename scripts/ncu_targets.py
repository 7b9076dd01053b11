"""The kernels of every configuration once, for ncu: `ncu --set full -k regex:... python scripts/ncu_targets.py`.

Two ws9 pairs through the fused path, one with the volumes materialised, one base_model pair, one post chain
(all stage images, both sides on one stream).  Full size (375x1242, D=228)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("STEREO_POST_SERIAL", "1")
os.environ.setdefault("STEREO_PAIR_LANES", "1")
from stereo_matching_b200 import _lib, checkpoint, synth          # noqa: E402
from stereo_matching_b200.models.model import base_model, base_model_ws_9     # noqa: E402

H, W, D = 375, 1242, int(sys.argv[1]) if len(sys.argv) > 1 else 228
dev = torch.device("cuda:0")
m = base_model_ws_9((None, None, 3))
m.set_layers(checkpoint.load_net_weights(os.path.join(ROOT, "tests", "golden", "new_checkpoint")))
ctx = m.context
pairs = [synth.make_pair(1000 + i, H, W, D) for i in range(2)]
d_left = torch.from_numpy(np.stack([p[0] for p in pairs])).to(dev)
d_right = torch.from_numpy(np.stack([p[1] for p in pairs])).to(dev)
d_ld = torch.empty((2, H, W), dtype=torch.uint8, device=dev)
d_rd, d_o = torch.empty_like(d_ld), torch.empty_like(d_ld)
d_lv = torch.empty((H, W, D), dtype=torch.float32, device=dev)
d_rv = torch.empty_like(d_lv)
prec = _lib.PREC_BF16X3
ctx.pairs_dev(m.net, d_left.data_ptr(), d_right.data_ptr(), 2, H, W, D, prec, d_ld.data_ptr(), d_rd.data_ptr(), d_o.data_ptr())
ctx.pairs_dev(m.net, d_left.data_ptr(), d_right.data_ptr(), 1, H, W, D, prec, d_ld.data_ptr(), d_rd.data_ptr(), None,
              d_lv.data_ptr(), d_rv.data_ptr())
ctx.sync()
d_st = torch.empty((8, H, W), dtype=torch.uint8, device=dev)
ctx.post_dev(d_lv.data_ptr(), d_rv.data_ptr(), d_ld.data_ptr(), d_rd.data_ptr(), H, W, D, [d_st[i].data_ptr() for i in range(8)])
ctx.sync()
bm = base_model((None, None, 3))
bm.set_layers(synth.make_base_model_weights())
bm.context.pairs_dev(bm.net, d_left.data_ptr(), d_right.data_ptr(), 1, H, W, D, prec, d_ld.data_ptr(), d_rd.data_ptr(), d_o.data_ptr())
bm.context.sync()
print("ok")
