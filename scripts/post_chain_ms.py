"""Device-resident timing of the post chain and of the whole pipeline on one full-size pair (bench.py's secondary
numbers without the rest of the bench): `python scripts/post_chain_ms.py [D]`."""
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
m = base_model_ws_9((None, None, 3))
m.set_layers(checkpoint.load_net_weights(os.path.join(ROOT, "tests", "golden", "new_checkpoint")))
ctx = m.context
dev = torch.device("cuda:0")
l, r, gt = synth.make_pair(1000, H, W, D)
d_left, d_right = torch.from_numpy(l).to(dev), torch.from_numpy(r).to(dev)
d_lv = torch.empty((H, W, D), dtype=torch.float32, device=dev)
d_rv = torch.empty_like(d_lv)
d_ld = torch.empty((H, W), dtype=torch.uint8, device=dev)
d_rd = torch.empty_like(d_ld)
d_stages = torch.empty((8, H, W), dtype=torch.uint8, device=dev)
stream = torch.cuda.Stream()
ctx.set_stream(stream.cuda_stream)
torch.cuda.synchronize()
prec = _lib.PREC_BF16X3
ctx.pairs_dev(m.net, d_left.data_ptr(), d_right.data_ptr(), 1, H, W, D, prec, d_ld.data_ptr(), d_rd.data_ptr(), None,
              d_lv.data_ptr(), d_rv.data_ptr())
ptrs = [d_stages[i].data_ptr() for i in range(8)]


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


post_ms = timed(lambda: ctx.post_dev(d_lv.data_ptr(), d_rv.data_ptr(), d_ld.data_ptr(), d_rd.data_ptr(), H, W, D, ptrs), 5)
d_np = torch.empty((4, H, W), dtype=torch.uint8, device=dev)
pp = _lib.default_post_params()
full_ms = timed(lambda: ctx.pairs_post_dev(m.net, d_left.data_ptr(), d_right.data_ptr(), 1, H, W, D, prec, pp,
                                           d_np.data_ptr()), 5)
d_l4, d_r4 = d_left.unsqueeze(0).repeat(4, 1, 1, 1).contiguous(), d_right.unsqueeze(0).repeat(4, 1, 1, 1).contiguous()
full4_ms = timed(lambda: ctx.pairs_post_dev(m.net, d_l4.data_ptr(), d_r4.data_ptr(), 4, H, W, D, prec, pp,
                                            d_np.data_ptr()), 3) / 4
# kernel classes of one chain, every kernel alone on the GPU (CUDA events around each launch group)
ctx.profile_enable(True)
ctx.post_dev(d_lv.data_ptr(), d_rv.data_ptr(), d_ld.data_ptr(), d_rd.data_ptr(), H, W, D, ptrs)
torch.cuda.synchronize()
classes = {k: round(v[0], 4) for k, v in ctx.profile_read().items() if v[1]}
ctx.profile_enable(False)
p8 = _lib.default_post_params()
p8.sgm_paths = 8          # not the reference (4-path): BASELINE.json configs[3]'s 8-path chain, labelled apart
post8_ms = timed(lambda: ctx.post_dev(d_lv.data_ptr(), d_rv.data_ptr(), d_ld.data_ptr(), d_rd.data_ptr(), H, W, D, ptrs,
                                      params=p8), 3)
print(json.dumps({"D": D, "post_chain_ms": post_ms, "post_chain_8path_ms": post8_ms, "kernel_class_ms": classes, "full_pipeline_ms": full_ms, "full_pipeline_ms_batch4": full4_ms}))
