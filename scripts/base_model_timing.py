"""Times base_model (win37) pairs at KITTI size: fp32 CUDA-core trunk vs the tensor-core trunk (conv_tcg)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from stereo_matching_b200 import _lib, synth  # noqa: E402
from stereo_matching_b200.models.model import base_model  # noqa: E402


def main():
    m = base_model((None, None, 3))
    m.set_layers(synth.make_base_model_weights())
    ctx = m.context
    D, h, w = 228, 375, 1242
    l, r, _ = synth.make_pair(7, h, w, D)
    out = {}
    for name, prec in (("fp32", _lib.PREC_FP32), ("bf16x3", _lib.PREC_BF16X3), ("bf16", _lib.PREC_BF16)):
        for _ in range(2):
            ctx.pairs(_lib.NET_WIN37, l, r, D, precision=prec)
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            ctx.pairs(_lib.NET_WIN37, l, r, D, precision=prec)
        out[name + "_ms_per_pair"] = (time.perf_counter() - t0) / reps * 1e3
        ctx.profile_enable(True)
        ctx.pairs(_lib.NET_WIN37, l, r, D, precision=prec)
        out[name + "_kernel_ms"] = {k: round(v[0], 4) for k, v in ctx.profile_read().items() if v[1]}
        ctx.profile_enable(False)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
