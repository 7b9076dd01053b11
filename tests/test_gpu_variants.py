"""Every selectable kernel variant of the fused path against the oracle.

The library picks its kernels once per process from environment variables (STEREO_CORR_KERNEL,
STEREO_CONV5_KERNEL, STEREO_PAIR_LANES; see DESIGN.md §4 - one predecessor per kernel is kept as a cross-check),
so each variant runs in
its own interpreter: a batch of three synthetic pairs through `stereo_pairs` (fused correlation + WTA,
LR check), maps written to an .npz.  Gates as in test_gpu_inference.py: >= 99.9 % of the WTA pixels
equal to the float64 oracle's, LR labels consistent with the maps.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import post_oracle as po
from oracle import stereo_oracle as so
from stereo_matching_b200 import synth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
H, W, D, N = 96, 400, 64, 3
WTA_AGREE = 0.999

CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib, checkpoint, synth
from stereo_matching_b200.models.model import base_model_ws_9
H, W, D, N = (int(v) for v in sys.argv[3:7])
m = base_model_ws_9((None, None, 3))
m.set_layers(checkpoint.load_net_weights(os.path.join(sys.argv[1], "tests", "golden", "new_checkpoint")))
pairs = [synth.make_pair(2000 + i, H, W, D) for i in range(N)]
left = np.stack([p[0] for p in pairs]); right = np.stack([p[1] for p in pairs])
out = m.context.pairs(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3, volumes=False, lr_check=True)
np.savez(sys.argv[2], left=out["left_disp"], right=out["right_disp"], lr=out["outlier"])
"""

VARIANTS = [
    {},                                       # defaults: conv_tc3, conv1_tc, corr_both, two lanes
    {"STEREO_CORR_KERNEL": "1"},              # per-tile correlation (corr_tc.cu)
    {"STEREO_CORR_SWEEPS": "2"},              # one sweep per volume (corr_sweep.cu) instead of one for both maps
    {"STEREO_PAIR_LANES": "1"},
    {"STEREO_PAIR_LANES": "4"},
]


@pytest.fixture(scope="module")
def reference_maps(ws9_layers):
    refs = []
    for i in range(N):
        l, r, _ = synth.make_pair(2000 + i, H, W, D)
        refs.append(so.infer_pair(l, r, ws9_layers, "win9_dep9", 9, D, np.float64))
    return refs


@pytest.mark.parametrize("env", VARIANTS, ids=lambda e: ",".join(f"{k[7:]}={v}" for k, v in e.items()) or "default")
def test_variant_matches_oracle(env, reference_maps, tmp_path):
    out = tmp_path / "maps.npz"
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", CHILD, ROOT, str(out), str(H), str(W), str(D), str(N)], env=e,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = np.load(out)
    for i, ref in enumerate(reference_maps):
        assert (got["left"][i] == ref["left_disp"]).mean() >= WTA_AGREE
        assert (got["right"][i] == ref["right_disp"]).mean() >= WTA_AGREE
        # the labels are an exact integer function of the two maps the same call returned
        lab = po.outlier_detection(got["left"][i].astype(np.float32), got["right"][i].astype(np.float32), D)
        assert np.array_equal(got["lr"][i], lab.astype(np.uint8))


VOL_CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib, checkpoint, synth
from stereo_matching_b200.models.model import base_model_ws_9
m = base_model_ws_9((None, None, 3))
m.set_layers(checkpoint.load_net_weights(os.path.join(sys.argv[1], "tests", "golden", "new_checkpoint")))
out = {}
for tag, (h, w, D) in {"a": (40, 300, 64), "b": (9, 1242, 228), "c": (13, 130, 256), "d": (5, 97, 4)}.items():
    l, r, _ = synth.make_pair(2500 + h, h, w, D)
    o = m.context.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    for k in ("left_cost", "right_cost", "left_disp", "right_disp"):
        out[tag + "_" + k] = o[k][0]
    lv, rv = m.context.cost_volume(np.random.default_rng(h).standard_normal((1, h, w, 64)).astype(np.float32),
                                   np.random.default_rng(w).standard_normal((1, h, w, 64)).astype(np.float32), D, _lib.PREC_BF16X3)
    out[tag + "_lv"], out[tag + "_rv"] = lv[0], rv[0]
np.savez(sys.argv[2], **out)
"""


def test_materialising_store_paths_agree_bit_for_bit(tmp_path):
    """The materialising correlation writes aligned volumes (D % 4 == 0) through one bulk copy per 32-pixel group of a
    shared-memory tile; STEREO_CORR_STORE=1 keeps the per-row stores.  The same accumulators leave by two routes: the
    volumes (with and without the fused WTA) and the maps must be equal bit for bit - image widths that end inside a
    32-pixel group, the widest range (256) and a tiny one."""
    outs = {}
    for name, env in (("bulk", {}), ("rows", {"STEREO_CORR_STORE": "1"})):
        e = dict(os.environ)
        e.update(env)
        out = tmp_path / (name + ".npz")
        r = subprocess.run([sys.executable, "-c", VOL_CHILD, ROOT, str(out)], env=e, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[name] = np.load(out)
    assert sorted(outs["bulk"].files) == sorted(outs["rows"].files) and len(outs["bulk"].files) == 24
    for k in outs["bulk"].files:
        assert np.array_equal(outs["bulk"][k], outs["rows"][k]), k
    lv = outs["bulk"]["b_left_cost"]
    assert np.array_equal(outs["bulk"]["b_left_disp"], np.argmax(lv, axis=2).astype(np.uint8))


PAD_CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib
ctx = _lib.Context(0)
out = {}
for D, h, w in ((201, 41, 263), (33, 30, 100), (255, 20, 70), (5, 12, 40)):
    rng = np.random.default_rng(D)
    lv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    rv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    ld, rd = ctx.wta(lv)[0], ctx.wta(rv)[0]
    for paths in (4, 8):
        p = _lib.default_post_params()
        p.sgm_paths = paths
        st = ctx.post(lv, rv, ld, rd, p)
        for k, v in st.items():
            out["%d_%d_%s" % (D, paths, k)] = v
np.savez(sys.argv[2], **out)
"""


def test_padded_sgm_rows_give_the_same_chain_bit_for_bit(tmp_path):
    """Disparity ranges that are not a multiple of 4 (the reference's default 201): the chain pads the rows of the SGM
    input to a multiple of 4 floats with "no such disparity" and runs the aligned SGM kernels on it (default) - every
    stage map and the final float map must equal the chain on unpadded rows (STEREO_POST_PAD=0, the unaligned code
    path of the same kernels) bit for bit, with 4 and with 8 paths."""
    outs = {}
    for name, env in (("padded", {}), ("plain", {"STEREO_POST_PAD": "0"})):
        e = dict(os.environ)
        e.update(env)
        out = tmp_path / (name + ".npz")
        r = subprocess.run([sys.executable, "-c", PAD_CHILD, ROOT, str(out)], env=e, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[name] = np.load(out)
    assert sorted(outs["padded"].files) == sorted(outs["plain"].files) and len(outs["plain"].files) == 4 * 2 * 9
    for k in outs["plain"].files:
        assert np.array_equal(outs["padded"][k], outs["plain"][k]), k


CBCA_CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib
ctx = _lib.Context(0)
out = {}
for D, h, w in ((228, 150, 61), (64, 70, 130), (4, 9, 5), (256, 5, 37)):
    rng = np.random.default_rng(D)
    # piecewise-constant guides: arms of every length; two iterations
    gl = np.repeat(np.repeat(rng.standard_normal((h // 3 + 1, w // 5 + 1)), 3, 0), 5, 1)[:h, :w].astype(np.float32) * 0.05
    gr = np.roll(gl, -3, axis=1)
    lv = rng.standard_normal((h, w, D)).astype(np.float32)
    rv = rng.standard_normal((h, w, D)).astype(np.float32)
    out["l_%d" % D], out["r_%d" % D] = ctx.nyu_cross_agg(gl, gr, lv, rv, 2)
np.savez(sys.argv[2], **out)
"""


def test_cbca_kernel_forms_agree_bit_for_bit(tmp_path):
    """The fused CBCA kernel in its default form (software-pipelined row loads) against the round-1 form
    (STEREO_CBCA_KERNEL=1: every row loaded and used at once): the same additions in the same order, so the bits must
    agree; bands of every length, images narrower than a block's window.  The default is checked against the C
    restatement in test_gpu_post.py."""
    maps = {}
    for name, env in (("pipelined", {}), ("round1", {"STEREO_CBCA_KERNEL": "1"})):
        e = dict(os.environ)
        e.update(env)
        out = tmp_path / (name + ".npz")
        r = subprocess.run([sys.executable, "-c", CBCA_CHILD, ROOT, str(out)], env=e, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        maps[name] = np.load(out)
    assert sorted(maps["pipelined"].files) == sorted(maps["round1"].files) and len(maps["round1"].files) == 8
    for k in maps["round1"].files:
        assert np.array_equal(maps["pipelined"][k], maps["round1"][k]), k


BASE_CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib, synth
from stereo_matching_b200.models.model import base_model
m = base_model((None, None, 3))
m.set_layers(synth.make_base_model_weights())
x = np.random.default_rng(11).standard_normal((2, 52, 300, 3)).astype(np.float32)
np.savez(sys.argv[2], pair=m.context.trunk(_lib.NET_WIN37, x, _lib.PREC_BF16X3),
         single=m.context.trunk(_lib.NET_WIN37, x[:1], _lib.PREC_BF16X3))
"""


def test_base_model_conv_kernel_families_agree(tmp_path):
    """STEREO_CONV5_KERNEL=1 (conv_tcg.cu, one MMA per tap) against the default conv_tcr.cu (each input row
    once): same products, different order of the fp32 accumulation, so equal within the tolerance of the split;
    both are checked against the float64 oracle."""
    from oracle import stereo_oracle as so
    outs = {}
    for name, env in (("row_once", {}), ("per_tap", {"STEREO_CONV5_KERNEL": "1"})):
        e = dict(os.environ)
        e.update(env)
        out = tmp_path / (name + ".npz")
        r = subprocess.run([sys.executable, "-c", BASE_CHILD, ROOT, str(out)], env=e, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[name] = np.load(out)
    a, b = outs["row_once"], outs["per_tap"]
    assert np.abs(a["pair"] - b["pair"]).max() / np.abs(b["pair"]).max() < 2e-5
    assert np.array_equal(a["single"][0], a["pair"][0])          # the batch size does not change an image's result
    layers = synth.make_base_model_weights()
    x = np.random.default_rng(11).standard_normal((2, 52, 300, 3)).astype(np.float32)
    ref = so.trunk_forward(x, layers, "win37_dep9", np.float64)
    for k in ("row_once", "per_tap"):
        assert np.abs(outs[k]["pair"] - ref).max() / np.abs(ref).max() < 6e-5


BOTH_CHILD = r"""
import sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from stereo_matching_b200 import _lib
ctx = _lib.Context(0)
out = {}
cases = {"kitti": (2, 7, 1242, 228), "one_block": (2, 5, 128, 64), "narrow": (1, 4, 97, 4), "two_blocks": (1, 3, 256, 256),
         "ragged": (2, 6, 130, 201), "d1": (1, 5, 700, 1), "wide": (1, 2, 1500, 256), "tiny": (1, 3, 5, 7)}
for tag, (n, h, w, D) in cases.items():
    for kind in ("ties", "normal"):
        rng = np.random.default_rng(h * 1000 + w)
        if kind == "ties":       # small integers: exact products, many equal scores -> the tie rule decides
            l = rng.integers(-1, 2, (n, h, w, 64)).astype(np.float32)
            r = rng.integers(-1, 2, (n, h, w, 64)).astype(np.float32)
            l[..., 8:] = 0; r[..., 8:] = 0
        else:
            l = rng.standard_normal((n, h, w, 64)).astype(np.float32)
            r = rng.standard_normal((n, h, w, 64)).astype(np.float32)
        for prec in (_lib.PREC_BF16X3, _lib.PREC_BF16):
            ld, rd = ctx.corr_wta(l, r, D, prec)
            out[f"{tag}_{kind}_{prec}_l"], out[f"{tag}_{kind}_{prec}_r"] = ld, rd
            if len(sys.argv) > 3:
                lv, rv = ctx.cost_volume(l, r, D, prec)
                out[f"{tag}_{kind}_{prec}_la"] = lv.argmax(-1).astype(np.uint8)
                out[f"{tag}_{kind}_{prec}_ra"] = rv.argmax(-1).astype(np.uint8)
np.savez(sys.argv[2], **out)
"""


def test_one_sweep_maps_equal_the_arg_max_of_the_volumes(tmp_path):
    """corr_both.cu reads the accumulators of the LEFT sweep along rows (left map) and along columns (right map).
    Both maps must equal numpy's argmax (first maximum) of the volumes the materialising kernel writes from the same
    accumulators, and the maps of the two-sweep kernel (STEREO_CORR_SWEEPS=2), bit for bit: images narrower than a
    block, ending on and just behind a block edge, D = 1 and D = 256, features with many exactly equal scores, and
    the right map's zero-filled slots x + d >= w."""
    runs = {}
    for tag, env, extra in (("both", {}, ["argmax"]), ("two", {"STEREO_CORR_SWEEPS": "2"}, [])):
        out = tmp_path / f"{tag}.npz"
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, "-c", BOTH_CHILD, ROOT, str(out)] + extra, env=e, capture_output=True,
                           text=True, timeout=900)
        assert r.returncode == 0, r.stderr[-2000:]
        runs[tag] = np.load(out)
    keys = [k for k in runs["two"].files]
    assert len(keys) == 8 * 2 * 2 * 2
    for k in keys:
        assert np.array_equal(runs["both"][k], runs["two"][k]), k
        assert np.array_equal(runs["both"][k], runs["both"][k + "a"]), k
