"""GPU parity tests of the post chain (P0-P9) through the C ABI, against
oracle/post_oracle.{c,py}.  The CUDA kernels perform the same fp32 operations in
the same order as the restated original, so per-kernel results are required to
be bit-identical; the whole-volume std (a float64 reduction in a different
order) is allowed 1 ulp, and chain outputs that depend on it are compared at
1e-5 / 99.9 %."""
import os

import cv2
import numpy as np
import pytest

from oracle import post_oracle as po
from oracle import stereo_oracle as so
from stereo_matching_b200 import _lib, synth
from stereo_matching_b200.models.model import base_model_ws_9

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(ws9_layers):
    m = base_model_ws_9((None, None, 3))
    m.set_layers(ws9_layers)
    return m.context


@pytest.fixture(scope="module")
def scene(ctx):
    """One synthetic pair through the GPU inference path: volumes + WTA maps."""
    D = 48
    l, r, gt = synth.make_pair(1007, 70, 230, D)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=True)
    return {"D": D, "gt": gt, "lv": out["left_cost"][0], "rv": out["right_cost"][0],
            "ld": out["left_disp"][0], "rd": out["right_disp"][0]}


def test_guide_image(ctx, scene, golden_dir):
    for img in (scene["ld"], cv2.imread(os.path.join(golden_dir, "results", "left_disp", "left_000002_10.png"), 0)):
        ref = po.normalise_guide(img)
        out = ctx.guide_image(img)
        np.testing.assert_array_max_ulp(out, ref, maxulp=1)


def test_volume_std(ctx, scene):
    ref = po.volume_std(scene["lv"])
    out = ctx.volume_std(scene["lv"])
    np.testing.assert_array_max_ulp(np.float32(out), np.float32(ref), maxulp=1)


def exact_params():
    p = _lib.default_post_params()
    p.cbca_exact = 1          # the original's strictly sequential summation order
    return p


@pytest.mark.parametrize("iters", [1, 2])
def test_cross_agg_bit_exact(ctx, scene, iters):
    l_img, r_img = po.normalise_guide(scene["ld"]), po.normalise_guide(scene["rd"])
    lu = np.ascontiguousarray(np.transpose(scene["lv"], (2, 0, 1)))
    ru = np.ascontiguousarray(np.transpose(scene["rv"], (2, 0, 1)))
    lref, rref = po.nyu_cross_agg(l_img, r_img, lu, ru, iters)
    lv, rv = ctx.nyu_cross_agg(l_img, r_img, scene["lv"], scene["rv"], iters, exact_params())
    assert np.array_equal(np.transpose(lv, (2, 0, 1)), lref)
    assert np.array_equal(np.transpose(rv, (2, 0, 1)), rref)


CBCA_SEPARABLE_TOL = 2e-6     # max|a-b| / max|b|: same sum, row sums associated first


@pytest.mark.parametrize("iters", [1, 2])
def test_cross_agg_separable_within_tolerance(ctx, scene, iters):
    l_img, r_img = po.normalise_guide(scene["ld"]), po.normalise_guide(scene["rd"])
    lu = np.ascontiguousarray(np.transpose(scene["lv"], (2, 0, 1)))
    ru = np.ascontiguousarray(np.transpose(scene["rv"], (2, 0, 1)))
    lref, rref = po.nyu_cross_agg(l_img, r_img, lu, ru, iters)
    lv, rv = ctx.nyu_cross_agg(l_img, r_img, scene["lv"], scene["rv"], iters)      # default: separable
    for out, ref in ((lv, lref), (rv, rref)):
        out = np.transpose(out, (2, 0, 1))
        assert np.abs(out - ref).max() / np.abs(ref).max() < CBCA_SEPARABLE_TOL
        assert (np.argmax(out, axis=0) == np.argmax(ref, axis=0)).mean() >= 0.999


@pytest.mark.parametrize("D,h,w", [(201, 23, 61), (256, 11, 37), (33, 70, 9)])
def test_cross_agg_fused_random_scenes(ctx, D, h, w):
    """The fused single-pass CBCA (default for D <= 256) on piecewise-constant guides with arms of every length,
    image widths that do not fill the last thread's pixels and heights below one band, against the C restatement."""
    rng = np.random.default_rng(D + h)
    gl = np.repeat(np.repeat(rng.standard_normal((h // 3 + 1, w // 4 + 1)), 3, 0), 4, 1)[:h, :w].astype(np.float32) * 0.05
    gr = np.roll(gl, -2, axis=1)
    lv = rng.standard_normal((h, w, D)).astype(np.float32)
    rv = rng.standard_normal((h, w, D)).astype(np.float32)
    lu, ru = (np.ascontiguousarray(np.transpose(v, (2, 0, 1))) for v in (lv, rv))
    lref, rref = po.nyu_cross_agg(gl, gr, lu, ru, 2)
    lo, ro = ctx.nyu_cross_agg(gl, gr, lv, rv, 2)
    for out, ref in ((lo, lref), (ro, rref)):
        assert np.abs(np.transpose(out, (2, 0, 1)) - ref).max() / np.abs(ref).max() < CBCA_SEPARABLE_TOL
    le, re_ = ctx.nyu_cross_agg(gl, gr, lv, rv, 2, exact_params())
    assert np.array_equal(np.transpose(le, (2, 0, 1)), lref) and np.array_equal(np.transpose(re_, (2, 0, 1)), rref)


@pytest.mark.parametrize("direction", [-1, 1])
def test_sgm_bit_exact(ctx, scene, direction):
    l_img, r_img = po.normalise_guide(scene["ld"]), po.normalise_guide(scene["rd"])
    vol = scene["lv"] if direction == -1 else scene["rv"]
    vol = (vol / po.volume_std(vol)) * np.float32(-1)
    ref = po.nyu_sgm(l_img, r_img, np.ascontiguousarray(np.transpose(vol, (2, 0, 1))), direction)
    out = ctx.nyu_sgm(l_img, r_img, vol, direction)
    assert np.array_equal(np.transpose(out, (2, 0, 1)), ref)


@pytest.mark.parametrize("D,h,w", [(33, 6, 40), (201, 6, 40), (228, 9, 45), (132, 7, 70), (260, 5, 41), (64, 13, 37),
                                   (4, 5, 9), (512, 3, 18)])
def test_sgm_ragged_disparity_ranges(ctx, D, h, w):
    """D % 4 != 0 runs the register-prefetch kernels, the rest the shared-memory ring kernels (sgm_ring.cu):
    scanline counts that do not fill the last CTA, stages cut short at the end of a scanline."""
    rng = np.random.default_rng(D)
    l_img = rng.standard_normal((h, w)).astype(np.float32) * 0.1
    r_img = rng.standard_normal((h, w)).astype(np.float32) * 0.1
    vol = rng.standard_normal((h, w, D)).astype(np.float32)
    for direction in (-1, 1):
        ref = po.nyu_sgm(l_img, r_img, np.ascontiguousarray(np.transpose(vol, (2, 0, 1))), direction)
        out = ctx.nyu_sgm(l_img, r_img, vol, direction)
        assert np.array_equal(np.transpose(out, (2, 0, 1)), ref)


def test_nyu_post_bit_exact_on_reference_maps(ctx, golden_dir):
    """The shipped CA2 maps of image 000002 through LR check, fills, median and
    bilateral (sub-pixel needs a volume: a synthetic one is used, identical on
    both sides)."""
    root = os.path.join(golden_dir, "results")
    l = cv2.imread(os.path.join(root, "nyu_cost_img_2", "000169_10.png"), 0).astype(np.float32)[:120]
    r = cv2.imread(os.path.join(root, "right_nyu_cost_img_2", "000169_10.png"), 0).astype(np.float32)[:120]
    h, w = l.shape
    D = 201
    vol = np.random.default_rng(0).random((h, w, D)).astype(np.float32)
    fin, outl = ctx.nyu_post(l, r, vol)
    fref, oref = po.nyu_post([l, r], np.ascontiguousarray(np.transpose(vol, (2, 0, 1))))
    assert np.array_equal(outl, oref)
    assert np.array_equal(fin, fref)
    ref_png = cv2.imread(os.path.join(root, "outlier", "000169_10.png"), 0)[:120]
    # rows cut at 120: the LR check only looks along rows, so the crop is still pinned by K5
    assert np.array_equal((outl * 127).astype(np.uint8), ref_png)


@pytest.mark.parametrize("exact", [1, 0])
def test_full_chain_vs_oracle(ctx, scene, exact):
    D = scene["D"]
    ref = po.post_chain(scene["lv"], scene["rv"], scene["ld"], scene["rd"])
    params = _lib.default_post_params()
    params.cbca_exact = exact
    out = ctx.post(scene["lv"], scene["rv"], scene["ld"], scene["rd"], params)
    if exact:
        # before the global std everything is bit exact
        assert np.array_equal(out["nyu_cost_img"], ref["nyu_cost_img"])
        assert np.array_equal(out["right_nyu_cost_img"], ref["right_nyu_cost_img"])
    else:
        assert (out["nyu_cost_img"] == ref["nyu_cost_img"]).mean() >= 0.999
        assert (out["right_nyu_cost_img"] == ref["right_nyu_cost_img"]).mean() >= 0.999
    for name in ("nyu_sgm_img", "right_nyu_sgm_img", "nyu_cost_img_2", "right_nyu_cost_img_2", "outlier"):
        assert (out[name] == ref[name]).mean() >= 0.999, name
    diff = np.abs(out["nyu_post_float"] - ref["nyu_post_float"])
    assert (diff <= 1e-3).mean() >= 0.999
    assert (out["nyu_post"] == ref["nyu_post"]).mean() >= 0.999
    e_ref = so.three_pixel_error(ref["nyu_post"], scene["gt"])
    e_out = so.three_pixel_error(out["nyu_post"], scene["gt"])
    assert abs(e_ref - e_out) <= 0.0005
    assert e_out < so.three_pixel_error(scene["ld"], scene["gt"])


@pytest.mark.parametrize("D,h,w", [(33, 6, 40), (201, 7, 23), (228, 9, 45), (64, 37, 13), (4, 5, 9), (512, 3, 18)])
def test_sgm_eight_paths_bit_exact(ctx, D, h, w):
    """sgm_paths = 8 (BASELINE.json configs[3]; NOT the reference, whose chain is 4-path, nyu.lua:59-64): the four
    diagonal paths added to the original four, against the C restatement with the same switch (which
    tests/test_post_oracle.py ties to plain Python loops).  Wide, tall and tiny images: diagonals of every length."""
    rng = np.random.default_rng(D + h)
    l_img = rng.standard_normal((h, w)).astype(np.float32) * 0.1
    r_img = rng.standard_normal((h, w)).astype(np.float32) * 0.1
    vol = rng.standard_normal((h, w, D)).astype(np.float32)
    p8 = _lib.default_post_params()
    p8.sgm_paths = 8
    for direction in (-1, 1):
        ref = po.nyu_sgm(l_img, r_img, np.ascontiguousarray(np.transpose(vol, (2, 0, 1))), direction, paths=8)
        out = ctx.nyu_sgm(l_img, r_img, vol, direction, p8)
        assert np.array_equal(np.transpose(out, (2, 0, 1)), ref)


def test_full_chain_eight_paths_vs_oracle(ctx, scene):
    """The whole chain with the 8-path SGM (labelled apart from the parity figures: no reference counterpart)."""
    ref = po.post_chain(scene["lv"], scene["rv"], scene["ld"], scene["rd"], sgm_paths=8)
    p8 = _lib.default_post_params()
    p8.sgm_paths = 8
    out = ctx.post(scene["lv"], scene["rv"], scene["ld"], scene["rd"], p8)
    four = ctx.post(scene["lv"], scene["rv"], scene["ld"], scene["rd"])
    assert not np.array_equal(out["nyu_sgm_img"], four["nyu_sgm_img"])
    for name in ("nyu_cost_img", "nyu_sgm_img", "right_nyu_sgm_img", "nyu_cost_img_2", "right_nyu_cost_img_2", "outlier",
                 "nyu_post"):
        assert (out[name] == ref[name]).mean() >= 0.999, name


def test_stage_maps_of_the_merged_passes_equal_the_stand_alone_kernels(ctx, scene):
    """The chain takes the stage image and the std of the first CBCA's output from ONE pass over the volume and lets
    the last SGM pass write the arg min map of its result.  Each must equal what the stand-alone entry points give
    on the same data: WTA of the exported CBCA volume, SGM of the explicitly scaled volume, its arg min."""
    l_img, r_img = ctx.guide_image(scene["ld"]), ctx.guide_image(scene["rd"])
    out = ctx.post(scene["lv"], scene["rv"], scene["ld"], scene["rd"])
    lu, ru = ctx.nyu_cross_agg(l_img, r_img, scene["lv"], scene["rv"], 1)
    assert np.array_equal(out["nyu_cost_img"], ctx.wta(lu)[0])
    assert np.array_equal(out["right_nyu_cost_img"], ctx.wta(ru)[0])
    for vol, direction, name in ((lu, -1, "nyu_sgm_img"), (ru, 1, "right_nyu_sgm_img")):
        sd = np.float32(ctx.volume_std(vol))
        sg = ctx.nyu_sgm(l_img, r_img, (vol / sd) * np.float32(-1), direction)
        assert np.array_equal(out[name], ctx.wta(sg, take_min=True)[0]), name
        assert np.array_equal(out[name], np.argmin(sg, axis=2).astype(np.uint8)), name


def test_post_error_behaviour(ctx, scene):
    p = _lib.default_post_params()
    assert (p.L1, p.sgm_i, p.nyu_cost_agg_1, p.post) == (5, 1, 1, 1)
    assert p.pi1 == pytest.approx(2.3) and p.blur_sigma == pytest.approx(5.99)
    big = np.zeros((2, 4, 300), np.float32)
    with pytest.raises(_lib.StereoError) as e:
        ctx.post(big, big, np.zeros((2, 4), np.uint8), np.zeros((2, 4), np.uint8))
    assert e.value.code == -1
    with pytest.raises(_lib.StereoError) as e:
        ctx.nyu_sgm(np.zeros((2, 4), np.float32), np.zeros((2, 4), np.float32), np.zeros((2, 4, 600), np.float32), -1)
    assert e.value.code == -5


def test_lua_surface_mirror(ctx, scene):
    """smooth.nyu.cross_agg / sgm / post with the reference's [1,D,H,W] layout."""
    from stereo_matching_b200 import post as post_mod
    l_img, r_img = po.normalise_guide(scene["ld"]), po.normalise_guide(scene["rd"])
    lu = np.ascontiguousarray(np.transpose(scene["lv"], (2, 0, 1)))[None]
    ru = np.ascontiguousarray(np.transpose(scene["rv"], (2, 0, 1)))[None]
    h, w = l_img.shape
    post_mod.nyu.params.cbca_exact = 1
    lo, ro = post_mod.nyu.cross_agg(l_img.reshape(1, 1, h, w), r_img.reshape(1, 1, h, w), lu, ru, 1)
    post_mod.nyu.params.cbca_exact = 0
    lref, rref = po.nyu_cross_agg(l_img, r_img, lu[0], ru[0], 1)
    assert lo.shape == lu.shape and np.array_equal(lo[0], lref) and np.array_equal(ro[0], rref)


@pytest.mark.parametrize("h,w,D,kw,kh,iters", [(13, 17, 6, 5, 5, 1), (24, 40, 33, 5, 5, 2), (9, 31, 64, 3, 7, 3), (60, 90, 201, 5, 5, 1)])
def test_box_agg_bit_exact(ctx, h, w, D, kw, kh, iters):
    rng = np.random.default_rng(h * w)
    vol = rng.standard_normal((h, w, D)).astype(np.float32)
    assert np.array_equal(ctx.box_agg(vol, kw, kh, iters), po.box_agg(vol, kw, kh, iters))


def test_staged_post_path_equals_fused_pipeline_and_runs_the_optional_branches(ctx):
    """post.py's staged path (taken for cost_agg / cost_agg_2 / confLoc) with those options off must reproduce the
    fused `stereo_post` stage PNGs bit for bit; with them on it must produce the extra stage images."""
    from stereo_matching_b200 import post as bpost
    h, w, D = 40, 96, 33
    rng = np.random.default_rng(5)
    lv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    rv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    ld, rd = ctx.wta(lv)[0], ctx.wta(rv)[0]
    params = _lib.default_post_params()
    opt = bpost.build_parser().parse_args([])
    fused = ctx.post(lv, rv, ld, rd, params)
    staged = bpost.run_staged(ctx, opt, params, lv, rv, ld, rd)
    for name in _lib.STAGE_NAMES:
        assert np.array_equal(fused[name], staged[name]), name
    opt = bpost.build_parser().parse_args(["--cost_agg", "1", "--cost_agg_2", "2", "--confLoc", "1", "--thres", "23"])
    extra = bpost.run_staged(ctx, opt, params, lv, rv, ld, rd)
    for name in ("cost_img", "post_cost_img", "right_post_cost_img", "nyu_post", "outlier"):
        assert extra[name].shape == (h, w) and extra[name].dtype == np.uint8
    agg = po.box_agg(lv, 5, 5, 1)
    ref = np.where(agg.max(axis=2) < 23, 255, agg.argmax(axis=2)).astype(np.uint8)
    assert np.array_equal(extra["cost_img"], ref)


@pytest.mark.parametrize("D,h,w", [(64, 72, 200), (201, 41, 263)])
def test_pairs_post_equals_inference_then_post(ctx, D, h, w):
    """`stereo_pairs_post` (volumes never leave the device) == `stereo_pairs` with exported volumes followed by
    `stereo_post` on them, bit for bit, for a batch of two pairs; also at the reference's default disp_range 201,
    whose volume rows are not 16-byte aligned (other SGM / WTA code paths), with odd image sizes."""
    pairs = [synth.make_pair(4100 + i, h, w, D) for i in range(2)]
    left = np.stack([p[0] for p in pairs])
    right = np.stack([p[1] for p in pairs])
    fused = ctx.pairs_post(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3)
    ref = ctx.pairs(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3, volumes=True)
    for i in range(2):
        st = ctx.post(ref["left_cost"][i], ref["right_cost"][i], ref["left_disp"][i], ref["right_disp"][i])
        assert np.array_equal(fused["nyu_post"][i], st["nyu_post"])
        assert np.array_equal(fused["outlier"][i], st["outlier"])
        assert np.array_equal(fused["left_disp"][i], ref["left_disp"][i])
    from stereo_matching_b200.metric import three_pixel_error
    gt = pairs[0][2]
    assert three_pixel_error(fused["nyu_post"][0].astype(np.int64), gt) <= three_pixel_error(ref["left_disp"][0].astype(np.int64), gt)


# ---- the benchmark size: 375 x 1242 at D = 228 (BASELINE.json) and at the reference's default D = 201 --------------
# The ring / band / partition logic of the default kernels (sgm_ring.cu stages over 1 242 steps and 375 scanlines, the
# row bands of cbca_fused.cu, the unaligned rows of D = 201) against the C restatement at the size they are timed at.
@pytest.fixture(scope="module", params=[228, 201])
def full_scene(request, ctx):
    D = request.param
    l, r, gt = synth.make_pair(1000, 375, 1242, D)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    sc = {"D": D, "gt": gt, "lv": out["left_cost"][0], "rv": out["right_cost"][0],
          "ld": out["left_disp"][0], "rd": out["right_disp"][0]}
    sc["ref"] = po.post_chain(sc["lv"], sc["rv"], sc["ld"], sc["rd"])
    return sc


@pytest.mark.parametrize("exact", [0, 1])
def test_full_size_chain_vs_oracle(ctx, full_scene, exact):
    """post.lua:116-243 on one full pair: every stage map and the final float map (gates of test_full_chain_vs_oracle)."""
    sc, ref = full_scene, full_scene["ref"]
    params = _lib.default_post_params()
    params.cbca_exact = exact
    out = ctx.post(sc["lv"], sc["rv"], sc["ld"], sc["rd"], params)
    for name in ("nyu_cost_img", "right_nyu_cost_img"):
        if exact:
            assert np.array_equal(out[name], ref[name]), name            # before the global std: bit exact
        else:
            assert (out[name] == ref[name]).mean() >= 0.999, name
    for name in ("nyu_sgm_img", "right_nyu_sgm_img", "nyu_cost_img_2", "right_nyu_cost_img_2", "outlier"):
        assert (out[name] == ref[name]).mean() >= 0.999, name
    diff = np.abs(out["nyu_post_float"] - ref["nyu_post_float"])
    assert (diff <= 1e-3).mean() >= 0.999
    assert (out["nyu_post"] == ref["nyu_post"]).mean() >= 0.999
    e_ref = so.three_pixel_error(ref["nyu_post"], sc["gt"])
    e_out = so.three_pixel_error(out["nyu_post"], sc["gt"])
    assert abs(e_ref - e_out) <= 0.0005
    assert e_out < so.three_pixel_error(sc["ld"], sc["gt"])


def test_full_size_sgm_bit_exact(ctx, full_scene):
    """smooth.nyu.sgm alone on the full-size volumes (nyu.lua:47-75), both matching directions, bit for bit."""
    sc = full_scene
    l_img, r_img = po.normalise_guide(sc["ld"]), po.normalise_guide(sc["rd"])
    for direction, vol in ((-1, sc["lv"]), (1, sc["rv"])):
        vol = (vol / po.volume_std(vol)) * np.float32(-1)
        ref = po.nyu_sgm(l_img, r_img, np.ascontiguousarray(np.transpose(vol, (2, 0, 1))), direction)
        out = ctx.nyu_sgm(l_img, r_img, vol, direction)
        assert np.array_equal(np.transpose(out, (2, 0, 1)), ref), direction
        del ref, out


def test_full_size_cross_agg(ctx, full_scene):
    """smooth.nyu.cross_agg alone on the full-size volumes (nyu.lua:24-45): the original's summation order bit for
    bit, the default single-pass kernel within its stated tolerance."""
    sc = full_scene
    l_img, r_img = po.normalise_guide(sc["ld"]), po.normalise_guide(sc["rd"])
    lu = np.ascontiguousarray(np.transpose(sc["lv"], (2, 0, 1)))
    ru = np.ascontiguousarray(np.transpose(sc["rv"], (2, 0, 1)))
    lref, rref = po.nyu_cross_agg(l_img, r_img, lu, ru, 1)
    del lu, ru
    le, re_ = ctx.nyu_cross_agg(l_img, r_img, sc["lv"], sc["rv"], 1, exact_params())
    assert np.array_equal(np.transpose(le, (2, 0, 1)), lref) and np.array_equal(np.transpose(re_, (2, 0, 1)), rref)
    del le, re_
    lo, ro = ctx.nyu_cross_agg(l_img, r_img, sc["lv"], sc["rv"], 1)
    for out, ref in ((lo, lref), (ro, rref)):
        assert np.abs(np.transpose(out, (2, 0, 1)) - ref).max() / np.abs(ref).max() < CBCA_SEPARABLE_TOL


def test_cli_workflow_inference_post_cal(tmp_path, capsys, ws9_layers):
    """The reference's workflow end to end through the three command lines: `inference.py` (maps + .bin volumes,
    inference.py:202-231) -> `post.lua` (reads them back, post.lua:78-92, writes the stage PNG directories,
    :124-132,158-166,177-185,237-241) -> `cal.py` (3-pixel error of nyu_post against gt).  The stage PNGs must be
    those of `ctx.post` on the same volumes, and cal.py's number that of the metric on them."""
    from stereo_matching_b200 import cal, inference
    from stereo_matching_b200 import post as bpost
    from stereo_matching_b200.metric import three_pixel_error
    res = tmp_path / "results"
    h, w, D = 60, 200, 40
    golden = os.path.join(os.path.dirname(__file__), "golden", "new_checkpoint")
    inference.main(["--model_dir", golden, "--out_dir", str(res), "--synthetic", "2", "--height", str(h),
                    "--width", str(w), "--disp_range", str(D)])
    text = capsys.readouterr().out
    assert "Mean Error:" in text
    n = bpost.run(bpost.build_parser().parse_args(["--saveDir", str(res), "--disp_range", str(D)]))
    assert n == 2
    m = base_model_ws_9((None, None, 3))
    m.set_layers(ws9_layers)
    errs = []
    for fid in (0, 1):
        l, r, gt = synth.make_pair(1000 + fid, h, w, D)
        out = m.context.pairs(_lib.NET_WIN9, l, r, D, precision=m.precision, volumes=True)
        lv = np.fromfile(res / "left_cost" / ("left_%06d_10_%d_%d_%d.bin" % (fid, h, w, D)), "<f4").reshape(h, w, D)
        assert np.array_equal(lv, out["left_cost"][0])
        assert np.array_equal(cv2.imread(str(res / "left_disp" / ("left_%06d_10.png" % fid)), 0), out["left_disp"][0])
        assert np.array_equal(cv2.imread(str(res / "gt" / ("true_%06d_10.png" % fid)), 0), gt)
        st = m.context.post(out["left_cost"][0], out["right_cost"][0], out["left_disp"][0], out["right_disp"][0])
        for name in _lib.STAGE_NAMES:
            png = cv2.imread(str(res / name / ("%06d_10.png" % fid)), cv2.IMREAD_GRAYSCALE)
            assert png is not None and np.array_equal(png, st[name]), name
        errs.append(three_pixel_error(st["nyu_post"], gt))
    mean = cal.run(str(res))
    assert mean == pytest.approx(np.mean(errs))
    assert mean <= float(text.split("Mean Error:")[1].split()[0]) + 0.01   # the chain does not lose to the raw WTA maps


def test_staged_path_follows_post_lua_order_with_optional_branches(ctx):
    """post.py's staged path with cost_agg / cost_agg_2 on against the restated post.lua (oracle post_chain with the
    same options): CBCA 2 runs before the second box aggregation (post.lua:168-186, then :188-221)."""
    from stereo_matching_b200 import post as bpost
    h, w, D = 40, 96, 33
    rng = np.random.default_rng(11)
    lv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    rv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    ld, rd = ctx.wta(lv)[0], ctx.wta(rv)[0]
    params = exact_params()
    opt = bpost.build_parser().parse_args(["--cost_agg", "1", "--cost_agg_2", "2"])
    out = bpost.run_staged(ctx, opt, params, lv, rv, ld, rd)
    ref = po.post_chain(lv, rv, ld, rd, cost_agg=1, cost_agg_2=2)
    for name in ("cost_img", "nyu_cost_img", "right_nyu_cost_img"):
        assert np.array_equal(out[name], ref[name]), name
    for name in ("nyu_sgm_img", "nyu_cost_img_2", "right_nyu_cost_img_2", "post_cost_img", "right_post_cost_img",
                 "nyu_post", "outlier"):
        assert (out[name] == ref[name]).mean() >= 0.999, name
