"""GPU parity tests of the inference path (T1-T5), all through the C ABI
(`stereo_matching_b200._lib` is a plain ctypes binding of include/stereo_b200.h).

Gates (BASELINE.json north_star): fp32 cost volume within 1e-4 of the float64
oracle relative to max|ref|; WTA maps bit-exact except at float ties,
>= 99.9 % agreement; 3-px error within 0.05 percentage points.
"""
import os

import cv2
import numpy as np
import pytest

from oracle import post_oracle as po
from oracle import stereo_oracle as so
from stereo_matching_b200 import _lib, synth
from stereo_matching_b200.models.model import base_model, base_model_ws_9

pytestmark = pytest.mark.gpu

COST_TOL = 1e-4      # north star: max|a-b| / max|b|
WTA_AGREE = 0.999


@pytest.fixture(scope="module")
def ws9(ws9_layers):
    # the fp32 CUDA-core path on purpose: the tcgen05 path (the default precision) has its own section below
    m = base_model_ws_9((None, None, 3), precision=_lib.PREC_FP32)
    m.set_layers(ws9_layers)
    return m


@pytest.fixture(scope="module")
def ctx(ws9):
    # the context that holds the ckpt-1000 weights
    return ws9.context


def rel_err(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / np.abs(b).max())


@pytest.mark.parametrize("shape,patch", [((7, 9), 9), ((33, 70), 9), ((20, 31), 37), ((375, 1242), 9)])
def test_normalize_pad(ctx, shape, patch):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
    ref = so.normalize_and_pad(img, patch)
    out = ctx.normalize_pad(img, patch)
    assert out.shape == ref.shape
    pad = (patch - 1) // 2
    assert np.all(out[0, :pad] == 0) and np.all(out[0, :, :pad] == 0)
    np.testing.assert_array_max_ulp(out, ref, maxulp=1)


@pytest.mark.parametrize("hw", [(12, 14), (24, 50), (41, 77), (9, 9)])
def test_trunk_ws9_vs_oracle(ws9, ws9_layers, hw):
    rng = np.random.default_rng(hw[0])
    x = rng.standard_normal((2,) + hw + (3,)).astype(np.float32)
    ref = so.trunk_forward(x, ws9_layers, "win9_dep9", np.float64)
    out = ws9(x, training=False)
    assert out.shape == ref.shape == (2, hw[0] - 8, hw[1] - 8, 64)
    assert rel_err(out, ref) < 1e-5


def test_trunk_win37_vs_oracle():
    layers = synth.make_base_model_weights()
    m = base_model((None, None, 3), precision=_lib.PREC_FP32)
    m.set_layers(layers)
    x = np.random.default_rng(3).standard_normal((1, 45, 60, 3)).astype(np.float32)
    ref = so.trunk_forward(x, layers, "win37_dep9", np.float64)
    out = m(x, training=False)
    assert out.shape == ref.shape == (1, 9, 24, 64)
    assert rel_err(out, ref) < 1e-5


@pytest.mark.parametrize("h,w,D", [(3, 10, 6), (4, 70, 12), (2, 64, 64), (3, 129, 201), (2, 300, 228), (2, 20, 33)])
def test_cost_volume_vs_oracle(ctx, h, w, D):
    rng = np.random.default_rng(w)
    L = rng.standard_normal((1, h, w, 64)).astype(np.float32)
    R = rng.standard_normal((1, h, w, 64)).astype(np.float32)
    lref, rref = so.cost_volumes(L, R, D, np.float64)
    lv, rv = ctx.cost_volume(L, R, D)
    assert rel_err(lv[0], lref) < COST_TOL and rel_err(rv[0], rref) < COST_TOL
    # the unwritten slots of inference.py:140-141 are exactly zero
    xs = np.arange(w)[:, None]
    ds = np.arange(D)[None, :]
    assert np.all(lv[0][:, xs < ds] == 0)
    assert np.all(rv[0][:, xs + ds >= w] == 0)
    # both volumes hold the same numbers (inference.py:149-150)
    for d in range(0, min(D, w), max(D // 5, 1)):
        assert np.array_equal(lv[0][:, d:, d], rv[0][:, :w - d, d])


def test_wta_matches_numpy_argmax(ctx):
    rng = np.random.default_rng(0)
    vol = rng.standard_normal((2, 5, 37, 201)).astype(np.float32)
    vol[0, 0, 0] = 0.0                    # all ties -> index 0
    vol[0, 0, 1] = -1.0
    vol[0, 0, 1, 7] = vol[0, 0, 1, 150] = 3.0   # tie -> lowest index
    assert np.array_equal(ctx.wta(vol), np.argmax(vol, axis=3).astype(np.uint8))
    assert np.array_equal(ctx.wta(vol, take_min=True), np.argmin(vol, axis=3).astype(np.uint8))


def test_wta_selects_zero_filled_slot_when_all_scores_negative(ctx):
    L = np.ones((1, 1, 5, 4), np.float32)
    R = -np.ones((1, 1, 5, 4), np.float32)
    ld, rd = ctx.corr_wta(L, R, 4)
    lref, rref = so.cost_volumes(L, R, 4)
    assert np.array_equal(ld[0], so.wta(lref)) and np.array_equal(rd[0], so.wta(rref))
    assert list(ld[0, 0]) == [1, 2, 3, 0, 0]


@pytest.mark.parametrize("h,w,D", [(6, 150, 64), (3, 200, 228), (2, 90, 201)])
def test_fused_corr_wta_equals_wta_of_volume(ctx, h, w, D):
    rng = np.random.default_rng(D)
    L = rng.standard_normal((2, h, w, 64)).astype(np.float32)
    R = rng.standard_normal((2, h, w, 64)).astype(np.float32)
    lv, rv = ctx.cost_volume(L, R, D)
    ld, rd = ctx.corr_wta(L, R, D)
    assert np.array_equal(ld, np.argmax(lv, axis=3).astype(np.uint8))
    assert np.array_equal(rd, np.argmax(rv, axis=3).astype(np.uint8))
    assert np.array_equal(ld, ctx.wta(lv)) and np.array_equal(rd, ctx.wta(rv))


def test_lr_check_bit_exact_on_reference_pngs(ctx, golden_dir):
    """K5 through the CUDA kernel: results/outlier of the reference."""
    root = os.path.join(golden_dir, "results")
    for i in ("000002", "000169", "000192"):
        l = cv2.imread(os.path.join(root, "nyu_cost_img_2", "%s_10.png" % i), 0)
        r = cv2.imread(os.path.join(root, "right_nyu_cost_img_2", "%s_10.png" % i), 0)
        o = cv2.imread(os.path.join(root, "outlier", "%s_10.png" % i), 0)
        out = ctx.lr_check(l, r, 201)[0]
        assert np.array_equal(out * 127, o), i
        assert np.array_equal(out, po.outlier_detection(l.astype(np.float32), r.astype(np.float32), 201).astype(np.uint8))


def test_pair_end_to_end_vs_oracle(ctx, ws9, ws9_layers):
    D = 64
    l, r, gt = synth.make_pair(1003, 96, 320, D)
    ref = so.infer_pair(l, r, ws9_layers, "win9_dep9", 9, D, np.float64)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=True)
    assert rel_err(out["left_cost"][0], ref["left_cost"]) < COST_TOL
    assert rel_err(out["right_cost"][0], ref["right_cost"]) < COST_TOL
    assert (out["left_disp"][0] == ref["left_disp"]).mean() >= WTA_AGREE
    assert (out["right_disp"][0] == ref["right_disp"]).mean() >= WTA_AGREE
    e_ref = so.three_pixel_error(ref["left_disp"], gt)
    e_out = so.three_pixel_error(out["left_disp"][0], gt)
    assert abs(e_ref - e_out) <= 0.0005            # 0.05 percentage points
    ref_out = po.outlier_detection(out["left_disp"][0].astype(np.float32), out["right_disp"][0].astype(np.float32), D)
    assert np.array_equal(out["outlier"][0], ref_out.astype(np.uint8))
    # fused path (no volumes) gives the same maps
    out2 = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=False)
    assert np.array_equal(out2["left_disp"], out["left_disp"]) and np.array_equal(out2["right_disp"], out["right_disp"])


def test_batch_of_pairs_equals_single_pairs(ctx, ws9):
    D = 40
    pairs = [synth.make_pair(1100 + i, 30, 100, D) for i in range(3)]
    L = np.stack([p[0] for p in pairs])
    R = np.stack([p[1] for p in pairs])
    out = ctx.pairs(_lib.NET_WIN9, L, R, D)
    for i in range(3):
        one = ctx.pairs(_lib.NET_WIN9, L[i], R[i], D)
        assert np.array_equal(out["left_disp"][i], one["left_disp"][0])
        assert np.array_equal(out["right_disp"][i], one["right_disp"][0])
        assert np.array_equal(out["outlier"][i], one["outlier"][0])


def test_full_size_properties(ctx, ws9):
    """BASELINE config 1 (375x1242, D=228) through size-independent properties:
    the two volumes are the same band indexed two ways, invalid slots are exactly
    zero, fused WTA equals WTA of the materialised volume, GT recovered."""
    D = 228
    l, r, gt = synth.make_pair(1000, 375, 1242, D)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=True)
    lv, rv = out["left_cost"][0], out["right_cost"][0]
    w = 1242
    for d in (0, 1, 57, 113, 200, 227):
        assert np.array_equal(lv[:, d:, d], rv[:, :w - d, d])
        assert np.all(lv[:, :d, d] == 0) and np.all(rv[:, w - d:, d] == 0)
    assert np.array_equal(out["left_disp"][0], np.argmax(lv, axis=2).astype(np.uint8))
    assert np.array_equal(out["right_disp"][0], np.argmax(rv, axis=2).astype(np.uint8))
    fused = ctx.pairs(_lib.NET_WIN9, l, r, D, volumes=False)
    assert np.array_equal(fused["left_disp"], out["left_disp"])
    assert so.three_pixel_error(out["left_disp"][0], gt) < 0.2
    # the float64 oracle on a strip of 8 rows (its padded input rows are 100..115)
    lp, rp = so.normalize_and_pad(l, 9), so.normalize_and_pad(r, 9)
    layers = ws9.layers
    lm = so.trunk_forward(lp[:, 100:116], layers, "win9_dep9", np.float64)
    rm = so.trunk_forward(rp[:, 100:116], layers, "win9_dep9", np.float64)
    lref, rref = so.cost_volumes(lm, rm, D, np.float64)
    assert rel_err(lv[100:108], lref) < COST_TOL and rel_err(rv[100:108], rref) < COST_TOL
    assert (out["left_disp"][0][100:108] == so.wta(lref)).mean() >= WTA_AGREE


def test_error_behaviour(ctx, ws9):
    with pytest.raises(_lib.StereoError) as e:
        ctx.cost_volume(np.zeros((1, 2, 8, 64), np.float32), np.zeros((1, 2, 8, 64), np.float32), 0)
    assert e.value.code == -1
    with pytest.raises(_lib.StereoError) as e:
        ctx.cost_volume(np.zeros((1, 2, 8, 6), np.float32), np.zeros((1, 2, 8, 6), np.float32), 4)
    assert e.value.code == -1
    with pytest.raises(_lib.StereoError) as e:
        ctx.cost_volume(np.zeros((1, 1, 8, 64), np.float32), np.zeros((1, 1, 8, 64), np.float32), 2000)
    assert e.value.code == -5
    with pytest.raises(_lib.StereoError) as e:
        ctx.trunk(_lib.NET_WIN9, np.zeros((1, 8, 30, 3), np.float32))
    assert e.value.code == -1
    fresh = _lib.Context(0)
    with pytest.raises(_lib.StereoError) as e:
        fresh.trunk(_lib.NET_WIN37, np.zeros((1, 40, 40, 3), np.float32))
    assert e.value.code == -3
    with pytest.raises(_lib.StereoError):
        fresh.set_weights(_lib.NET_WIN9, np.zeros(10, np.float32))
    fresh.close()
    with pytest.raises(ValueError):
        ws9(np.zeros((1, 8, 8, 3), np.float32))
    with pytest.raises(NotImplementedError):
        ws9(np.zeros((1, 20, 20, 3), np.float32), training=True)


# ---- tensor-core trunk (tcgen05 / TMEM), STEREO_PREC_BF16X3 and STEREO_PREC_BF16 ----------------
BF16X3_TOL = 2e-5     # 3-pass bf16 split: SURVEY section 7 measured 5.6e-6 on ckpt-1000
BF16_TOL = 2e-2       # single bf16 pass: its own, much looser tolerance (does NOT meet the 1e-4 gate)


@pytest.mark.parametrize("hw", [(12, 14), (24, 50), (20, 300), (75, 140), (11, 139)])
def test_trunk_tensor_core_bf16x3_vs_oracle(ctx, ws9_layers, hw):
    rng = np.random.default_rng(hw[1])
    x = rng.standard_normal((2,) + hw + (3,)).astype(np.float32)
    ref = so.trunk_forward(x, ws9_layers, "win9_dep9", np.float64)
    out = ctx.trunk(_lib.NET_WIN9, x, _lib.PREC_BF16X3)
    assert out.shape == ref.shape
    assert rel_err(out, ref) < BF16X3_TOL
    fp32 = ctx.trunk(_lib.NET_WIN9, x, _lib.PREC_FP32)
    assert rel_err(out, fp32.astype(np.float64)) < BF16X3_TOL


def test_trunk_tensor_core_bf16_reports_its_own_tolerance(ctx, ws9_layers):
    x = np.random.default_rng(9).standard_normal((1, 40, 150, 3)).astype(np.float32)
    ref = so.trunk_forward(x, ws9_layers, "win9_dep9", np.float64)
    out = ctx.trunk(_lib.NET_WIN9, x, _lib.PREC_BF16)
    err = rel_err(out, ref)
    assert 1e-5 < err < BF16_TOL        # clearly worse than the split, clearly not garbage


def test_pair_end_to_end_tensor_core(ctx, ws9_layers):
    D = 64
    l, r, gt = synth.make_pair(1003, 96, 320, D)
    ref = so.infer_pair(l, r, ws9_layers, "win9_dep9", 9, D, np.float64)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    assert rel_err(out["left_cost"][0], ref["left_cost"]) < COST_TOL
    assert rel_err(out["right_cost"][0], ref["right_cost"]) < COST_TOL
    assert (out["left_disp"][0] == ref["left_disp"]).mean() >= WTA_AGREE
    assert (out["right_disp"][0] == ref["right_disp"]).mean() >= WTA_AGREE
    assert abs(so.three_pixel_error(ref["left_disp"], gt) - so.three_pixel_error(out["left_disp"][0], gt)) <= 0.0005


@pytest.fixture(scope="module")
def base_layers(ctx):
    layers = synth.make_base_model_weights()
    m = base_model((None, None, 3))
    m.set_layers(layers)              # loads STEREO_NET_WIN37 weights into the shared context
    return layers


# nine 5x5 layers with K = 800-1600 per output (ws9: four 3x3 layers, K = 576): the dropped lo*lo terms of the
# split add up to ~3e-5 of max|ref| on the features; the cost volume gate (1e-4) is checked further down
BASE_BF16X3_TOL = 6e-5


@pytest.mark.parametrize("hw", [(45, 60), (37, 37), (40, 200), (52, 300)])
def test_base_model_trunk_tensor_core_bf16x3_vs_oracle(ctx, base_layers, hw):
    """base_model (5x5, 32/64 channels, models/model.py:5-28) through conv_tcg: layers 2-9 on tcgen05."""
    x = np.random.default_rng(hw[1]).standard_normal((2,) + hw + (3,)).astype(np.float32)
    ref = so.trunk_forward(x, base_layers, "win37_dep9", np.float64)
    out = ctx.trunk(_lib.NET_WIN37, x, _lib.PREC_BF16X3)
    assert out.shape == ref.shape == (2, hw[0] - 36, hw[1] - 36, 64)
    assert rel_err(out, ref) < BASE_BF16X3_TOL
    fp32 = ctx.trunk(_lib.NET_WIN37, x, _lib.PREC_FP32)
    assert rel_err(out, fp32.astype(np.float64)) < BASE_BF16X3_TOL


def test_base_model_batched_pairs_equal_single_pairs(ctx, base_layers):
    """Three pairs in one call alternate over two streams with their own padded images, activations and feature
    planes: every pair's maps and volumes must be those of a call with that pair alone."""
    D = 40
    pairs = [synth.make_pair(1300 + i, 38, 150, D) for i in range(3)]
    left, right = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    both = ctx.pairs(_lib.NET_WIN37, left, right, D, precision=_lib.PREC_BF16X3, volumes=True)
    for i in range(3):
        one = ctx.pairs(_lib.NET_WIN37, left[i], right[i], D, precision=_lib.PREC_BF16X3, volumes=True)
        for k in ("left_disp", "right_disp", "outlier", "left_cost", "right_cost"):
            assert np.array_equal(both[k][i], one[k][0]), (i, k)


def test_base_model_trunk_tensor_core_many_bands_vs_fp32_path(ctx, base_layers):
    """A larger input (four 128-pixel strips, a dozen row bands per image) against the fp32 CUDA-core path, which the
    small cases above tie to the float64 oracle."""
    x = np.random.default_rng(21).standard_normal((2, 130, 420, 3)).astype(np.float32)
    fp32 = ctx.trunk(_lib.NET_WIN37, x, _lib.PREC_FP32)
    out = ctx.trunk(_lib.NET_WIN37, x, _lib.PREC_BF16X3)
    assert out.shape == fp32.shape == (2, 94, 384, 64)
    assert rel_err(out, fp32.astype(np.float64)) < BASE_BF16X3_TOL


def test_base_model_full_size_strip_vs_oracle(ctx, base_layers):
    """base_model at the benchmark size (375x1242, D = 228, models/model.py:5-28) through the tensor-core kernels:
    band symmetry and zero fill of the whole volumes, fused WTA == WTA of the materialised volume, and the float64
    oracle on strips of rows at the top, in the middle and at the bottom of the image (padding rows included)."""
    D = 228
    h, w = 375, 1242
    l, r, gt = synth.make_pair(1000, h, w, D)
    out = ctx.pairs(_lib.NET_WIN37, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    lv, rv = out["left_cost"][0], out["right_cost"][0]
    for d in (0, 1, 113, 227):
        assert np.array_equal(lv[:, d:, d], rv[:, :w - d, d])
        assert np.all(lv[:, :d, d] == 0) and np.all(rv[:, w - d:, d] == 0)
    assert np.array_equal(out["left_disp"][0], np.argmax(lv, axis=2).astype(np.uint8))
    assert np.array_equal(out["right_disp"][0], np.argmax(rv, axis=2).astype(np.uint8))
    lp, rp = so.normalize_and_pad(l, 37), so.normalize_and_pad(r, 37)
    rows = 3
    for y0 in (0, 186, h - rows):
        lm = so.trunk_forward(lp[:, y0:y0 + rows + 36], base_layers, "win37_dep9", np.float64)
        rm = so.trunk_forward(rp[:, y0:y0 + rows + 36], base_layers, "win37_dep9", np.float64)
        lref, rref = so.cost_volumes(lm, rm, D, np.float64)
        assert rel_err(lv[y0:y0 + rows], lref) < COST_TOL and rel_err(rv[y0:y0 + rows], rref) < COST_TOL, y0
        assert (out["left_disp"][0][y0:y0 + rows] == so.wta(lref)).mean() >= WTA_AGREE, y0


def test_default_precision_is_the_tensor_core_path(ws9_layers):
    """`base_model_ws_9(...)` without a precision runs the tcgen05 kernels (split bf16, fp32 accumulate)."""
    m = base_model_ws_9((None, None, 3))
    assert m.precision == _lib.PREC_BF16X3
    m.set_layers(ws9_layers)
    x = np.random.default_rng(8).standard_normal((2, 30, 150, 3)).astype(np.float32)
    before = m.context.launch_count
    out = m(x, training=False)
    assert m.context.launch_count > before
    assert rel_err(out, so.trunk_forward(x, ws9_layers, "win9_dep9", np.float64)) < 2e-5
    assert np.array_equal(out, m.context.trunk(_lib.NET_WIN9, x, _lib.PREC_BF16X3))


def test_base_model_trunk_tensor_core_bf16_reports_its_own_tolerance(ctx, base_layers):
    x = np.random.default_rng(4).standard_normal((1, 45, 170, 3)).astype(np.float32)
    ref = so.trunk_forward(x, base_layers, "win37_dep9", np.float64)
    err = rel_err(ctx.trunk(_lib.NET_WIN37, x, _lib.PREC_BF16), ref)
    assert 1e-5 < err < BF16_TOL


def test_base_model_pair_end_to_end_tensor_core(ctx, base_layers):
    D = 48
    l, r, gt = synth.make_pair(1007, 40, 200, D)
    ref = so.infer_pair(l, r, base_layers, "win37_dep9", 37, D, np.float64)
    out = ctx.pairs(_lib.NET_WIN37, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    assert rel_err(out["left_cost"][0], ref["left_cost"]) < COST_TOL
    assert rel_err(out["right_cost"][0], ref["right_cost"]) < COST_TOL
    assert (out["left_disp"][0] == ref["left_disp"]).mean() >= WTA_AGREE
    assert (out["right_disp"][0] == ref["right_disp"]).mean() >= WTA_AGREE
    fp32 = ctx.pairs(_lib.NET_WIN37, l, r, D, precision=_lib.PREC_FP32)
    assert (out["left_disp"][0] == fp32["left_disp"][0]).mean() >= WTA_AGREE


# ---- tensor-core correlation (tcgen05), STEREO_PREC_BF16X3 ---------------------------------------
@pytest.mark.parametrize("h,w,D", [(3, 10, 6), (2, 300, 228), (3, 129, 201), (2, 20, 33), (2, 150, 257), (4, 260, 64)])
def test_cost_volume_tensor_core_vs_oracle(ctx, h, w, D):
    rng = np.random.default_rng(w + D)
    L = rng.standard_normal((2, h, w, 64)).astype(np.float32)
    R = rng.standard_normal((2, h, w, 64)).astype(np.float32)
    lv, rv = ctx.cost_volume(L, R, D, _lib.PREC_BF16X3)
    for i in range(2):
        lref, rref = so.cost_volumes(L[i:i + 1], R[i:i + 1], D, np.float64)
        assert rel_err(lv[i], lref) < COST_TOL and rel_err(rv[i], rref) < COST_TOL
    xs = np.arange(w)[:, None]
    ds = np.arange(D)[None, :]
    assert np.all(lv[0][:, xs < ds] == 0)            # zero-filled slots are exact zeros
    assert np.all(rv[0][:, xs + ds >= w] == 0)
    scale = np.abs(lv).max()
    for d in range(0, min(D, w), max(D // 5, 1)):
        assert np.abs(lv[0][:, d:, d] - rv[0][:, :w - d, d]).max() <= 1e-6 * scale
    if D > 256:
        # disparity maps are bytes like the reference's PNGs: a wider range is refused, not wrapped
        with pytest.raises(_lib.StereoError) as e:
            ctx.corr_wta(L, R, D, _lib.PREC_BF16X3)
        assert e.value.code == -1
        return
    # fused WTA sees exactly the values that the materialising variant stores
    ld, rd = ctx.corr_wta(L, R, D, _lib.PREC_BF16X3)
    assert np.array_equal(ld, np.argmax(lv, axis=3).astype(np.uint8))
    assert np.array_equal(rd, np.argmax(rv, axis=3).astype(np.uint8))


def test_tensor_core_wta_zero_fill_and_ties(ctx):
    L = np.ones((1, 1, 5, 64), np.float32)
    R = -np.ones((1, 1, 5, 64), np.float32)
    ld, rd = ctx.corr_wta(L, R, 4, _lib.PREC_BF16X3)
    lref, rref = so.cost_volumes(L, R, 4)
    assert np.array_equal(ld[0], so.wta(lref)) and np.array_equal(rd[0], so.wta(rref))
    Z = np.zeros((1, 2, 140, 64), np.float32)      # all ties -> disparity 0 everywhere
    ld, rd = ctx.corr_wta(Z, Z, 100, _lib.PREC_BF16X3)
    assert not ld.any() and not rd.any()


def test_tensor_core_correlation_falls_back_to_cuda_cores_for_wide_ranges(ctx):
    rng = np.random.default_rng(1)
    L = rng.standard_normal((1, 1, 400, 64)).astype(np.float32)
    R = rng.standard_normal((1, 1, 400, 64)).astype(np.float32)
    lv, rv = ctx.cost_volume(L, R, 300, _lib.PREC_BF16X3)      # D > 257: fp32 CUDA-core kernel
    lref, rref = so.cost_volumes(L, R, 300, np.float64)
    assert rel_err(lv[0], lref) < 1e-5 and rel_err(rv[0], rref) < 1e-5


def test_full_size_tensor_core_pipeline(ctx):
    """375x1242, D=228 through trunk + correlation on the tensor cores vs the fp32 CUDA-core path."""
    D = 228
    l, r, gt = synth.make_pair(1001, 375, 1242, D)
    a = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_FP32, volumes=True)
    b = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    assert rel_err(b["left_cost"][0], a["left_cost"][0].astype(np.float64)) < COST_TOL
    assert rel_err(b["right_cost"][0], a["right_cost"][0].astype(np.float64)) < COST_TOL
    assert (a["left_disp"] == b["left_disp"]).mean() >= WTA_AGREE
    assert (a["right_disp"] == b["right_disp"]).mean() >= WTA_AGREE
    assert np.array_equal(b["left_disp"][0], np.argmax(b["left_cost"][0], axis=2).astype(np.uint8))
    c = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=False)
    assert np.array_equal(c["left_disp"], b["left_disp"]) and np.array_equal(c["right_disp"], b["right_disp"])
    assert abs(so.three_pixel_error(a["left_disp"][0], gt) - so.three_pixel_error(b["left_disp"][0], gt)) <= 0.0005


@pytest.mark.parametrize("D", [4, 64, 228, 256])
def test_wta_vector_path_matches_numpy(ctx, D):
    rng = np.random.default_rng(D)
    vol = rng.standard_normal((1, 7, 53, D)).astype(np.float32)
    vol[0, 0, 0] = 0.0
    vol[0, 0, 1, D // 3] = vol[0, 0, 1, D - 1] = 9.0        # tie -> lowest index
    assert np.array_equal(ctx.wta(vol), np.argmax(vol, axis=3).astype(np.uint8))
    assert np.array_equal(ctx.wta(vol, take_min=True), np.argmin(vol, axis=3).astype(np.uint8))


@pytest.mark.parametrize("h,w,D", [(374, 1238, 201), (376, 1241, 228), (375, 1242, 201)])
def test_kitti_sizes_tensor_core(ctx, ws9, h, w, D):
    """The three image sizes of the shipped results (K8), reference-default and benchmark disparity ranges,
    through the tensor-core kernels: band symmetry, exact zero fill, fused WTA == WTA of the materialised
    volume, and the float64 oracle on a strip of rows."""
    l, r, gt = synth.make_pair(3000 + h, h, w, D)
    out = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=True)
    lv, rv = out["left_cost"][0], out["right_cost"][0]
    for d in (0, 1, D // 2, D - 1):
        assert np.array_equal(lv[:, d:, d], rv[:, :w - d, d])
        assert np.all(lv[:, :d, d] == 0) and np.all(rv[:, w - d:, d] == 0)
    assert np.array_equal(out["left_disp"][0], np.argmax(lv, axis=2).astype(np.uint8))
    assert np.array_equal(out["right_disp"][0], np.argmax(rv, axis=2).astype(np.uint8))
    fused = ctx.pairs(_lib.NET_WIN9, l, r, D, precision=_lib.PREC_BF16X3, volumes=False)
    assert np.array_equal(fused["left_disp"], out["left_disp"]) and np.array_equal(fused["right_disp"], out["right_disp"])
    lab = po.outlier_detection(fused["left_disp"][0].astype(np.float32), fused["right_disp"][0].astype(np.float32), D)
    assert np.array_equal(fused["outlier"][0], lab.astype(np.uint8))
    y0 = h - 8                                   # the last 8 rows: bottom padding included
    lp, rp = so.normalize_and_pad(l, 9), so.normalize_and_pad(r, 9)
    lm = so.trunk_forward(lp[:, y0:y0 + 16], ws9.layers, "win9_dep9", np.float64)
    rm = so.trunk_forward(rp[:, y0:y0 + 16], ws9.layers, "win9_dep9", np.float64)
    lref, rref = so.cost_volumes(lm, rm, D, np.float64)
    assert rel_err(lv[y0:], lref) < COST_TOL and rel_err(rv[y0:], rref) < COST_TOL
    assert (out["left_disp"][0][y0:] == so.wta(lref)).mean() >= WTA_AGREE


def test_two_contexts_on_one_gpu_are_independent(ws9_layers):
    """One context per thread is the threading model: two contexts driven from two threads at the
    same time give the results of a single one."""
    import threading
    D = 33
    pairs = [synth.make_pair(5200 + i, 60, 180, D) for i in range(4)]
    left = np.stack([p[0] for p in pairs])
    right = np.stack([p[1] for p in pairs])
    models = []
    for _ in range(2):
        m = base_model_ws_9((None, None, 3))
        m._ctx = _lib.Context(0)                      # a context of its own instead of the shared default
        m.set_layers(ws9_layers)
        models.append(m)
    ref = models[0].context.pairs(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3)
    outs = [None, None]

    def work(k):
        for _ in range(3):
            outs[k] = models[k].context.pairs(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3)

    threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for k in range(2):
        for name in ("left_disp", "right_disp", "outlier"):
            assert np.array_equal(outs[k][name], ref[name]), (k, name)
        models[k].context.close()


def test_batched_tensor_core_pairs_with_volumes_equal_single_pairs(ctx):
    """Pairs of a batch run on alternating streams with their own buffers: with the volumes exported too, every
    pair must equal the same pair processed alone."""
    D = 40
    pairs = [synth.make_pair(5300 + i, 50, 300, D) for i in range(3)]
    left = np.stack([p[0] for p in pairs])
    right = np.stack([p[1] for p in pairs])
    batch = ctx.pairs(_lib.NET_WIN9, left, right, D, precision=_lib.PREC_BF16X3, volumes=True)
    for i in range(3):
        one = ctx.pairs(_lib.NET_WIN9, left[i], right[i], D, precision=_lib.PREC_BF16X3, volumes=True)
        for name in ("left_disp", "right_disp", "outlier", "left_cost", "right_cost"):
            assert np.array_equal(batch[name][i], one[name][0]), (i, name)


@pytest.mark.parametrize("net,n,h,w", [(_lib.NET_WIN9, 9, 21, 300), (_lib.NET_WIN9, 11, 14, 131), (_lib.NET_WIN37, 9, 40, 150)])
def test_grouped_fused_correlation_equals_single_pairs(ctx, base_layers, net, n, h, w):
    """Fused path (no volumes): the features of up to four consecutive pairs stay in a lane's planar buffer
    (left and right of a pair as neighbouring images) and one correlation launch serves them.  Full groups on both
    lanes and a partial last group: every pair's maps must be those of a call with that pair alone."""
    D = 40
    pairs = [synth.make_pair(6100 + i, h, w, D) for i in range(n)]
    left, right = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    batch = ctx.pairs(net, left, right, D, precision=_lib.PREC_BF16X3, volumes=False)
    for i in range(n):
        one = ctx.pairs(net, left[i], right[i], D, precision=_lib.PREC_BF16X3, volumes=False)
        for name in ("left_disp", "right_disp", "outlier"):
            assert np.array_equal(batch[name][i], one[name][0]), (i, name)


def test_inference_cli_sharded_over_two_ranks(tmp_path):
    """`torchrun --nproc-per-node 2 -m stereo_matching_b200.inference`: the pair list is sharded round-robin over the
    ranks (shard.py; one GPU per rank, here both on the one GPU of the box), every rank writes the files of its pairs,
    rank 0 prints the mean over all of them - the same files and the same mean as a single process."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    golden = os.path.join(root, "tests", "golden", "new_checkpoint")
    common = ["--model_dir", golden, "--synthetic", "3", "--height", "40", "--width", "150", "--disp_range", "32"]
    outs = {}
    for name, launcher in (("one", [sys.executable, "-m", "stereo_matching_b200.inference"]),
                           ("two", [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                                    "--master-addr", "127.0.0.1", "--master-port", "29533", "-m",
                                    "stereo_matching_b200.inference"])):
        out_dir = tmp_path / name
        r = subprocess.run(launcher + common + ["--out_dir", str(out_dir)], capture_output=True, text=True, timeout=600, cwd=root)
        assert r.returncode == 0, r.stderr[-2000:]
        means = [float(l.split("Mean Error:")[1]) for l in r.stdout.splitlines() if "Mean Error:" in l]
        assert len(means) == 1                                   # rank 0 only
        outs[name] = (out_dir, means[0])
    assert outs["one"][1] == pytest.approx(outs["two"][1])
    for fid in range(3):
        for sub, pat in (("left_disp", "left_%06d_10.png"), ("right_disp", "right_%06d_10.png"), ("gt", "true_%06d_10.png")):
            a = cv2.imread(str(outs["one"][0] / sub / (pat % fid)), 0)
            b = cv2.imread(str(outs["two"][0] / sub / (pat % fid)), 0)
            assert a is not None and b is not None and np.array_equal(a, b), (sub, fid)
        name = "left_%06d_10_40_150_32.bin" % fid
        assert (outs["one"][0] / "left_cost" / name).read_bytes() == (outs["two"][0] / "left_cost" / name).read_bytes()
