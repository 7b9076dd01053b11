"""Pins for the CPU oracle of the inference path (oracle/stereo_oracle.py):
K2/K3/K4 metric known answers, K7/K8/K9 facts about the shipped maps, the
cost-volume indexing of inference.py:139-150 and a drift fixture."""
import json
import os

import cv2
import numpy as np
import pytest

from oracle import stereo_oracle as so
from stereo_matching_b200 import synth

from conftest import REFERENCE


def _kat(golden_dir):
    with open(os.path.join(golden_dir, "kat.json")) as f:
        return json.load(f)


def test_metric_known_answers_committed_images(golden_dir):
    kat = _kat(golden_dir)
    for i, vals in kat["per_image"].items():
        gt = cv2.imread(os.path.join(golden_dir, "results", "gt", "true_%s_10.png" % i), 0)
        l = cv2.imread(os.path.join(golden_dir, "results", "left_disp", "left_%s_10.png" % i), 0)
        p = cv2.imread(os.path.join(golden_dir, "results", "nyu_post", "%s_10.png" % i), 0)
        assert so.three_pixel_error(l, gt) == pytest.approx(vals["left_disp"], abs=1e-12)
        assert so.three_pixel_error(p, gt) == pytest.approx(vals["nyu_post"], abs=1e-12)
    assert kat["per_image"]["000002"]["left_disp"] == pytest.approx(0.337698, abs=1e-6)


def test_readme_numbers(golden_dir):
    kat = _kat(golden_dir)
    # README.md:62,76 "12.47" and "3.32" of the reference
    assert kat["mean_3px"]["left_disp"] == pytest.approx(0.12471779202002124, abs=1e-12)
    assert kat["mean_3px"]["nyu_post"] == pytest.approx(0.033224348286030894, abs=1e-12)
    assert kat["wta_invalid_band_violations"] == {"left": 0, "right": 0}      # K7
    assert len(kat["ids"]) == 40 and kat["ids"][:4] == [2, 5, 6, 14]          # K9
    assert sorted({tuple(s) for s in kat["shapes"].values()}) == [(374, 1238), (375, 1242), (376, 1241)]


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference checkout not mounted")
def test_readme_numbers_from_all_40_reference_pngs():
    res = os.path.join(REFERENCE, "results")
    tot = 0.0
    names = sorted(os.listdir(os.path.join(res, "gt")))
    for n in names:
        i = n.split("_")[1]
        gt = cv2.imread(os.path.join(res, "gt", n), 0)
        p = cv2.imread(os.path.join(res, "nyu_post", "%s_10.png" % i), 0)
        tot += so.three_pixel_error(p, gt)
    assert tot / len(names) == pytest.approx(0.033224348286030894, abs=1e-12)


def test_cost_volume_indexing_and_zero_fill():
    rng = np.random.default_rng(0)
    L = rng.standard_normal((1, 3, 10, 4))
    R = rng.standard_normal((1, 3, 10, 4))
    D = 6
    lv, rv = so.cost_volumes(L, R, D)
    for y in range(3):
        for x in range(10):
            for d in range(D):
                el = float(L[0, y, x] @ R[0, y, x - d]) if x >= d else 0.0
                er = float(L[0, y, x + d] @ R[0, y, x]) if x + d < 10 else 0.0
                assert lv[y, x, d] == pytest.approx(el, abs=1e-12)
                assert rv[y, x, d] == pytest.approx(er, abs=1e-12)
    assert lv[0, 2, 3] == 0.0 and rv[0, 9, 1] == 0.0


def test_disp_range_wider_than_image():
    L = np.ones((1, 2, 4, 2))
    lv, rv = so.cost_volumes(L, L, 7)
    assert lv.shape == (2, 4, 7) and np.all(lv[:, :, 4:] == 0)


def test_wta_zero_fill_and_first_max():
    # all valid scores negative -> the zero-filled slot x+1 wins
    L = np.ones((1, 1, 5, 2))
    R = -np.ones((1, 1, 5, 2))
    lv, rv = so.cost_volumes(L, R, 4)
    assert list(so.wta(lv)[0]) == [1, 2, 3, 0, 0]
    # ties -> lowest index
    v = np.zeros((1, 1, 3))
    assert so.wta(v)[0, 0] == 0


def test_normalize_and_pad():
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (7, 9, 3), dtype=np.uint8)
    out = so.normalize_and_pad(img, 9)
    assert out.shape == (1, 15, 17, 3) and out.dtype == np.float32
    inner = out[0, 4:11, 4:13]
    assert abs(float(inner.astype(np.float64).mean())) < 1e-6
    assert float(inner.astype(np.float64).std()) == pytest.approx(1.0, abs=1e-6)
    assert np.all(out[0, :4] == 0) and np.all(out[0, :, :4] == 0)


def test_trunk_shapes(ws9_layers):
    x = np.zeros((1, 12, 14, 3), np.float32)
    y = so.trunk_forward(x, ws9_layers, "win9_dep9", np.float32)
    assert y.shape == (1, 4, 6, 64)
    y37 = so.trunk_forward(np.zeros((1, 38, 40, 3), np.float32), synth.make_base_model_weights(), "win37_dep9", np.float32)
    assert y37.shape == (1, 2, 4, 64)


def test_oracle_drift_fixture(golden_dir, ws9_layers):
    g = np.load(os.path.join(golden_dir, "oracle_small.npz"))
    l, r, gt = synth.make_pair(4242, 20, 44, 12)
    assert np.array_equal(l, g["left"]) and np.array_equal(r, g["right"]) and np.array_equal(gt, g["gt"])
    o = so.infer_pair(l, r, ws9_layers, disp_range=12, dtype=np.float64)
    scale = np.abs(g["left_cost"]).max()
    assert np.abs(o["left_cost"] - g["left_cost"]).max() / scale < 1e-6
    assert np.abs(o["right_cost"] - g["right_cost"]).max() / scale < 1e-6
    assert np.array_equal(o["left_disp"], g["left_disp"])
    assert np.array_equal(o["right_disp"], g["right_disp"])
    o32 = so.infer_pair(l, r, ws9_layers, disp_range=12, dtype=np.float32)
    assert np.abs(o32["left_cost"] - g["left_cost"]).max() / scale < 1e-5


def test_synthetic_pair_ground_truth_is_consistent():
    l, r, gt = synth.make_pair(1000, 60, 200, 64)
    assert l.shape == (60, 200, 3) and l.dtype == np.uint8 and gt.dtype == np.uint8
    ys, xs = np.nonzero(gt)
    d = gt[ys, xs].astype(int)
    assert np.all(xs - d >= 0)
    # a visible pixel shows the same colour in both views
    assert np.array_equal(l[ys, xs], r[ys, xs - d])
    assert (gt > 0).mean() > 0.5


def test_banded_cost_volumes_equal_the_reference_loop():
    """The vectorised CPU variant bench.py times beside the reference's D-loop gives the same volumes."""
    rng = np.random.default_rng(5)
    L = rng.standard_normal((1, 3, 37, 64)).astype(np.float32)
    R = rng.standard_normal((1, 3, 37, 64)).astype(np.float32)
    for D in (6, 37, 50):
        lref, rref = so.cost_volumes(L, R, D, np.float64)
        lb, rb = so.cost_volumes_banded(L, R, D, np.float64)
        assert np.allclose(lb, lref, rtol=0, atol=1e-12) and np.allclose(rb, rref, rtol=0, atol=1e-12)
        assert np.array_equal(lb == 0, lref == 0)


def test_right_volume_is_the_sheared_left_volume_and_its_wta_is_a_column_scan():
    """The identity behind the one-sweep correlation kernel (csrc/corr_both.cu), on the reference loop's own
    output: right[x][d] = left[x + d][d] wherever x + d < W and 0.0 elsewhere (inference.py:160-183 writes both
    volumes from the same `inner`), so the right disparity of pixel x is the arg max over the rows x, x + 1, ... of the
    banded matrix S[xl][xr] = <L[xl], R[xr]> - with the zero-filled slots taking part, lowest disparity on ties."""
    rng = np.random.default_rng(5)
    W, D = 23, 9
    L = rng.integers(-2, 3, (1, 2, W, 3)).astype(np.float64)      # small integers: exact products, many ties
    R = rng.integers(-2, 3, (1, 2, W, 3)).astype(np.float64)
    lv, rv = so.cost_volumes(L, R, D)
    for y in range(2):
        S = L[0, y] @ R[0, y].T                                   # S[xl][xr]
        for x in range(W):
            for d in range(D):
                assert rv[y, x, d] == (lv[y, x + d, d] if x + d < W else 0.0)
                assert rv[y, x, d] == (S[x + d, x] if x + d < W else 0.0)
            # column scan with the running best seeded by nothing, first maximum wins; rows past the image are 0.0
            best, best_d = -np.inf, 0
            for d in range(D):
                v = S[x + d, x] if x + d < W else 0.0
                if v > best:
                    best, best_d = v, d
            assert so.wta(rv)[y, x] == best_d
