"""Pins for the CPU oracle of the post chain (oracle/post_oracle.{c,py}).

K5: outlier_detection is bit-exact against the PNG triples the reference ships.
K6: occlusion + mismatch fill + median + bilateral reproduce results/nyu_post to
    within +-1 on >= 99.5 % of pixels (the residue is the sub-pixel step, whose
    input volume is not shipped).
"""
import glob
import os

import cv2
import numpy as np
import pytest

from oracle import post_oracle as po

from conftest import REFERENCE


def _triples(root):
    ids = sorted(os.path.basename(f)[:6] for f in glob.glob(os.path.join(root, "outlier", "*.png")))
    for i in ids:
        yield (i,
               cv2.imread(os.path.join(root, "nyu_cost_img_2", "%s_10.png" % i), 0),
               cv2.imread(os.path.join(root, "right_nyu_cost_img_2", "%s_10.png" % i), 0),
               cv2.imread(os.path.join(root, "outlier", "%s_10.png" % i), 0))


def test_outlier_detection_bit_exact_committed(golden_dir):
    n = 0
    for i, l, r, o in _triples(os.path.join(golden_dir, "results")):
        out = po.outlier_detection(l.astype(np.float32), r.astype(np.float32), 201)
        assert np.array_equal((out * 127).astype(np.uint8), o), i
        n += 1
    assert n == 3


@pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="reference checkout not mounted")
def test_outlier_detection_bit_exact_all_40():
    n = px = 0
    for i, l, r, o in _triples(os.path.join(REFERENCE, "results")):
        out = po.outlier_detection(l.astype(np.float32), r.astype(np.float32), 201)
        assert np.array_equal((out * 127).astype(np.uint8), o), i
        n += 1
        px += o.size
    assert n == 40 and px == 18623378


def test_fill_median_bilateral_soft_pin(golden_dir):
    root = os.path.join(golden_dir, "results")
    for i, l, r, o in _triples(root):
        ref = cv2.imread(os.path.join(root, "nyu_post", "%s_10.png" % i), 0).astype(int)
        d0 = l.astype(np.float32)
        outlier = po.outlier_detection(d0, r.astype(np.float32), 201)
        d = po.interpolate_occlusion(d0, outlier)
        d = po.interpolate_mismatch(d, outlier)
        d = po.median2d(d, 5)
        d = po.mean2d(d, po.gaussian(5.99), 5.0)
        diff = np.abs(d.astype(np.uint8).astype(int) - ref)
        assert (diff <= 1).mean() >= 0.995, i
        assert diff.max() <= 8, i


def test_cross_arms_small():
    img = np.zeros((3, 12), np.float32)
    img[:, 8:] = 1.0
    a = po.cross(img, L1=5, tau1=0.03)
    # pixel (1,5): left arm runs to distance 4 (exclusive end 0); right arm stops
    # at the first differing pixel x=8 (exclusive end 8)
    assert a[0, 1, 5] == 0 and a[1, 1, 5] == 8
    # distance 1 is always accepted, even across an edge
    assert a[1, 1, 7] == 9
    # image border: exclusive end is -1 / H
    assert a[0, 1, 0] == -1 and a[2, 0, 3] == -1 and a[3, 2, 3] == 3


def test_cbca_constant_volume_is_fixed_point():
    rng = np.random.default_rng(0)
    img = rng.random((9, 20)).astype(np.float32)
    arms = po.cross(img)
    vol = np.full((5, 9, 20), 2.5, np.float32)
    for direction in (-1, 1):
        out = po.cbca(arms, arms, vol, direction)
        assert np.allclose(out, 2.5)


def test_sgm_first_pixels_and_constant_volume():
    h, w, d = 4, 6, 5
    img = np.zeros((h, w), np.float32)
    vol = np.zeros((h, w, d), np.float32)
    out = po.sgm2(img, img, vol, -1)
    assert np.all(out == 0)
    rng = np.random.default_rng(1)
    vol = rng.random((h, w, d)).astype(np.float32)
    out = po.sgm2(img, img, vol, -1)
    # each of the 4 paths adds C + (aggregated - min) >= C, and corner pixels
    # start two paths with exactly C
    assert np.all(out >= 4 * vol - 1e-5)
    assert out.shape == (h, w, d)


def test_subpixel_parabola():
    vol = np.zeros((5, 1, 1), np.float32)
    vol[:, 0, 0] = [4, 1, 0.5, 2, 9]
    d = po.subpixel_enchancement(np.array([[2.0]], np.float32), vol, 5)
    cn, cz, cp = 1.0, 0.5, 2.0
    assert d[0, 0] == pytest.approx(2 - (cp - cn) / (2 * (cp + cn - 2 * cz)))
    # border disparities are left alone
    assert po.subpixel_enchancement(np.array([[0.0]], np.float32), vol, 5)[0, 0] == 0
    assert po.subpixel_enchancement(np.array([[4.0]], np.float32), vol, 5)[0, 0] == 4


def test_median_border_uses_in_image_samples():
    img = np.arange(12, dtype=np.float32).reshape(3, 4)
    out = po.median2d(img, 5)
    # pixel (0,0): window covers rows 0-2, cols 0-2 -> 9 samples, element 4 of the sort
    assert out[0, 0] == np.sort(img[:3, :3].ravel())[4]


def test_full_chain_runs_and_improves(ws9_layers):
    from oracle import stereo_oracle as so
    from stereo_matching_b200 import synth
    l, r, gt = synth.make_pair(1001, 60, 200, 32)
    o = so.infer_pair(l, r, ws9_layers, disp_range=32, dtype=np.float32)
    st = po.post_chain(o["left_cost"], o["right_cost"], o["left_disp"], o["right_disp"])
    e0 = so.three_pixel_error(o["left_disp"], gt)
    e1 = so.three_pixel_error(st["nyu_post"], gt)
    assert e1 < e0
    assert set(np.unique(st["outlier"])) <= {0, 127, 254}


def test_box_agg_is_count_exclude_pad_average_pooling():
    """post.lua:98-106: nn.SpatialAveragePooling(5, 5, 1, 1, 2, 2):setCountExcludePad() == torch avg_pool2d with
    count_include_pad=False (independent implementation; summation order differs, hence allclose)."""
    import torch
    rng = np.random.default_rng(11)
    vol = rng.standard_normal((13, 17, 6)).astype(np.float32)
    for kw, kh in ((5, 5), (3, 7), (1, 1)):
        got = po.box_agg(vol, kw, kh)
        ref = torch.nn.functional.avg_pool2d(torch.tensor(vol).permute(2, 0, 1)[None], (kh, kw), 1, ((kh - 1) // 2, (kw - 1) // 2),
                                             count_include_pad=False)[0].permute(1, 2, 0).numpy()
        assert np.allclose(got, ref, rtol=1e-5, atol=1e-6)
    assert np.array_equal(po.box_agg(vol, 1, 1), vol)
    twice = po.box_agg(vol, 5, 5, iters=2)
    assert np.array_equal(twice, po.box_agg(po.box_agg(vol, 5, 5), 5, 5))
    const = np.full((9, 9, 3), 2.5, np.float32)
    assert np.array_equal(po.box_agg(const, 5, 5, iters=3), const)        # borders divide by the in-image count
