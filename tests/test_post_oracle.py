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


def _sgm_python(x0, x1, vol, direction, paths, p=po.NYU):
    """One-pixel-at-a-time Python loops over explicit paths (a second, independent restatement of the recurrence in
    SURVEY.md section 8(a) P3), float32 operation for operation."""
    f = np.float32
    h, w, D = vol.shape
    out = np.zeros_like(vol)
    dirs = [(1, 0), (-1, 0), (0, 1), (0, -1), (1, 1), (-1, 1), (1, -1), (-1, -1)][:paths]
    pi1, pi2, tau, a1, q1, q2 = (f(p[k]) for k in ("pi1", "pi2", "tau_so", "alpha1", "sgm_q1", "sgm_q2"))
    for sd, (dx, dy) in enumerate(dirs):
        Lr = np.zeros_like(vol)
        ys = range(h) if dy >= 0 else range(h - 1, -1, -1)
        xs = range(w) if dx >= 0 else range(w - 1, -1, -1)
        for y in ys:
            for x in xs:
                px, py = x - dx, y - dy
                if not (0 <= px < w and 0 <= py < h):
                    Lr[y, x] = vol[y, x]
                    continue
                prev = Lr[py, px]
                pmin = prev.min()
                D1 = abs(x0[y, x] - x0[py, px])
                for d in range(D):
                    xx = x + d * direction
                    if xx < 0 or xx >= w or xx - dx < 0 or xx - dx >= w:
                        D2 = f(10)
                    else:
                        D2 = abs(x1[y, xx] - x1[py, xx - dx])
                    if D1 < tau and D2 < tau:
                        P1, P2 = pi1, pi2
                    elif D1 > tau and D2 > tau:
                        P1, P2 = pi1 / (q1 * q2), pi2 / (q1 * q2)
                    else:
                        P1, P2 = pi1 / q1, pi2 / q1
                    c = min(prev[d], pmin + P2)
                    if d > 0:
                        c = min(c, prev[d - 1] + (P1 / a1 if sd == 2 else P1))
                    if d + 1 < D:
                        c = min(c, prev[d + 1] + (P1 / a1 if sd == 3 else P1))
                    Lr[y, x, d] = vol[y, x, d] + c - pmin
        out = out + Lr
    return out


@pytest.mark.parametrize("paths", [4, 8])
@pytest.mark.parametrize("direction", [-1, 1])
def test_sgm_c_restatement_equals_python_loops(paths, direction):
    """The C restatement (scanline order, OpenMP) against plain Python loops, bit for bit; `paths=8` is the
    extension named in BASELINE.json configs[3] - it has no counterpart in the reference (nyu.lua:59-64 is 4-path)."""
    rng = np.random.default_rng(paths + direction)
    h, w, D = 7, 9, 6
    x0 = (rng.standard_normal((h, w)) * 0.08).astype(np.float32)
    x1 = (rng.standard_normal((h, w)) * 0.08).astype(np.float32)
    vol = rng.standard_normal((h, w, D)).astype(np.float32)
    assert np.array_equal(po.sgm2(x0, x1, vol, direction, paths=paths), _sgm_python(x0, x1, vol, direction, paths))


def test_post_chain_optional_branches_follow_post_lua_order():
    """cost_agg_2 runs AFTER nyu_cost_agg_2 (post.lua:168-186 then :188-221): the box filter and CBCA do not commute."""
    rng = np.random.default_rng(3)
    h, w, D = 24, 40, 12
    lv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    rv = (rng.standard_normal((h, w, D)) * 10 + 20).astype(np.float32)
    ld, rd = lv.argmax(2).astype(np.uint8), rv.argmax(2).astype(np.uint8)
    base = po.post_chain(lv, rv, ld, rd)
    ext = po.post_chain(lv, rv, ld, rd, cost_agg=1, cost_agg_2=2)
    for name in ("nyu_cost_img", "nyu_sgm_img", "nyu_cost_img_2", "right_nyu_cost_img_2"):
        assert np.array_equal(base[name], ext[name]), name              # stages before the second box filter
    l_img, r_img = po.normalise_guide(ld), po.normalise_guide(rd)
    want = po.box_agg(np.transpose(base["left_vol"], (1, 2, 0)), 5, 5, 2)
    assert np.array_equal(np.transpose(ext["left_vol"], (1, 2, 0)), want)
    assert np.array_equal(ext["post_cost_img"], want.argmin(2).astype(np.uint8))
    assert np.array_equal(ext["cost_img"], po.box_agg(lv, 5, 5, 1).argmax(2).astype(np.uint8))
