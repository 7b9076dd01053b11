"""Real-data path: preprocess.py against a loop restatement of
kitti2015_gene_loc_1.m, data_handler.Data_handler against a loop restatement of data_handler.py:92-143, both on
a small KITTI-shaped tree written to a temporary directory (synthetic pairs, 16-bit ground-truth PNGs)."""
import os

import cv2
import numpy as np
import pytest

from stereo_matching_b200 import preprocess
from stereo_matching_b200.data_handler import Data_handler, soft_label_rows

from kitti_tree import H, W, D, PSZ, HR, N_FILES, TR, VAL, build


@pytest.fixture(scope="module")
def kitti(tmp_path_factory):
    return build(tmp_path_factory.mktemp("kitti"))


def _loop_locations(raw, file_id):
    """kitti2015_gene_loc_1.m:52-68 as written: column-major find, 1-based, round half away from zero."""
    disp = raw.astype(np.float64) / 256
    disp[raw == 0] = -1
    h, w = disp.shape
    out = []
    for c in range(1, w + 1):
        for r in range(1, h + 1):
            v = disp[r - 1, c - 1]
            if v == -1:
                continue
            t = c - v
            rx = np.floor(t + 0.5) if t >= 0 else -np.floor(-t + 0.5)
            ll = c + PSZ <= w and c - PSZ > 0 and r + PSZ <= h and r - PSZ > 0
            rr = rx - HR - PSZ > 0 and rx + HR + PSZ <= w and r - PSZ > 0 and r + PSZ <= h
            if ll and rr:
                out.append([file_id, 1, c, r, rx])
    return np.array(out, np.float32).reshape(-1, 5)


def test_preprocess_files_match_loop_restatement(kitti):
    perm = np.fromfile(os.path.join(kitti["util"], "myPerm.bin"), "<f4").astype(int)
    assert sorted(perm) == list(range(N_FILES))
    assert np.array_equal(perm, np.random.default_rng(123).permutation(N_FILES))
    for name, ids, count in (("tr", perm[:TR], kitti["counts"][0]), ("val", perm[TR:TR + VAL], kitti["counts"][1])):
        got = np.fromfile("%s/%s_%d_%d_%d.bin" % (kitti["util"], name, len(ids), PSZ, HR), "<f4").reshape(-1, 5)
        ref = np.concatenate([_loop_locations(kitti["raw"][i], i) for i in ids])
        assert count == len(ref) > 100
        assert np.array_equal(got, ref)


def test_disp_read_matches_reference_convention(kitti):
    d = preprocess.disp_read(os.path.join(kitti["root"], "disp_noc_0", "000000_10.png"))
    raw = kitti["raw"][0]
    assert d.dtype == np.float64 and np.all(d[raw == 0] == -1) and np.array_equal(d[raw > 0], raw[raw > 0] / 256)


def _loop_patches(dh, locs, val):
    """data_handler.py:92-101 (validation) / :135-142 (training) patch by patch."""
    hp, hr = dh.half_patch, dh.half_range
    left = np.zeros((len(locs), dh.patch_size, dh.patch_size, dh.num_channels), np.float32)
    right = np.zeros((len(locs), dh.patch_size, dh.patch_size + dh.disp_range - 1, dh.num_channels), np.float32)
    for i, (img_id, kind, cx, cy, rcx) in enumerate(locs):
        lp = dh.ldata[img_id][cy - hp:cy + hp + 1, cx - hp:cx + hp + 1, :]
        if kind == 1:
            left[i] = lp
            right[i] = dh.rdata[img_id][cy - hp:cy + hp + 1, rcx - hp - hr:rcx + hp + hr + 1, :]
        else:
            left[i] = lp if val else np.transpose(lp, (1, 0, 2))
            right[i] = np.transpose(dh.rdata[img_id][cy - hp - hr:cy + hp + hr + 1, rcx - hp:rcx + hp + 1, :], (1, 0, 2))
    return left, right


@pytest.mark.parametrize("version,channels", [("kitti2015", 3), ("kitti2012", 1)])
def test_data_handler_matches_loop_restatement(kitti, tmp_path, version, channels):
    np.random.seed(123)
    dh = Data_handler(version, kitti["root"], kitti["util"], TR, VAL, 50, 16, 2 * PSZ + 1, 2 * HR + 1,
                      val_npy=str(tmp_path / "val.npy"))
    assert dh.num_channels == channels
    # permutations drawn like the reference: two shuffles of the global generator after seed(123)
    np.random.seed(123)
    tp, vp = np.arange(len(dh.tr_loc)), np.arange(len(dh.val_loc))
    np.random.shuffle(tp)
    np.random.shuffle(vp)
    assert np.array_equal(dh.tr_perm, tp) and np.array_equal(dh.val_perm, vp)
    # images: per-image normalisation over all pixels and channels
    fn = int(dh.file_ids[0])
    assert abs(float(dh.ldata[fn].mean())) < 1e-5 and abs(float(dh.ldata[fn].std()) - 1) < 1e-5
    # validation set and val.npy
    vl, vr, vt = dh.evaluate()
    rl, rr = _loop_patches(dh, dh.val_loc[dh.val_perm[:50]], val=True)
    assert np.array_equal(vl, rl) and np.array_equal(vr, rr)
    assert np.array_equal(vt, soft_label_rows(50, 2 * HR + 1)) and vt[0, HR] == 0.5 and abs(vt[0].sum() - 1) < 1e-6
    assert np.array_equal(np.load(tmp_path / "val.npy"), np.unique(dh.val_loc[dh.val_perm[:50], 0]))
    # training batches, across the end of the location list
    n = len(dh.tr_perm)
    dh.tr_ptr = n - 21
    for expect_start, expect_epoch in ((n - 21, 0), (n - 5, 1), (11, 1)):
        assert dh.tr_ptr == expect_start
        bl, br, bt = dh.next_batch()
        order = [(expect_start + k) % n for k in range(16)]
        rl, rr = _loop_patches(dh, dh.tr_loc[dh.tr_perm[order]], val=False)
        assert np.array_equal(bl, rl) and np.array_equal(br, rr) and bt.shape == (16, 2 * HR + 1)
        assert dh.curr_epoch == expect_epoch or expect_start == n - 5
    assert dh.curr_epoch == 1 and dh.tr_ptr == 27


def test_data_handler_vertical_locations(kitti, tmp_path):
    """Type-2 (vertical) locations are commented out upstream but the handler still honours them."""
    np.random.seed(1)
    util = str(tmp_path / "util_hr5")
    preprocess.gene_loc(TR, VAL, PSZ, 5, util, 123, kitti["root"], n_files=N_FILES)
    dh = Data_handler("kitti2015", kitti["root"], util, TR, VAL, 4, 4, 2 * PSZ + 1, 11, val_npy="")
    fn = int(dh.file_ids[0])
    locs = np.array([[fn, 2, 100, 20, 90], [fn, 1, 120, 22, 100], [fn, 2, 60, 15, 55]])
    for val in (False, True):
        gl, gr = dh._gather(locs, val=val)
        rl, rr = _loop_patches(dh, locs, val=val)
        assert np.array_equal(gl, rl) and np.array_equal(gr, rr)


def test_validate_on_training_set_keeps_reference_double_shift(kitti):
    """num_val_img == 0: val_loc IS tr_loc and the reference subtracts 1 twice (data_handler.py:35-47)."""
    np.random.seed(5)
    raw = np.fromfile("%s/tr_%d_%d_%d.bin" % (kitti["util"], TR, PSZ, HR), "<f4").reshape(-1, 5).astype(int)
    # the double shift moves every validation patch one pixel up and left of its training twin
    dh = Data_handler("kitti2015", kitti["root"], kitti["util"], TR, 0, 8, 4, 2 * PSZ + 1, 2 * HR + 1, val_npy="")
    assert dh.val_loc is dh.tr_loc
    assert np.array_equal(dh.tr_loc[:, 2:], raw[:, 2:] - 2) and np.array_equal(dh.tr_loc[:, :2], raw[:, :2])


def test_out_of_image_location_is_an_error(kitti):
    np.random.seed(1)
    dh = Data_handler("kitti2015", kitti["root"], kitti["util"], TR, VAL, 4, 4, 2 * PSZ + 1, 2 * HR + 1, val_npy="")
    fn = int(dh.file_ids[0])
    for bad in ([fn, 1, 3, 20, 100], [fn, 1, 100, 20, 10], [fn, 1, 100, H - 2, 60], [fn, 1, W - 2, 20, 100]):
        with pytest.raises(ValueError):
            dh._gather(np.array([bad]))


def test_inference_reader_kitti2012_replicates_the_grey_channel(kitti):
    """inference.py:105-107,129-130: the grey image is broadcast into the 3-channel tower input."""
    import argparse
    from stereo_matching_b200.inference import _read_kitti
    flags = argparse.Namespace(data_root=kitti["root"], data_version="kitti2012")
    l, r, gt = _read_kitti(flags, 2)
    assert l.shape == (H, W, 3) and l.dtype == np.uint8 and gt is None       # no disp_noc/ in this tree
    assert np.array_equal(l[:, :, 0], l[:, :, 1]) and np.array_equal(l[:, :, 1], l[:, :, 2])
    grey = cv2.imread(os.path.join(kitti["root"], "image_0", "000002_10.png"), cv2.IMREAD_GRAYSCALE)
    assert np.array_equal(l[:, :, 0], grey)
    # same normalised input as normalising the grey image alone
    a, b = l.astype(np.float64), grey.astype(np.float64)
    assert abs(a.mean() - b.mean()) < 1e-12 and abs(a.std() - b.std()) < 1e-12
    flags.data_version = "kitti2015"
    l3, r3, gt3 = _read_kitti(flags, 2)
    raw = kitti["raw"][2]
    assert np.array_equal(gt3, (raw >> 8).astype(np.uint8))                 # cv2 GRAYSCALE of a 16-bit PNG = high byte
    assert np.array_equal(l3[:, :, 1], grey)


def test_preprocess_cli_reuses_an_existing_permutation(kitti, tmp_path):
    """`python -m stereo_matching_b200.preprocess --perm myPerm.bin`: location files regenerated for a given split."""
    out = tmp_path / "again"
    rc = preprocess.main(["--data_root", kitti["root"], "--save_dir", str(out), "--tr_num", str(TR), "--val_num", str(VAL),
                          "--psz", str(PSZ), "--half_range", str(HR), "--n_files", str(N_FILES),
                          "--perm", os.path.join(kitti["util"], "myPerm.bin")])
    assert rc == 0
    for name in ("myPerm.bin", "tr_%d_%d_%d.bin" % (TR, PSZ, HR), "val_%d_%d_%d.bin" % (VAL, PSZ, HR)):
        assert open(out / name, "rb").read() == open(os.path.join(kitti["util"], name), "rb").read()
