"""A small KITTI-shaped tree (synthetic pairs, 16-bit ground-truth PNGs) for the real-data tests."""
import os

import cv2
import numpy as np

from stereo_matching_b200 import preprocess, synth

H, W, D, PSZ, HR = 40, 260, 41, 4, 20
N_FILES, TR, VAL = 5, 3, 2


def build(root):
    for sub in ("image_2", "image_3", "image_0", "image_1", "disp_noc_0"):
        os.makedirs(root / sub)
    gts = {}
    for i in range(N_FILES):
        l, r, gt = synth.make_pair(300 + i, H, W, D)
        cv2.imwrite(str(root / "image_2" / ("%06d_10.png" % i)), l[:, :, ::-1])
        cv2.imwrite(str(root / "image_3" / ("%06d_10.png" % i)), r[:, :, ::-1])
        cv2.imwrite(str(root / "image_0" / ("%06d_10.png" % i)), l[:, :, 1])
        cv2.imwrite(str(root / "image_1" / ("%06d_10.png" % i)), r[:, :, 1])
        # KITTI stores disparity * 256 in 16 bits; add a half-pixel fraction to some pixels to exercise the rounding
        frac = (np.arange(W)[None, :] % 3 == 0) * 128
        raw = gt.astype(np.uint16) * 256 + ((gt > 0) * frac).astype(np.uint16)
        cv2.imwrite(str(root / "disp_noc_0" / ("%06d_10.png" % i)), raw)
        gts[i] = raw
    util = root / "util"
    counts = preprocess.gene_loc(TR, VAL, PSZ, HR, str(util), 123, str(root), n_files=N_FILES)
    return {"root": str(root), "util": str(util), "raw": gts, "counts": counts}
