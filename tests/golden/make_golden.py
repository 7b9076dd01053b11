"""Regenerate tests/golden from the read-only reference checkout.

    python tests/golden/make_golden.py [/root/reference]

/root/reference does not exist on the GPU box, so the fixtures the tests need
are copied (data files only - checkpoint tensors and result PNGs, no source)
or derived here and committed:

  new_checkpoint/        the shipped TensorBundle checkpoint ckpt-1000 (weight source, K1)
  results/<stage>/       stage PNGs of three validation images, one per image size (K5, K6, K8)
  kat.json               known answers computed from ALL 40 shipped images with the
                         reference's own formulas (K2, K3, K4, K7, K9)
  oracle_small.npz       oracle outputs on one tiny synthetic pair, to detect oracle drift
"""
import glob
import json
import os
import shutil
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

KEEP_IDS = ["000002", "000169", "000192"]      # 375x1242, 374x1238, 376x1241
STAGES = {"gt": "true_%s_10.png", "left_disp": "left_%s_10.png", "right_disp": "right_%s_10.png",
          "nyu_cost_img_2": "%s_10.png", "right_nyu_cost_img_2": "%s_10.png",
          "outlier": "%s_10.png", "nyu_post": "%s_10.png"}


def three_px(pred, gt):   # inference.py:185-196 / cal.py:17-28
    valid = (gt != 0).astype("float")
    return float((np.abs(pred * valid - gt) > 3).sum() / valid.sum())


def main(ref):
    res = os.path.join(ref, "results")
    ck_dst = os.path.join(HERE, "new_checkpoint")
    os.makedirs(ck_dst, exist_ok=True)
    for name in ("checkpoint", "ckpt-1000.index", "ckpt-1000.data-00000-of-00001"):
        shutil.copyfile(os.path.join(ref, "new_checkpoint", name), os.path.join(ck_dst, name))
    for stage, pat in STAGES.items():
        os.makedirs(os.path.join(HERE, "results", stage), exist_ok=True)
        for i in KEEP_IDS:
            shutil.copyfile(os.path.join(res, stage, pat % i), os.path.join(HERE, "results", stage, pat % i))

    ids = sorted(os.path.basename(f)[5:11] for f in glob.glob(os.path.join(res, "gt", "*.png")))
    kat = {"ids": [int(i) for i in ids], "per_image": {}, "shapes": {}}
    means = {k: 0.0 for k in ("left_disp", "nyu_cost_img", "nyu_sgm_img", "nyu_cost_img_2", "nyu_post")}
    viol_l = viol_r = 0
    for i in ids:
        gt = cv2.imread(os.path.join(res, "gt", "true_%s_10.png" % i), cv2.IMREAD_GRAYSCALE)
        kat["shapes"][i] = list(gt.shape)
        for k in means:
            name = ("left_%s_10.png" if k == "left_disp" else "%s_10.png") % i
            pred = cv2.imread(os.path.join(res, k, name), cv2.IMREAD_GRAYSCALE)
            e = three_px(pred, gt)
            means[k] += e / len(ids)
            if i in KEEP_IDS:
                kat["per_image"].setdefault(i, {})[k] = e
        l = cv2.imread(os.path.join(res, "left_disp", "left_%s_10.png" % i), cv2.IMREAD_GRAYSCALE).astype(int)
        r = cv2.imread(os.path.join(res, "right_disp", "right_%s_10.png" % i), cv2.IMREAD_GRAYSCALE).astype(int)
        xs = np.arange(l.shape[1])[None, :]
        viol_l += int((l > xs).sum())
        viol_r += int((r > l.shape[1] - 1 - xs).sum())
    kat["mean_3px"] = means
    kat["wta_invalid_band_violations"] = {"left": viol_l, "right": viol_r}
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1, sort_keys=True)

    from oracle import stereo_oracle as so
    from stereo_matching_b200 import checkpoint, synth
    layers = checkpoint.load_net_weights(ck_dst)
    l, r, gt = synth.make_pair(4242, 20, 44, 12)
    o = so.infer_pair(l, r, layers, disp_range=12, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "oracle_small.npz"), left=l, right=r, gt=gt,
                        lmap=o["lmap"].astype(np.float32), left_cost=o["left_cost"].astype(np.float32),
                        right_cost=o["right_cost"].astype(np.float32),
                        left_disp=o["left_disp"], right_disp=o["right_disp"])
    print(json.dumps(means, indent=1))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
