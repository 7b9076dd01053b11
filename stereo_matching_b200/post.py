"""Drop-in for /root/reference/post.lua and smooth/nyu.lua.

    python -m stereo_matching_b200.post --saveDir results --disp_range 201

walks `<saveDir>/left_cost/*.bin` like post.lua:78-92 (file name
left_<id>_10_<H>_<W>_<D>.bin), loads the matching right volume and the two WTA
PNGs post.lua uses as guide images, runs CBCA -> SGM -> CBCA -> LR check /
interpolation / sub-pixel / median / bilateral on the GPU and writes the same
stage PNGs (nyu_cost_img, right_nyu_cost_img, nyu_sgm_img, right_nyu_sgm_img,
nyu_cost_img_2, right_nyu_cost_img_2, nyu_post, outlier) as `%06d_10.png`.

`nyu` mirrors the Lua table `smooth.nyu`: same three functions, same argument
order, volumes in the reference's [1, D, H, W] layout (transposed on the host to
the library's [H, W, D]).
"""
from __future__ import annotations

import argparse
import os

import numpy as np

from . import _lib
from .models.model import default_context


def _hwd(vol_1dhw: np.ndarray) -> np.ndarray:
    v = np.asarray(vol_1dhw, np.float32)
    if v.ndim == 4:
        v = v[0]
    return np.ascontiguousarray(np.transpose(v, (1, 2, 0)))


def _1dhw(vol_hwd: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(np.transpose(vol_hwd, (2, 0, 1)))[None]


class _Nyu:
    """smooth.nyu of /root/reference/smooth/nyu.lua."""

    def __init__(self, device: int = 0):
        self.device = device
        self.params = _lib.default_post_params()

    @property
    def ctx(self):
        return default_context(self.device)

    def cross_agg(self, l_img, r_img, left_vol, right_vol, iter):        # nyu.lua:24
        l = np.asarray(l_img, np.float32).reshape(np.asarray(l_img).shape[-2:])
        r = np.asarray(r_img, np.float32).reshape(np.asarray(r_img).shape[-2:])
        lv, rv = self.ctx.nyu_cross_agg(l, r, _hwd(left_vol), _hwd(right_vol), iter, self.params)
        return _1dhw(lv), _1dhw(rv)

    def sgm(self, l_img, r_img, vol, direction):                          # nyu.lua:47
        l = np.asarray(l_img, np.float32).reshape(np.asarray(l_img).shape[-2:])
        r = np.asarray(r_img, np.float32).reshape(np.asarray(r_img).shape[-2:])
        return _1dhw(self.ctx.nyu_sgm(l, r, _hwd(vol), direction, self.params))

    def post(self, disp, left_vol):                                       # nyu.lua:91
        d0 = np.asarray(disp[0], np.float32).reshape(np.asarray(disp[0]).shape[-2:])
        d1 = np.asarray(disp[1], np.float32).reshape(np.asarray(disp[1]).shape[-2:])
        return self.ctx.nyu_post(d0, d1, _hwd(left_vol), self.params)


nyu = _Nyu()


def fromfile(fname: str) -> np.ndarray:
    """post.lua:57-63: float32 file, NaN -> 1e38."""
    x = np.fromfile(fname, "<f4")
    x[np.isnan(x)] = np.float32(1e38)
    return x


def build_parser():
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    # option names and defaults of post.lua:10-49 (unused reference options are accepted and ignored)
    p.add_argument("--model", default="split_win19_dep9")
    p.add_argument("-g", "--gpuid", type=int, default=0)
    p.add_argument("--feat_dim", type=int, default=64)
    p.add_argument("--data_version", default="kitti2015")
    p.add_argument("--data_root", default="")
    p.add_argument("--saveDir", default="results")
    p.add_argument("--disp_range", type=int, default=201)
    p.add_argument("--postprocess", type=int, default=1)
    p.add_argument("--sgm", type=int, default=1)
    p.add_argument("--post", type=int, default=1)
    p.add_argument("--nyu_cost_agg_1", type=int, default=1)
    p.add_argument("--nyu_cost_agg_2", type=int, default=1)
    p.add_argument("--cost_agg", type=int, default=0)
    p.add_argument("--cost_agg_2", type=int, default=0)
    p.add_argument("--confLoc", type=int, default=0)
    p.add_argument("--cost_w", type=int, default=5)
    p.add_argument("--cost_h", type=int, default=5)
    p.add_argument("--thres", type=float, default=23)
    p.add_argument("--in_dir", default=None, help="where left_cost/ etc. live (default: --saveDir, like post.lua's ./results)")
    return p


def run_staged(ctx, opt, params, unary, right_unary, l_u8, r_u8):
    """post.lua:95-243 stage by stage through the separate entry points - the path taken when one of the
    options that post.lua leaves off by default is set: `cost_agg` / `cost_agg_2` (box aggregation,
    post.lua:98-115,188-221) or `confLoc` (score threshold, post.lua:108-110,139-142).  Volumes are [H][W][D];
    every arithmetic step is the device kernel of the fused pipeline or the same fp32 operation in numpy."""
    l_img, r_img = ctx.guide_image(l_u8), ctx.guide_image(r_u8)           # post.lua:91-94
    st = {}
    lu, ru = np.asarray(unary, np.float32), np.asarray(right_unary, np.float32)
    if opt.cost_agg > 0:                                                   # post.lua:98-115
        cost_vol = ctx.box_agg(lu, opt.cost_w, opt.cost_h, opt.cost_agg)
        pred = ctx.wta(cost_vol)[0]
        if opt.confLoc == 1:
            pred = np.where(cost_vol.max(axis=2) < np.float32(opt.thres), 255, pred).astype(np.uint8)   # (256 - 1):byte()
        st["cost_img"] = pred
        if opt.nyu_cost_agg_1 <= 0:
            lu = cost_vol                                                  # post.lua:134 (the right volume stays raw)
    if opt.nyu_cost_agg_1 > 0:                                             # post.lua:116-132
        lu, ru = ctx.nyu_cross_agg(l_img, r_img, lu, ru, opt.nyu_cost_agg_1, params)
        st["nyu_cost_img"], st["right_nyu_cost_img"] = ctx.wta(lu)[0], ctx.wta(ru)[0]
    if opt.confLoc == 1:                                                   # post.lua:139-142
        lu = np.where(lu < np.float32(opt.thres), np.float32(0), lu)
        ru = np.where(ru < np.float32(opt.thres), np.float32(0), ru)
    lu = (lu / np.float32(ctx.volume_std(lu))).astype(np.float32)         # post.lua:146-147
    ru = (ru / np.float32(ctx.volume_std(ru))).astype(np.float32)
    if opt.sgm == 1:                                                       # post.lua:149-166
        lu = ctx.nyu_sgm(l_img, r_img, lu * np.float32(-1), -1, params)
        ru = ctx.nyu_sgm(l_img, r_img, ru * np.float32(-1), 1, params)
        st["nyu_sgm_img"], st["right_nyu_sgm_img"] = ctx.wta(lu, take_min=True)[0], ctx.wta(ru, take_min=True)[0]
    if opt.nyu_cost_agg_2 > 0:                                             # post.lua:168-185
        lu, ru = ctx.nyu_cross_agg(l_img, r_img, lu, ru, opt.nyu_cost_agg_2, params)
        st["nyu_cost_img_2"], st["right_nyu_cost_img_2"] = ctx.wta(lu, take_min=True)[0], ctx.wta(ru, take_min=True)[0]
    if opt.cost_agg_2 > 0:                                                 # post.lua:188-221
        lu = ctx.box_agg(lu, opt.cost_w, opt.cost_h, opt.cost_agg_2)
        ru = ctx.box_agg(ru, opt.cost_w, opt.cost_h, opt.cost_agg_2)
        st["post_cost_img"] = ctx.wta(lu, take_min=True)[0]
        st["right_post_cost_img"] = ctx.wta(ru)[0]                         # `ru:max(1)` in the original (post.lua:214)
    if opt.post == 1:                                                      # post.lua:223-242
        d0 = ctx.wta(lu, take_min=True)[0].astype(np.float32)
        d1 = ctx.wta(ru, take_min=True)[0].astype(np.float32)
        final, outlier = ctx.nyu_post(d0, d1, lu, params)
        st["nyu_post"] = final.astype(np.uint8)                            # :byte() truncation
        st["outlier"] = (outlier * 127).astype(np.uint8)
        st["nyu_post_float"] = final
    return st


def run(opt) -> int:
    import cv2
    in_dir = opt.in_dir or opt.saveDir
    ctx = default_context(opt.gpuid)
    params = _lib.default_post_params()
    params.sgm, params.post = opt.sgm, opt.post
    params.nyu_cost_agg_1, params.nyu_cost_agg_2 = opt.nyu_cost_agg_1, opt.nyu_cost_agg_2
    left_cost_url = os.path.join(in_dir, "left_cost")
    n = 0
    for file in sorted(os.listdir(left_cost_url)):
        if ".bin" not in file:
            continue
        s = file.split("_")                                     # post.lua:81-88
        img_h, img_w, d = int(s[3]), int(s[4]), int(s[5].split(".")[0])
        file_id = int(s[1])
        right_cost_path = "right_" + "_".join(s[1:])
        if d != opt.disp_range:
            raise SystemExit("%s holds D=%d but --disp_range is %d (post.lua views the file with opt.disp_range)"
                             % (file, d, opt.disp_range))
        unary = fromfile(os.path.join(left_cost_url, file)).reshape(img_h, img_w, d)
        right_unary = fromfile(os.path.join(in_dir, "right_cost", right_cost_path)).reshape(img_h, img_w, d)
        l_img = cv2.imread(os.path.join(in_dir, "left_disp", "left_%06d_10.png" % file_id), cv2.IMREAD_GRAYSCALE)
        r_img = cv2.imread(os.path.join(in_dir, "right_disp", "right_%06d_10.png" % file_id), cv2.IMREAD_GRAYSCALE)
        if opt.postprocess != 1:
            continue
        if opt.cost_agg or opt.cost_agg_2 or opt.confLoc:
            stages = run_staged(ctx, opt, params, unary, right_unary, l_img, r_img)     # non-default branches
        else:
            stages = ctx.post(unary, right_unary, l_img, r_img, params)
        for name in ("cost_img", "post_cost_img", "right_post_cost_img"):
            if name in stages:
                os.makedirs(os.path.join(opt.saveDir, name), exist_ok=True)
                cv2.imwrite(os.path.join(opt.saveDir, name, "%06d_10.png" % file_id), stages[name])
        for name in _lib.STAGE_NAMES:
            if name not in stages:
                continue
            skip = (name.endswith("nyu_cost_img") and not opt.nyu_cost_agg_1) or \
                   ("sgm" in name and not opt.sgm) or (name.endswith("_2") and not opt.nyu_cost_agg_2) or \
                   (name in ("nyu_post", "outlier") and not opt.post)
            if skip:
                continue
            os.makedirs(os.path.join(opt.saveDir, name), exist_ok=True)
            cv2.imwrite(os.path.join(opt.saveDir, name, "%06d_10.png" % file_id), stages[name])
        print("writing NYU post image done..")
        n += 1
    return n


def main(argv=None):
    run(build_parser().parse_args(argv))


if __name__ == "__main__":
    main()
