"""TEST INFRASTRUCTURE ONLY - CPU restatement of /root/reference/post.lua and
smooth/nyu.lua on top of `post_oracle.c` (the mc-cnn kernels restated in C).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module.  Parity status is described in post_oracle.c's header.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpost_oracle.so")
_lib = None

# smooth/nyu.lua:3-22
NYU = dict(L1=5, tau1=0.03, sgm_i=1, pi1=2.3, pi2=24.25, tau_so=0.08, alpha1=1.75,
           sgm_q1=3.0, sgm_q2=2.0, occlusion=1, subpixel=1, median=1, bilateral=1,
           blur_sigma=5.99, blur_t=5.0)


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or \
            os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "post_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def num_threads() -> int:
    return int(lib().oracle_num_threads())


# ---- thin wrappers, one per adcensus call site --------------------------------
def cross(img, L1=NYU["L1"], tau1=NYU["tau1"]):
    img = _f32(img)
    h, w = img.shape
    out = np.empty((4, h, w), np.float32)
    lib().oracle_cross(_fp(img), _fp(out), h, w, int(L1), ctypes.c_float(tau1))
    return out


def cbca(x0c, x1c, vol, direction):
    vol = _f32(vol)
    d, h, w = vol.shape
    out = np.empty_like(vol)
    lib().oracle_cbca(_fp(_f32(x0c)), _fp(_f32(x1c)), _fp(vol), _fp(out), d, h, w, int(direction))
    return out


def sgm2(x0, x1, vol_hwd, direction, p=NYU, paths=4):
    vol = _f32(vol_hwd)
    h, w, d = vol.shape
    out = np.zeros_like(vol)
    lib().oracle_sgm2_paths(_fp(_f32(x0)), _fp(_f32(x1)), _fp(vol), _fp(out), h, w, d,
                            ctypes.c_float(p["pi1"]), ctypes.c_float(p["pi2"]), ctypes.c_float(p["tau_so"]),
                            ctypes.c_float(p["alpha1"]), ctypes.c_float(p["sgm_q1"]),
                            ctypes.c_float(p["sgm_q2"]), int(direction), int(paths))
    return out


def outlier_detection(d0, d1, disp_max):
    d0, d1 = _f32(d0), _f32(d1)
    h, w = d0.shape
    out = np.zeros((h, w), np.float32)
    lib().oracle_outlier_detection(_fp(d0), _fp(d1), _fp(out), h, w, int(disp_max))
    return out


def interpolate_occlusion(d0, outlier):
    d0, outlier = _f32(d0), _f32(outlier)
    h, w = d0.shape
    out = np.empty_like(d0)
    lib().oracle_interpolate_occlusion(_fp(d0), _fp(outlier), _fp(out), h, w)
    return out


def interpolate_mismatch(d0, outlier):
    d0, outlier = _f32(d0), _f32(outlier)
    h, w = d0.shape
    out = np.empty_like(d0)
    lib().oracle_interpolate_mismatch(_fp(d0), _fp(outlier), _fp(out), h, w)
    return out


def subpixel_enchancement(d0, vol_dhw, disp_max):
    d0, vol = _f32(d0), _f32(vol_dhw)
    h, w = d0.shape
    out = np.empty_like(d0)
    lib().oracle_subpixel(_fp(d0), _fp(vol), _fp(out), h, w, int(disp_max))
    return out


def median2d(img, kernel_size=5):
    img = _f32(img)
    h, w = img.shape
    out = np.empty_like(img)
    lib().oracle_median2d(_fp(img), _fp(out), h, w, int(kernel_size))
    return out


def gaussian(sigma):
    """smooth/nyu.lua:77-89 (double precision, then :cuda() -> float32)."""
    kr = int(math.ceil(sigma * 3))
    ks = kr * 2 + 1
    k = np.empty((ks, ks), np.float64)
    for i in range(ks):
        for j in range(ks):
            y = i - kr
            x = j - kr
            k[i, j] = math.exp(-(x * x + y * y) / (2 * sigma * sigma))
    return k.astype(np.float32)


def mean2d(img, kernel, alpha2):
    img, kernel = _f32(img), _f32(kernel)
    h, w = img.shape
    out = np.empty_like(img)
    lib().oracle_mean2d(_fp(img), _fp(kernel), _fp(out), h, w, kernel.shape[0] // 2,
                        ctypes.c_float(alpha2))
    return out


# ---- smooth/nyu.lua ------------------------------------------------------------
def nyu_cross_agg(l_img, r_img, left_vol, right_vol, iters):
    """smooth/nyu.lua:24-45.  Volumes are [D,H,W]; both calls pass (x0c, x1c)."""
    x0c = cross(l_img)
    x1c = cross(r_img)
    for _ in range(iters):
        left_vol = cbca(x0c, x1c, left_vol, -1)
        right_vol = cbca(x0c, x1c, right_vol, 1)
    return left_vol, right_vol


def nyu_sgm(l_img, r_img, vol_dhw, direction, p=NYU, paths=4):
    """smooth/nyu.lua:47-75: permute to [H,W,D], sgm_i sweeps, result = out / 4.
    `paths=8` (not in the reference, see oracle_sgm2_paths) adds the four diagonals and divides by 8."""
    if paths not in (4, 8):
        raise ValueError("paths must be 4 or 8")
    vol = np.ascontiguousarray(np.transpose(vol_dhw, (1, 2, 0)))
    out = None
    for _ in range(p["sgm_i"]):
        out = sgm2(l_img, r_img, vol, direction, p, paths)
        vol = out / np.float32(paths)
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1))) / np.float32(paths)


def nyu_post(disp, left_vol_dhw, p=NYU):
    """smooth/nyu.lua:91-115. disp = [left, right] float maps (0-based)."""
    max_disp = left_vol_dhw.shape[0]
    d0 = _f32(disp[0])
    outlier = outlier_detection(d0, disp[1], max_disp)
    if p["occlusion"] == 1:
        d0 = interpolate_occlusion(d0, outlier)
        d0 = interpolate_mismatch(d0, outlier)
    if p["subpixel"] == 1:
        d0 = subpixel_enchancement(d0, left_vol_dhw, max_disp)
    if p["median"] == 1:
        d0 = median2d(d0, 5)
    if p["bilateral"] == 1:
        d0 = mean2d(d0, gaussian(p["blur_sigma"]), p["blur_t"])
    return d0, outlier


# ---- post.lua ------------------------------------------------------------------
def normalise_guide(img_u8):
    """post.lua:91-94: byte image -> float, subtract mean, divide by torch's
    UNBIASED std.  (THC accumulates in fp32 in an unspecified order; double
    accumulation rounded to fp32 is used here.)"""
    x = img_u8.astype(np.float32)
    mean = np.float32(x.astype(np.float64).mean())
    x = x + (-mean)
    std = np.float32(x.astype(np.float64).std(ddof=1))
    return x / std


def volume_std(vol):
    """post.lua:146-147 `lu:std()`: unbiased std over the whole volume."""
    return np.float32(vol.astype(np.float64).std(ddof=1))


def post_chain(left_cost_hwd, right_cost_hwd, left_disp_u8, right_disp_u8,
               nyu_cost_agg_1=1, sgm=1, nyu_cost_agg_2=1, post=1, p=NYU,
               cost_agg=0, cost_agg_2=0, confLoc=0, cost_w=5, cost_h=5, thres=23.0, sgm_paths=4):
    """post.lua:78-243.  Inputs are what inference.py wrote: the two [H,W,D] float32
    volumes and the two uint8 WTA maps, which post.lua loads as its guide images
    (post.lua:91-92).  Returns every stage PNG as a uint8 array plus the final float
    map.  The optional branches (`cost_agg`, `confLoc`, `cost_agg_2`; off by default in
    post.lua:37-49) run in post.lua's order: box aggregation of the left volume
    (:97-115), CBCA 1 (:116-137), threshold (:139-142), std (:146-147), SGM (:149-166),
    CBCA 2 (:168-186), box aggregation 2 (:188-221), nyu.post (:223-242).
    `sgm_paths=8` is NOT the reference (its chain is 4-path): see nyu_sgm."""
    stages = {}
    unary = _f32(left_cost_hwd)
    right_unary = _f32(right_cost_hwd)
    unary = np.where(np.isnan(unary), np.float32(1e38), unary)         # post.lua:57-63
    right_unary = np.where(np.isnan(right_unary), np.float32(1e38), right_unary)
    l_img = normalise_guide(left_disp_u8)
    r_img = normalise_guide(right_disp_u8)
    cost_vol = None
    if cost_agg > 0:                                                     # post.lua:97-115
        cost_vol = box_agg(unary, cost_w, cost_h, cost_agg)
        pred = np.argmax(cost_vol, axis=2)
        if confLoc == 1:
            pred = np.where(cost_vol.max(axis=2) < np.float32(thres), 255, pred)    # 256 - 1, :byte()
        stages["cost_img"] = pred.astype(np.uint8)
    if nyu_cost_agg_1 > 0:
        lu = np.ascontiguousarray(np.transpose(unary, (2, 0, 1)))      # post.lua:119-120
        ru = np.ascontiguousarray(np.transpose(right_unary, (2, 0, 1)))
        lu, ru = nyu_cross_agg(l_img, r_img, lu, ru, nyu_cost_agg_1)
        stages["nyu_cost_img"] = np.argmax(lu, axis=0).astype(np.uint8)      # post.lua:125-127
        stages["right_nyu_cost_img"] = np.argmax(ru, axis=0).astype(np.uint8)
    else:
        # post.lua:134-135 takes cost_vol / right_cost_vol; the latter is never assigned in post.lua, so the
        # reference only runs this branch for the left volume: the raw right volume stands in here
        lu = np.ascontiguousarray(np.transpose(cost_vol if cost_vol is not None else unary, (2, 0, 1)))
        ru = np.ascontiguousarray(np.transpose(right_unary, (2, 0, 1)))
    if confLoc == 1:                                                     # post.lua:139-142
        lu = np.where(lu < np.float32(thres), np.float32(0), lu)
        ru = np.where(ru < np.float32(thres), np.float32(0), ru)
    lu = lu / volume_std(lu)                                             # post.lua:146-147
    ru = ru / volume_std(ru)
    if sgm == 1:
        lu = lu * np.float32(-1)                                         # post.lua:152-153
        ru = ru * np.float32(-1)
        lu = nyu_sgm(l_img, r_img, lu, -1, p, sgm_paths)
        ru = nyu_sgm(l_img, r_img, ru, 1, p, sgm_paths)
        stages["nyu_sgm_img"] = np.argmin(lu, axis=0).astype(np.uint8)
        stages["right_nyu_sgm_img"] = np.argmin(ru, axis=0).astype(np.uint8)
    if nyu_cost_agg_2 > 0:                                               # post.lua:168-186
        lu, ru = nyu_cross_agg(l_img, r_img, lu, ru, nyu_cost_agg_2)
        stages["nyu_cost_img_2"] = np.argmin(lu, axis=0).astype(np.uint8)
        stages["right_nyu_cost_img_2"] = np.argmin(ru, axis=0).astype(np.uint8)
    if cost_agg_2 > 0:                                                   # post.lua:188-221
        lu = np.ascontiguousarray(np.transpose(box_agg(np.transpose(lu, (1, 2, 0)), cost_w, cost_h, cost_agg_2), (2, 0, 1)))
        ru = np.ascontiguousarray(np.transpose(box_agg(np.transpose(ru, (1, 2, 0)), cost_w, cost_h, cost_agg_2), (2, 0, 1)))
        stages["post_cost_img"] = np.argmin(lu, axis=0).astype(np.uint8)
        stages["right_post_cost_img"] = np.argmax(ru, axis=0).astype(np.uint8)     # `ru:max(1)`, post.lua:214
    if post == 1:
        disp = [np.argmin(lu, axis=0).astype(np.float32), np.argmin(ru, axis=0).astype(np.float32)]
        final, outlier = nyu_post(disp, lu, p)
        stages["nyu_post_float"] = final
        stages["nyu_post"] = final.astype(np.uint8)                      # :byte() truncation, post.lua:238
        stages["outlier"] = (outlier * 127).astype(np.uint8)             # post.lua:241
    stages["left_vol"] = lu
    stages["right_vol"] = ru
    return stages


def box_agg(vol, kw=5, kh=5, iters=1):
    """post.lua:98-106,188-203: `iters` passes of the count-exclude-pad average pooling; vol is [H][W][D]."""
    cur = _f32(vol)
    h, w, d = cur.shape
    for _ in range(int(iters)):
        out = np.empty_like(cur)
        lib().oracle_box_agg(_fp(cur), _fp(out), h, w, d, int(kw), int(kh))
        cur = out
    return cur
