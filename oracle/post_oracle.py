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


def sgm2(x0, x1, vol_hwd, direction, p=NYU):
    vol = _f32(vol_hwd)
    h, w, d = vol.shape
    out = np.zeros_like(vol)
    lib().oracle_sgm2(_fp(_f32(x0)), _fp(_f32(x1)), _fp(vol), _fp(out), h, w, d,
                      ctypes.c_float(p["pi1"]), ctypes.c_float(p["pi2"]), ctypes.c_float(p["tau_so"]),
                      ctypes.c_float(p["alpha1"]), ctypes.c_float(p["sgm_q1"]),
                      ctypes.c_float(p["sgm_q2"]), int(direction))
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


def nyu_sgm(l_img, r_img, vol_dhw, direction, p=NYU):
    """smooth/nyu.lua:47-75: permute to [H,W,D], sgm_i sweeps, result = out / 4."""
    vol = np.ascontiguousarray(np.transpose(vol_dhw, (1, 2, 0)))
    out = None
    for _ in range(p["sgm_i"]):
        out = sgm2(l_img, r_img, vol, direction, p)
        vol = out / np.float32(4)
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1))) / np.float32(4)


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
               nyu_cost_agg_1=1, sgm=1, nyu_cost_agg_2=1, post=1, p=NYU):
    """post.lua:78-243 with default options.  Inputs are what inference.py wrote:
    the two [H,W,D] float32 volumes and the two uint8 WTA maps, which post.lua
    loads as its guide images (post.lua:91-92).  Returns every stage PNG as a
    uint8 array plus the final float map."""
    stages = {}
    unary = _f32(left_cost_hwd)
    right_unary = _f32(right_cost_hwd)
    unary = np.where(np.isnan(unary), np.float32(1e38), unary)         # post.lua:57-63
    right_unary = np.where(np.isnan(right_unary), np.float32(1e38), right_unary)
    l_img = normalise_guide(left_disp_u8)
    r_img = normalise_guide(right_disp_u8)
    lu = np.ascontiguousarray(np.transpose(unary, (2, 0, 1)))          # post.lua:119-120
    ru = np.ascontiguousarray(np.transpose(right_unary, (2, 0, 1)))
    if nyu_cost_agg_1 > 0:
        lu, ru = nyu_cross_agg(l_img, r_img, lu, ru, nyu_cost_agg_1)
        stages["nyu_cost_img"] = np.argmax(lu, axis=0).astype(np.uint8)      # post.lua:125-127
        stages["right_nyu_cost_img"] = np.argmax(ru, axis=0).astype(np.uint8)
    lu = lu / volume_std(lu)                                             # post.lua:146-147
    ru = ru / volume_std(ru)
    if sgm == 1:
        lu = lu * np.float32(-1)                                         # post.lua:152-153
        ru = ru * np.float32(-1)
        lu = nyu_sgm(l_img, r_img, lu, -1, p)
        ru = nyu_sgm(l_img, r_img, ru, 1, p)
        stages["nyu_sgm_img"] = np.argmin(lu, axis=0).astype(np.uint8)
        stages["right_nyu_sgm_img"] = np.argmin(ru, axis=0).astype(np.uint8)
    if nyu_cost_agg_2 > 0:
        lu, ru = nyu_cross_agg(l_img, r_img, lu, ru, nyu_cost_agg_2)
        stages["nyu_cost_img_2"] = np.argmin(lu, axis=0).astype(np.uint8)
        stages["right_nyu_cost_img_2"] = np.argmin(ru, axis=0).astype(np.uint8)
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
