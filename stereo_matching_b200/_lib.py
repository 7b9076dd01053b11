"""ctypes binding of libstereo_b200.so (include/stereo_b200.h).

There is deliberately no fallback: if the CUDA library is missing or fails,
every call raises.  Nothing here imports torch; buffers are numpy arrays
(host entry points) or raw device addresses (`*_dev` entry points).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# STEREO_B200_LIB: another build of the library (A/B runs of compile-time variants, scripts/build_variant.py)
LIB_PATH = os.environ.get("STEREO_B200_LIB") or os.path.join(HERE, "libstereo_b200.so")

NET_WIN9, NET_WIN37 = 0, 1
PREC_FP32, PREC_BF16X3, PREC_BF16 = 0, 1, 2
NET_BY_NAME = {"win9_dep9": NET_WIN9, "win37_dep9": NET_WIN37}

_c_ctx = ctypes.c_void_p
_i = ctypes.c_int
_p = ctypes.c_void_p



class PostParams(ctypes.Structure):
    """stereo_post_params: smooth/nyu.lua:3-22 constants + post.lua:37-40 options."""
    _fields_ = [("L1", _i), ("tau1", ctypes.c_float), ("sgm_i", _i),
                ("pi1", ctypes.c_float), ("pi2", ctypes.c_float), ("tau_so", ctypes.c_float),
                ("alpha1", ctypes.c_float), ("sgm_q1", ctypes.c_float), ("sgm_q2", ctypes.c_float),
                ("occlusion", _i), ("subpixel", _i), ("median", _i), ("bilateral", _i),
                ("blur_sigma", ctypes.c_double), ("blur_t", ctypes.c_float),
                ("nyu_cost_agg_1", _i), ("sgm", _i), ("nyu_cost_agg_2", _i), ("post", _i), ("cbca_exact", _i), ("sgm_paths", _i)]


_pp = ctypes.POINTER(PostParams)
STAGE_NAMES = ("nyu_cost_img", "right_nyu_cost_img", "nyu_sgm_img", "right_nyu_sgm_img",
               "nyu_cost_img_2", "right_nyu_cost_img_2", "nyu_post", "outlier")

# name -> argtypes (restype int unless listed in _RESTYPES); kept in one place so that
# tests can check that every symbol the header declares is exported and bound
SIGNATURES = {
    "stereo_last_error": [],
    "stereo_version": [],
    "stereo_create": [_i, ctypes.POINTER(_c_ctx)],
    "stereo_destroy": [_c_ctx],
    "stereo_set_stream": [_c_ctx, _p],
    "stereo_sync": [_c_ctx],
    "stereo_launch_count": [_c_ctx],
    "stereo_set_weights": [_c_ctx, _i, _p, ctypes.c_size_t],
    "stereo_normalize_pad": [_c_ctx, _p, _i, _i, _i, _i, _p],
    "stereo_normalize_pad_dev": [_c_ctx, _p, _i, _i, _i, _i, _p],
    "stereo_trunk": [_c_ctx, _i, _p, _i, _i, _i, _p, _i],
    "stereo_trunk_dev": [_c_ctx, _i, _p, _i, _i, _i, _p, _i],
    "stereo_cost_volume": [_c_ctx, _p, _p, _i, _i, _i, _i, _i, _p, _p, _i],
    "stereo_cost_volume_dev": [_c_ctx, _p, _p, _i, _i, _i, _i, _i, _p, _p, _i],
    "stereo_wta": [_c_ctx, _p, _i, _i, _i, _i, _i, _p],
    "stereo_wta_dev": [_c_ctx, _p, _i, _i, _i, _i, _i, _p],
    "stereo_corr_wta": [_c_ctx, _p, _p, _i, _i, _i, _i, _i, _p, _p, _i],
    "stereo_corr_wta_dev": [_c_ctx, _p, _p, _i, _i, _i, _i, _i, _p, _p, _i],
    "stereo_lr_check": [_c_ctx, _p, _p, _i, _i, _i, _i, _p],
    "stereo_lr_check_dev": [_c_ctx, _p, _p, _i, _i, _i, _i, _p],
    "stereo_pairs": [_c_ctx, _i, _p, _p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p],
    "stereo_pairs_dev": [_c_ctx, _i, _p, _p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p],
    "stereo_post_default_params": [_pp],
    "stereo_guide_image": [_c_ctx, _p, _i, _i, _p],
    "stereo_nyu_cross_agg": [_c_ctx, _p, _p, _p, _p, _i, _i, _i, _i, _pp],
    "stereo_nyu_sgm": [_c_ctx, _p, _p, _p, _i, _i, _i, _i, _pp],
    "stereo_nyu_post": [_c_ctx, _p, _p, _p, _i, _i, _i, _pp, _p, _p],
    "stereo_volume_std": [_c_ctx, _p, ctypes.c_size_t, _p],
    "stereo_volume_std_static": [_p, ctypes.c_size_t, _p, _p],
    "stereo_post": [_c_ctx, _p, _p, _p, _p, _i, _i, _i, _pp, ctypes.POINTER(_p), _p],
    "stereo_post_dev": [_c_ctx, _p, _p, _p, _p, _i, _i, _i, _pp, ctypes.POINTER(_p), _p],
    "stereo_profile_enable": [_c_ctx, _i],
    "stereo_profile_read": [_c_ctx, _i, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)],
    "stereo_profile_name": [_i],
    "stereo_pairs_post": [_c_ctx, _i, _p, _p, _i, _i, _i, _i, _i, _pp, _p, _p, _p],
    "stereo_pairs_post_dev": [_c_ctx, _i, _p, _p, _i, _i, _i, _i, _i, _pp, _p, _p, _p],
    "stereo_box_agg": [_c_ctx, _p, _i, _i, _i, _i, _i, _i, _p],
    "stereo_box_agg_dev": [_c_ctx, _p, _i, _i, _i, _i, _i, _i, _p, _p],
    "stereo_train_create": [_c_ctx, _i, _i, _i, ctypes.POINTER(ctypes.c_void_p)],
    "stereo_train_destroy": [_p],
    "stereo_train_param_count": [_p],
    "stereo_train_stat_count": [_p],
    "stereo_train_patch_size": [_p],
    "stereo_train_set_state": [_p, _p, _p, _p, _p, ctypes.c_longlong],
    "stereo_train_get_state": [_p, _p, _p, _p, _p, ctypes.POINTER(ctypes.c_longlong)],
    "stereo_train_forward_backward": [_p, _p, _p, _p, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(_i), _p],
    "stereo_train_grads_dev": [_p, ctypes.POINTER(ctypes.c_void_p)],
    "stereo_train_apply": [_p, _p, ctypes.c_float],
}
_RESTYPES = {"stereo_last_error": ctypes.c_char_p, "stereo_launch_count": ctypes.c_int64,
             "stereo_profile_name": ctypes.c_char_p,
             "stereo_train_param_count": ctypes.c_size_t, "stereo_train_stat_count": ctypes.c_size_t}
KERNEL_CLASSES = 10

_lib = None


class StereoError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libstereo_b200 error {code}: {message}")
        self.code = code


def load():
    """dlopen the library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing - build it with `python -m stereo_matching_b200.build`; "
                "there is no CPU fallback")
        lib = ctypes.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = _RESTYPES.get(name, ctypes.c_int)
        _lib = lib
    return _lib


def check(rc: int):
    if rc != 0:
        raise StereoError(rc, load().stereo_last_error().decode("utf-8", "replace"))


def _ptr(a):
    """numpy array -> host address; int -> device address; None -> NULL."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return a.ctypes.data
    return int(a)


def _arr(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def default_post_params() -> PostParams:
    p = PostParams()
    check(load().stereo_post_default_params(ctypes.byref(p)))
    return p


class Context:
    """One `stereo_ctx`: one GPU, one stream.  Not thread-safe."""

    def __init__(self, device: int = 0):
        self._lib = load()
        h = _c_ctx()
        check(self._lib.stereo_create(int(device), ctypes.byref(h)))
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.stereo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- plumbing --
    def set_stream(self, cuda_stream: int | None):
        check(self._lib.stereo_set_stream(self._h, cuda_stream))

    def sync(self):
        check(self._lib.stereo_sync(self._h))

    @property
    def launch_count(self) -> int:
        return int(self._lib.stereo_launch_count(self._h))

    def profile_enable(self, on: bool = True):
        check(self._lib.stereo_profile_enable(self._h, int(bool(on))))

    def profile_read(self):
        """{kernel class name: (total_ms, launches)} since profile_enable(True)."""
        out = {}
        for k in range(KERNEL_CLASSES):
            ms, cnt = ctypes.c_double(), ctypes.c_int64()
            check(self._lib.stereo_profile_read(self._h, k, ctypes.byref(ms), ctypes.byref(cnt)))
            out[self._lib.stereo_profile_name(k).decode()] = (ms.value, cnt.value)
        return out

    def set_weights(self, net: int, packed: np.ndarray):
        packed = _arr(packed, np.float32)
        check(self._lib.stereo_set_weights(self._h, int(net), packed.ctypes.data, packed.size))

    # -- host-pointer entry points (numpy in, numpy out) --
    def normalize_pad(self, rgb: np.ndarray, patch: int) -> np.ndarray:
        rgb = _arr(rgb, np.uint8)
        if rgb.ndim == 3:
            rgb = rgb[None]
        n, h, w, c = rgb.shape
        if c != 3:
            raise ValueError("expected [n,h,w,3] uint8")
        out = np.empty((n, h + patch - 1, w + patch - 1, 3), np.float32)
        check(self._lib.stereo_normalize_pad(self._h, rgb.ctypes.data, n, h, w, int(patch), out.ctypes.data))
        return out

    def trunk(self, net: int, x: np.ndarray, precision: int = PREC_FP32) -> np.ndarray:
        x = _arr(x, np.float32)
        if x.ndim != 4 or x.shape[3] != 3:
            raise ValueError("expected [n,h,w,3] float32")
        n, h, w, _ = x.shape
        shrink = 8 if net == NET_WIN9 else 36
        out = np.empty((n, max(h - shrink, 0), max(w - shrink, 0), 64), np.float32)
        check(self._lib.stereo_trunk(self._h, int(net), x.ctypes.data, n, h, w, out.ctypes.data, int(precision)))
        return out

    def cost_volume(self, lmap: np.ndarray, rmap: np.ndarray, disp_range: int, precision: int = PREC_FP32):
        lmap, rmap = _arr(lmap, np.float32), _arr(rmap, np.float32)
        if lmap.shape != rmap.shape or lmap.ndim != 4:
            raise ValueError("expected two [n,h,w,c] float32 maps of equal shape")
        n, h, w, c = lmap.shape
        lv = np.empty((n, h, w, disp_range), np.float32)
        rv = np.empty((n, h, w, disp_range), np.float32)
        check(self._lib.stereo_cost_volume(self._h, lmap.ctypes.data, rmap.ctypes.data, n, h, w, c,
                                           int(disp_range), lv.ctypes.data, rv.ctypes.data, int(precision)))
        return lv, rv

    def wta(self, vol: np.ndarray, take_min: bool = False) -> np.ndarray:
        vol = _arr(vol, np.float32)
        if vol.ndim == 3:
            vol = vol[None]
        n, h, w, d = vol.shape
        out = np.empty((n, h, w), np.uint8)
        check(self._lib.stereo_wta(self._h, vol.ctypes.data, n, h, w, d, int(bool(take_min)), out.ctypes.data))
        return out

    def corr_wta(self, lmap: np.ndarray, rmap: np.ndarray, disp_range: int, precision: int = PREC_FP32):
        lmap, rmap = _arr(lmap, np.float32), _arr(rmap, np.float32)
        if lmap.shape != rmap.shape or lmap.ndim != 4:
            raise ValueError("expected two [n,h,w,c] float32 maps of equal shape")
        n, h, w, c = lmap.shape
        ld = np.empty((n, h, w), np.uint8)
        rd = np.empty((n, h, w), np.uint8)
        check(self._lib.stereo_corr_wta(self._h, lmap.ctypes.data, rmap.ctypes.data, n, h, w, c,
                                        int(disp_range), ld.ctypes.data, rd.ctypes.data, int(precision)))
        return ld, rd

    def lr_check(self, d0: np.ndarray, d1: np.ndarray, disp_range: int) -> np.ndarray:
        d0, d1 = _arr(d0, np.uint8), _arr(d1, np.uint8)
        if d0.ndim == 2:
            d0, d1 = d0[None], d1[None]
        n, h, w = d0.shape
        out = np.empty((n, h, w), np.uint8)
        check(self._lib.stereo_lr_check(self._h, d0.ctypes.data, d1.ctypes.data, n, h, w, int(disp_range),
                                        out.ctypes.data))
        return out

    def pairs(self, net: int, left_rgb: np.ndarray, right_rgb: np.ndarray, disp_range: int,
              precision: int = PREC_FP32, lr_check: bool = True, volumes: bool = False):
        left_rgb, right_rgb = _arr(left_rgb, np.uint8), _arr(right_rgb, np.uint8)
        if left_rgb.ndim == 3:
            left_rgb, right_rgb = left_rgb[None], right_rgb[None]
        if left_rgb.shape != right_rgb.shape or left_rgb.shape[3] != 3:
            raise ValueError("expected two [n,h,w,3] uint8 batches of equal shape")
        n, h, w, _ = left_rgb.shape
        out = {"left_disp": np.empty((n, h, w), np.uint8), "right_disp": np.empty((n, h, w), np.uint8)}
        if lr_check:
            out["outlier"] = np.empty((n, h, w), np.uint8)
        if volumes:
            out["left_cost"] = np.empty((n, h, w, disp_range), np.float32)
            out["right_cost"] = np.empty((n, h, w, disp_range), np.float32)
        check(self._lib.stereo_pairs(self._h, int(net), left_rgb.ctypes.data, right_rgb.ctypes.data, n, h, w,
                                     int(disp_range), int(precision), out["left_disp"].ctypes.data,
                                     out["right_disp"].ctypes.data, _ptr(out.get("outlier")),
                                     _ptr(out.get("left_cost")), _ptr(out.get("right_cost"))))
        return out

    # -- post chain (post.lua / smooth/nyu.lua); volumes are [h][w][D] float32 --
    def guide_image(self, img_u8: np.ndarray) -> np.ndarray:
        img = _arr(img_u8, np.uint8)
        h, w = img.shape
        out = np.empty((h, w), np.float32)
        check(self._lib.stereo_guide_image(self._h, img.ctypes.data, h, w, out.ctypes.data))
        return out

    def box_agg(self, vol: np.ndarray, kw: int = 5, kh: int = 5, iters: int = 1) -> np.ndarray:
        vol = _arr(vol, np.float32)
        h, w, d = vol.shape
        out = np.empty_like(vol)
        check(self._lib.stereo_box_agg(self._h, vol.ctypes.data, h, w, d, int(kw), int(kh), int(iters), out.ctypes.data))
        return out

    def volume_std(self, vol: np.ndarray) -> float:
        vol = _arr(vol, np.float32)
        out = np.empty(1, np.float32)
        check(self._lib.stereo_volume_std(self._h, vol.ctypes.data, vol.size, out.ctypes.data))
        return out[0]

    def nyu_cross_agg(self, l_img, r_img, left_vol, right_vol, iters: int, params: "PostParams | None" = None):
        params = params or default_post_params()
        l_img, r_img = _arr(l_img, np.float32), _arr(r_img, np.float32)
        lv, rv = np.array(left_vol, np.float32, order="C"), np.array(right_vol, np.float32, order="C")
        h, w, d = lv.shape
        check(self._lib.stereo_nyu_cross_agg(self._h, l_img.ctypes.data, r_img.ctypes.data, lv.ctypes.data,
                                             rv.ctypes.data, h, w, d, int(iters), ctypes.byref(params)))
        return lv, rv

    def nyu_sgm(self, l_img, r_img, vol, direction: int, params: "PostParams | None" = None):
        params = params or default_post_params()
        l_img, r_img = _arr(l_img, np.float32), _arr(r_img, np.float32)
        v = np.array(vol, np.float32, order="C")
        h, w, d = v.shape
        check(self._lib.stereo_nyu_sgm(self._h, l_img.ctypes.data, r_img.ctypes.data, v.ctypes.data, h, w, d,
                                       int(direction), ctypes.byref(params)))
        return v

    def nyu_post(self, disp_left, disp_right, left_vol, params: "PostParams | None" = None):
        params = params or default_post_params()
        d0, d1 = _arr(disp_left, np.float32), _arr(disp_right, np.float32)
        v = _arr(left_vol, np.float32)
        h, w, d = v.shape
        final = np.empty((h, w), np.float32)
        outlier = np.empty((h, w), np.float32)
        check(self._lib.stereo_nyu_post(self._h, d0.ctypes.data, d1.ctypes.data, v.ctypes.data, h, w, d,
                                        ctypes.byref(params), final.ctypes.data, outlier.ctypes.data))
        return final, outlier

    def post(self, left_cost, right_cost, left_disp, right_disp, params: "PostParams | None" = None):
        """post.lua:78-243 for one pair -> {stage name: uint8 map} + 'nyu_post_float'."""
        params = params or default_post_params()
        lv, rv = _arr(left_cost, np.float32), _arr(right_cost, np.float32)
        ld, rd = _arr(left_disp, np.uint8), _arr(right_disp, np.uint8)
        h, w, d = lv.shape
        stages = {n: np.empty((h, w), np.uint8) for n in STAGE_NAMES}
        ptrs = (_p * 8)(*[stages[n].ctypes.data for n in STAGE_NAMES])
        final = np.empty((h, w), np.float32)
        check(self._lib.stereo_post(self._h, lv.ctypes.data, rv.ctypes.data, ld.ctypes.data, rd.ctypes.data,
                                    h, w, d, ctypes.byref(params), ptrs, final.ctypes.data))
        stages["nyu_post_float"] = final
        return stages

    def pairs_post(self, net: int, left_rgb, right_rgb, disp_range: int, precision: int = PREC_FP32,
                   params: "PostParams | None" = None):
        """inference.py + post.lua for a batch of pairs, volumes kept on the device -> final maps."""
        params = params or default_post_params()
        left_rgb, right_rgb = _arr(left_rgb, np.uint8), _arr(right_rgb, np.uint8)
        if left_rgb.ndim == 3:
            left_rgb, right_rgb = left_rgb[None], right_rgb[None]
        n, h, w, _ = left_rgb.shape
        out = {k: np.empty((n, h, w), np.uint8) for k in ("nyu_post", "outlier", "left_disp")}
        check(self._lib.stereo_pairs_post(self._h, int(net), left_rgb.ctypes.data, right_rgb.ctypes.data, n, h, w,
                                          int(disp_range), int(precision), ctypes.byref(params), out["nyu_post"].ctypes.data,
                                          out["outlier"].ctypes.data, out["left_disp"].ctypes.data))
        return out

    def post_dev(self, left_cost, right_cost, left_disp, right_disp, h, w, disp_range, stage_ptrs,
                 final=None, params: "PostParams | None" = None):
        params = params or default_post_params()
        ptrs = (_p * 8)(*[(_ptr(s) if s is not None else None) for s in stage_ptrs])
        check(self._lib.stereo_post_dev(self._h, _ptr(left_cost), _ptr(right_cost), _ptr(left_disp),
                                        _ptr(right_disp), h, w, int(disp_range), ctypes.byref(params), ptrs,
                                        _ptr(final)))

    # -- device-pointer entry points (raw addresses; asynchronous) --
    def pairs_dev(self, net, left_rgb, right_rgb, n, h, w, disp_range, precision, left_disp, right_disp,
                  outlier=None, left_vol=None, right_vol=None):
        check(self._lib.stereo_pairs_dev(self._h, int(net), _ptr(left_rgb), _ptr(right_rgb), n, h, w,
                                         int(disp_range), int(precision), _ptr(left_disp), _ptr(right_disp),
                                         _ptr(outlier), _ptr(left_vol), _ptr(right_vol)))

    def pairs_post_dev(self, net, left_rgb, right_rgb, n, h, w, disp_range, precision, params, nyu_post, outlier=None,
                       left_wta=None):
        check(self._lib.stereo_pairs_post_dev(self._h, int(net), _ptr(left_rgb), _ptr(right_rgb), n, h, w, int(disp_range),
                                              int(precision), ctypes.byref(params), _ptr(nyu_post), _ptr(outlier),
                                              _ptr(left_wta)))

    def normalize_pad_dev(self, rgb, n, h, w, patch, out):
        check(self._lib.stereo_normalize_pad_dev(self._h, _ptr(rgb), n, h, w, int(patch), _ptr(out)))

    def trunk_dev(self, net, x, n, h, w, out, precision=PREC_FP32):
        check(self._lib.stereo_trunk_dev(self._h, int(net), _ptr(x), n, h, w, _ptr(out), int(precision)))

    def cost_volume_dev(self, lmap, rmap, n, h, w, c, disp_range, left_vol, right_vol, precision=PREC_FP32):
        check(self._lib.stereo_cost_volume_dev(self._h, _ptr(lmap), _ptr(rmap), n, h, w, c, int(disp_range),
                                               _ptr(left_vol), _ptr(right_vol), int(precision)))

    def wta_dev(self, vol, n, h, w, disp_range, take_min, disp):
        check(self._lib.stereo_wta_dev(self._h, _ptr(vol), n, h, w, int(disp_range), int(bool(take_min)), _ptr(disp)))

    def corr_wta_dev(self, lmap, rmap, n, h, w, c, disp_range, left_disp, right_disp, precision=PREC_FP32):
        check(self._lib.stereo_corr_wta_dev(self._h, _ptr(lmap), _ptr(rmap), n, h, w, c, int(disp_range),
                                            _ptr(left_disp), _ptr(right_disp), int(precision)))

    def lr_check_dev(self, d0, d1, n, h, w, disp_range, outlier):
        check(self._lib.stereo_lr_check_dev(self._h, _ptr(d0), _ptr(d1), n, h, w, int(disp_range), _ptr(outlier)))


class Trainer:
    """One `stereo_trainer`: patch training of a trunk on the context's GPU (reference train.py:92-112).

    `params` / `stats` are flat float32 arrays: per layer kernel [k][k][Cin][Cout], bias, gamma, beta
    (Keras' trainable_variables order), and per layer moving mean, moving variance."""

    def __init__(self, ctx: "Context", net: int, batch: int, disp_range: int):
        self._lib = load()
        self.ctx = ctx
        h = ctypes.c_void_p()
        check(self._lib.stereo_train_create(ctx._h, int(net), int(batch), int(disp_range), ctypes.byref(h)))
        self._h = h
        self.batch, self.disp_range = int(batch), int(disp_range)
        self.n_params = int(self._lib.stereo_train_param_count(h))
        self.n_stats = int(self._lib.stereo_train_stat_count(h))
        self.patch = int(self._lib.stereo_train_patch_size(h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.stereo_train_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_state(self, params, stats, adam_m=None, adam_v=None, step: int = 0):
        params, stats = _arr(params, np.float32), _arr(stats, np.float32)
        if params.size != self.n_params or stats.size != self.n_stats:
            raise ValueError(f"expected {self.n_params} parameters and {self.n_stats} statistics")
        m = None if adam_m is None else _arr(adam_m, np.float32)
        v = None if adam_v is None else _arr(adam_v, np.float32)
        check(self._lib.stereo_train_set_state(self._h, params.ctypes.data, stats.ctypes.data, _ptr(m), _ptr(v), int(step)))

    def get_state(self):
        out = {"params": np.empty(self.n_params, np.float32), "stats": np.empty(self.n_stats, np.float32),
               "adam_m": np.empty(self.n_params, np.float32), "adam_v": np.empty(self.n_params, np.float32)}
        step = ctypes.c_longlong(0)
        check(self._lib.stereo_train_get_state(self._h, out["params"].ctypes.data, out["stats"].ctypes.data,
                                               out["adam_m"].ctypes.data, out["adam_v"].ctypes.data, ctypes.byref(step)))
        out["step"] = int(step.value)
        return out

    def forward_backward(self, left, right, targets, want_grads: bool = False):
        left, right, targets = _arr(left, np.float32), _arr(right, np.float32), _arr(targets, np.float32)
        p, B, D = self.patch, self.batch, self.disp_range
        if left.shape != (B, p, p, 3) or right.shape != (B, p, p + D - 1, 3) or targets.shape != (B, D):
            raise ValueError(f"expected left {(B, p, p, 3)}, right {(B, p, p + D - 1, 3)}, targets {(B, D)}")
        loss, correct = ctypes.c_float(0), ctypes.c_int(0)
        grads = np.empty(self.n_params, np.float32) if want_grads else None
        check(self._lib.stereo_train_forward_backward(self._h, left.ctypes.data, right.ctypes.data, targets.ctypes.data,
                                                      ctypes.byref(loss), ctypes.byref(correct), _ptr(grads)))
        return float(loss.value), int(correct.value), grads

    def grads_dev(self) -> int:
        """Device address of the flat gradient array (for an in-place all-reduce)."""
        ptr = ctypes.c_void_p()
        check(self._lib.stereo_train_grads_dev(self._h, ctypes.byref(ptr)))
        return int(ptr.value)

    def apply(self, lr: float, grads_dev=None):
        check(self._lib.stereo_train_apply(self._h, _ptr(grads_dev), float(lr)))
