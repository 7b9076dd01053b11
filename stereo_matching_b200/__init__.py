"""stereo_matching_b200 - the Luo/Schwing/Urtasun stereo hot path of
datvuthanh/Stereo-Matching on hand-written sm_100a CUDA kernels.

Python is the host language of the reference, so the host side here is Python
too: a ctypes binding of the C-ABI library `libstereo_b200.so`
(include/stereo_b200.h) plus mirrors of the reference's own surface
(`models.model.base_model_ws_9`, `inference`, `post`).  The package directory is
spelled with an underscore because `stereo-matching_b200` is not importable.
"""
from . import checkpoint, metric, synth  # noqa: F401
from ._lib import Context, StereoError, NET_WIN9, NET_WIN37, PREC_FP32, PREC_BF16X3, PREC_BF16  # noqa: F401

__all__ = ["Context", "StereoError", "checkpoint", "metric", "synth",
           "NET_WIN9", "NET_WIN37", "PREC_FP32", "PREC_BF16X3", "PREC_BF16"]
