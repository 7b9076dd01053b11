"""Drop-in for /root/reference/models/model.py: the same two factories,
`base_model(input_shape)` (:5-28) and `base_model_ws_9(input_shape)` (:30-41),
returning an object that is called like the Keras model at
/root/reference/inference.py:136-137 - `y = model(x, training=False)` with `x`
float32 NHWC `[N,h,w,3]` - but runs on libstereo_b200's sm_100a kernels.

Weights come from the reference's own TensorBundle checkpoints via
`model.load_checkpoint(model_dir)` (the `ckpt.restore(manager.latest_checkpoint)`
of inference.py:88-91) or from `model.set_layers(layers)`.

BatchNormalization folding.  Every block of the reference is
conv('valid') -> [ReLU] -> BN(inference).  BN is the affine map s*a + t with
s = gamma / sqrt(moving_variance + 1e-3), t = beta - s * moving_mean.  Because
no padding happens inside the network, the affine map of block i folds exactly
(in real arithmetic) into the NEXT convolution:
    conv_{i+1}(s*a + t) = conv(W * s[cin], a) + (b + sum_{taps,cin} W * t[cin]),
and the last BN, which follows a linear conv, folds into that conv's output
channels.  The fold is done here in float64 and rounded once to float32.
"""
from __future__ import annotations

import os
from typing import Dict, List

import numpy as np

from .. import _lib, checkpoint

BN_EPS = 1e-3   # Keras BatchNormalization default (models/model.py uses the defaults)

_SPECS = {
    # name: (net id, kernel size, channels, total valid-conv shrink)
    "base_model_ws_9": (_lib.NET_WIN9, 3, [3, 64, 64, 64, 64], 8),
    "base_model": (_lib.NET_WIN37, 5, [3, 32, 32, 32, 64, 64, 64, 64, 64, 64], 36),
}

_default_context: Dict[int, "_lib.Context"] = {}


def default_context(device: int = 0) -> "_lib.Context":
    ctx = _default_context.get(device)
    if ctx is None:
        ctx = _default_context[device] = _lib.Context(device)
    return ctx


def fold_batchnorm(layers: List[Dict[str, np.ndarray]], ks: int, chans: List[int]) -> np.ndarray:
    """Fold the BN layers into the convolutions and pack for stereo_set_weights:
    per conv layer, kernel [k][k][Cin][Cout] then bias [Cout]."""
    n_conv = len(chans) - 1
    if len(layers) != 2 * n_conv:
        raise ValueError(f"expected {2 * n_conv} weight groups (conv, bn alternating), got {len(layers)}")
    packed = []
    s_prev = t_prev = None
    for i in range(n_conv):
        conv, bn = layers[2 * i], layers[2 * i + 1]
        k = np.asarray(conv["kernel"], np.float64)
        b = np.asarray(conv["bias"], np.float64)
        if k.shape != (ks, ks, chans[i], chans[i + 1]):
            raise ValueError(f"conv {i}: kernel shape {k.shape}, expected {(ks, ks, chans[i], chans[i + 1])}")
        s = np.asarray(bn["gamma"], np.float64) / np.sqrt(np.asarray(bn["moving_variance"], np.float64) + BN_EPS)
        t = np.asarray(bn["beta"], np.float64) - s * np.asarray(bn["moving_mean"], np.float64)
        if s_prev is not None:       # BN of the previous block acts on this conv's input
            b = b + np.einsum("hwio,i->o", k, t_prev)
            k = k * s_prev[None, None, :, None]
        if i == n_conv - 1:          # no ReLU after the last conv: its own BN folds in too
            k = k * s[None, None, None, :]
            b = b * s + t
        packed.append(k.astype(np.float32).ravel())
        packed.append(b.astype(np.float32).ravel())
        s_prev, t_prev = s, t
    return np.concatenate(packed)


class StereoTrunk:
    """Callable stand-in for the Keras `Model` the reference factories return."""

    def __init__(self, name: str, input_shape=(None, None, 3), device: int = 0, precision: int = _lib.PREC_BF16X3):
        # default: the tcgen05 kernels with split-bf16 operands (meets the 1e-4 cost-volume gate);
        # precision=_lib.PREC_FP32 selects the CUDA-core fp32 kernels
        if tuple(input_shape)[-1] != 3:
            raise ValueError("the reference trunks take 3-channel input (models/model.py)")
        self.name = name
        self.net, self.kernel_size, self.channels, self.shrink = _SPECS[name]
        self.input_shape = tuple(input_shape)
        self.precision = precision
        self.device = device
        self.layers = None
        self._ctx = None

    # -- weights --
    @property
    def context(self) -> "_lib.Context":
        if self._ctx is None:
            self._ctx = default_context(self.device)
        return self._ctx

    def set_layers(self, layers):
        """`layers`: [conv, bn, conv, bn, ...] dicts as in checkpoint.load_net_weights."""
        packed = fold_batchnorm(layers, self.kernel_size, self.channels)
        self.context.set_weights(self.net, packed)
        self.layers = layers
        return self

    def load_checkpoint(self, model_dir: str, verbose: bool = True) -> str:
        """Restore from `<model_dir>/checkpoint` -> latest TensorBundle
        (inference.py:86-95).  Returns the restored prefix."""
        prefix = checkpoint.latest_checkpoint(model_dir) if os.path.isdir(model_dir) else model_dir
        if prefix is None:
            raise FileNotFoundError(f"no checkpoint found in {model_dir}")
        self.set_layers(checkpoint.load_net_weights(prefix))
        if verbose:
            print("Restored from {}".format(prefix))     # inference.py:93
        return prefix

    # -- forward --
    def __call__(self, x, training: bool = False):
        if training:
            raise NotImplementedError("training=True (batch-statistics BatchNorm, train.py:100-101) runs through "
                                      "stereo_matching_b200.train / _lib.Trainer, which owns the device state")
        if self.layers is None:
            raise RuntimeError("weights not set: call load_checkpoint(model_dir) or set_layers(layers) first")
        x = np.asarray(x)
        if x.ndim != 4 or x.shape[3] != 3:
            raise ValueError(f"expected NHWC input [N,h,w,3], got {x.shape}")
        if x.shape[1] <= self.shrink or x.shape[2] <= self.shrink:
            raise ValueError(f"input {x.shape[1]}x{x.shape[2]} is smaller than the receptive field {self.shrink + 1}")
        return self.context.trunk(self.net, x.astype(np.float32, copy=False), self.precision)

    predict = __call__


def base_model(input_shape=(None, None, 3), **kw) -> StereoTrunk:
    """/root/reference/models/model.py:5-28 - nine 5x5 valid convs, 37x37 receptive field."""
    return StereoTrunk("base_model", input_shape, **kw)


def base_model_ws_9(input_shape=(None, None, 3), **kw) -> StereoTrunk:
    """/root/reference/models/model.py:30-41 - four 3x3 valid convs, 9x9 receptive field."""
    return StereoTrunk("base_model_ws_9", input_shape, **kw)
