"""CPU restatement of the reference's training step.  TEST INFRASTRUCTURE ONLY: imported by tests/ and by
bench/smoke checks, never by the product path.

Follows /root/reference/train.py:92-112 (two tower passes in training mode, `map_inner_product`
train.py:11-17, `softmax_cross_entropy_with_logits` + `reduce_mean` train.py:106, Adam train.py:80,112),
/root/reference/models/model.py:30-41 (conv 'valid' -> ReLU -> BatchNormalization, Keras defaults
momentum 0.99, epsilon 1e-3) with torch autograd in float64 (or float32).

Parity status: **unpinned** - TensorFlow is not installable here and the reference ships no training
fixtures (loss curves, gradients); the CUDA path is checked against this restatement only.  Keras'
BatchNormalization normalises with the biased batch variance and (fused path) also feeds the biased
variance into the moving average; both are restated that way.

Parameter layout (flat, Keras `trainable_variables` order): per layer kernel [k][k][Cin][Cout], bias,
gamma, beta.  Statistics: per layer moving mean, moving variance.
"""
from typing import Dict, List, Tuple

import numpy as np
import torch

BN_EPS = 1e-3
BN_MOMENTUM = 0.99

SPECS = {  # name -> (kernel size, channels)
    "win9_dep9": (3, [3, 64, 64, 64, 64]),
    "win37_dep9": (5, [3, 32, 32, 32, 64, 64, 64, 64, 64, 64]),
}


def param_layout(net: str) -> Tuple[List[Dict[str, slice]], int, int]:
    ks, ch = SPECS[net]
    layers, po, so = [], 0, 0
    for l in range(len(ch) - 1):
        cin, cout = ch[l], ch[l + 1]
        d = {"kernel": slice(po, po + ks * ks * cin * cout)}
        po += ks * ks * cin * cout
        for name in ("bias", "gamma", "beta"):
            d[name] = slice(po, po + cout)
            po += cout
        d["mean"] = slice(so, so + cout)
        d["var"] = slice(so + cout, so + 2 * cout)
        so += 2 * cout
        d["shape"] = (ks, ks, cin, cout)
        layers.append(d)
    return layers, po, so


def init_params(net: str, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Glorot-uniform kernels, zero biases, gamma 1, beta 0, moving mean 0 / variance 1 (Keras defaults)."""
    layers, n_p, n_s = param_layout(net)
    rng = np.random.default_rng(seed)
    p, s = np.zeros(n_p, np.float32), np.zeros(n_s, np.float32)
    for d in layers:
        kh, kw, cin, cout = d["shape"]
        lim = np.sqrt(6.0 / (kh * kw * cin + kh * kw * cout))
        p[d["kernel"]] = rng.uniform(-lim, lim, kh * kw * cin * cout).astype(np.float32)
        p[d["gamma"]] = 1.0
        s[d["var"]] = 1.0
    return p, s


def _tower(x: torch.Tensor, params: torch.Tensor, stats: np.ndarray, net: str, update_stats: bool):
    """x: [B, h, w, 3] NHWC.  Returns [B, h', w', C]; updates `stats` in place like Keras does."""
    layers, _, _ = param_layout(net)
    y = x.permute(0, 3, 1, 2)                                   # NCHW for torch
    for li, d in enumerate(layers):
        kh, kw, cin, cout = d["shape"]
        w = params[d["kernel"]].reshape(kh, kw, cin, cout).permute(3, 2, 0, 1)
        y = torch.nn.functional.conv2d(y, w, params[d["bias"]])
        if li < len(layers) - 1:
            y = torch.relu(y)
        mean = y.mean(dim=(0, 2, 3))
        var = y.var(dim=(0, 2, 3), unbiased=False)
        if update_stats:
            stats[d["mean"]] = stats[d["mean"]] * BN_MOMENTUM + mean.detach().numpy().astype(np.float32) * np.float32(1 - BN_MOMENTUM)
            stats[d["var"]] = stats[d["var"]] * BN_MOMENTUM + var.detach().numpy().astype(np.float32) * np.float32(1 - BN_MOMENTUM)
        xhat = (y - mean[None, :, None, None]) / torch.sqrt(var[None, :, None, None] + BN_EPS)
        y = params[d["gamma"]][None, :, None, None] * xhat + params[d["beta"]][None, :, None, None]
    return y.permute(0, 2, 3, 1)


def forward_backward(params: np.ndarray, stats: np.ndarray, left: np.ndarray, right: np.ndarray, targets: np.ndarray,
                     net: str = "win9_dep9", dtype=torch.float64):
    """One forward/backward of train.py:98-110.  Returns loss, correct3, grads (flat float64), logits; `stats`
    (float32, flat) is updated in place by the left then the right tower pass."""
    p = torch.tensor(np.asarray(params), dtype=dtype, requires_grad=True)
    lf = _tower(torch.tensor(np.asarray(left), dtype=dtype), p, stats, net, True)          # [B,1,1,C]
    rf = _tower(torch.tensor(np.asarray(right), dtype=dtype), p, stats, net, True)         # [B,1,D,C]
    logits = torch.einsum("bc,bdc->bd", lf[:, 0, 0, :], rf[:, 0, :, :])                    # train.py:11-17
    t = torch.tensor(np.asarray(targets), dtype=dtype)
    loss = (-(t * torch.log_softmax(logits, dim=1)).sum(dim=1)).mean()                     # train.py:106
    loss.backward()
    pred = logits.argmax(dim=1).numpy()
    label = np.asarray(targets).argmax(axis=1)                                             # train.py:96
    correct3 = int((np.abs(pred - label) <= 3).sum())                                      # train.py:117-119
    return float(loss.item()), correct3, p.grad.numpy().astype(np.float64), logits.detach().numpy()


def adam_step(params: np.ndarray, grads: np.ndarray, m: np.ndarray, v: np.ndarray, step: int, lr: float,
              b1: float = 0.9, b2: float = 0.999, eps: float = 1e-7):
    """tf.keras.optimizers.Adam (train.py:80,112), `step` = 1 for the first update.  Float64 arithmetic."""
    g = grads.astype(np.float64)
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    lr_t = lr * np.sqrt(1 - b2 ** step) / (1 - b1 ** step)
    return params - lr_t * m / (np.sqrt(v) + eps), m, v


def learning_rate(step: int, base: float = 1e-3) -> float:
    """train.py:125-130: divided by 5 at step 24000 and every 8000 steps after that."""
    if step < 24000:
        return base
    return base / 5.0 ** (1 + (step - 24000) // 8000)


def soft_labels(batch: int, disp_range: int) -> np.ndarray:
    """data_handler.py:78-85: 5-tap soft label centred on disp_range // 2."""
    t = np.zeros((batch, disp_range), np.float32)
    for k, wgt in enumerate([0.05, 0.2, 0.5, 0.2, 0.05]):
        t[:, disp_range // 2 - 2 + k] = wgt
    return t
