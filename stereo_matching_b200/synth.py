"""Synthetic KITTI-shaped stereo pairs with exact ground truth.

KITTI images, `val.npy` and the cost volumes are not shipped with the
reference (/root/reference/kitti2015/README.md is a wget stub), so every
parity test and benchmark in this repository runs on scenes made here.

A scene is a stack of fronto-parallel layers.  Layer k carries a multi-octave
blurred-noise RGB texture, a rectangular support in left-image coordinates and
an integer disparity d_k; layers are painted back to front with increasing
disparity, so occlusions are exact.  The left view shows a layer at x, the
right view shows it at x - d_k.  Ground truth follows KITTI `disp_noc`
(the file /root/reference/inference.py:104 reads): uint8 disparity of the left
pixel, 0 where the pixel is not visible in the right view.
"""
from __future__ import annotations

import numpy as np


def _texture(rng: np.random.Generator, h: int, w: int) -> np.ndarray:
    from scipy.ndimage import gaussian_filter

    acc = np.zeros((h, w, 3), np.float32)
    for sigma in (1.0, 2.0, 4.0, 8.0):
        noise = rng.random((h, w, 3), dtype=np.float32)
        blurred = gaussian_filter(noise, sigma=(sigma, sigma, 0), mode="wrap")
        blurred -= blurred.mean()
        blurred /= blurred.std() + 1e-12
        acc += blurred
    acc -= acc.min()
    acc *= 255.0 / max(float(acc.max()), 1e-12)
    return acc


def make_pair(seed: int, h: int = 375, w: int = 1242, disp_range: int = 228,
              n_layers: int = 8):
    """Return (left_u8[h,w,3], right_u8[h,w,3], gt_u8[h,w]) for `seed`."""
    rng = np.random.default_rng(seed)
    dmax = min(disp_range - 1, 200)
    dmax = max(dmax, 1)
    we = w + dmax                      # left-coordinate canvas, extended to the right
    disps = [min(5, dmax)] + sorted(int(v) for v in rng.integers(1, dmax + 1, size=n_layers))
    left_owner = np.zeros((h, we), np.int32)
    tex = []
    masks = []
    for k, d in enumerate(disps):
        tex.append(_texture(rng, h, we))
        m = np.zeros((h, we), bool)
        if k == 0:
            m[:] = True
        else:
            rh = int(rng.integers(max(h // 6, 2), max(h // 2, 3)))
            rw = int(rng.integers(max(w // 10, 2), max(w // 3, 3)))
            y0 = int(rng.integers(0, max(h - rh, 1)))
            x0 = int(rng.integers(0, max(we - rw, 1)))
            m[y0:y0 + rh, x0:x0 + rw] = True
        masks.append(m)
        left_owner[m] = k              # painter's order: later layers are nearer
    left = np.zeros((h, w, 3), np.float32)
    right = np.zeros((h, w, 3), np.float32)
    right_owner = np.zeros((h, w), np.int32)
    for k, d in enumerate(disps):
        sel = left_owner[:, :w] == k
        left[sel] = tex[k][:, :w][sel]
        # right pixel xr sees the layer point whose left coordinate is xr + d
        mk = masks[k][:, d:d + w]
        right[mk] = tex[k][:, d:d + w][mk]
        right_owner[mk] = k
    gt = np.zeros((h, w), np.uint8)
    xs = np.arange(w)[None, :].repeat(h, 0)
    ys = np.arange(h)[:, None].repeat(w, 1)
    for k, d in enumerate(disps):
        sel = left_owner[:, :w] == k
        xr = xs - d
        ok = sel & (xr >= 0)
        vis = np.zeros_like(ok)
        vis[ok] = right_owner[ys[ok], xr[ok]] == k
        gt[vis] = d
    return (np.clip(np.rint(left), 0, 255).astype(np.uint8),
            np.clip(np.rint(right), 0, 255).astype(np.uint8), gt)


def make_base_model_weights(seed: int = 7):
    """Seeded stand-in weights for `base_model` (37x37), for which the
    reference ships no checkpoint (/root/reference/models/model.py:5-28).

    Same list-of-dicts layout as `checkpoint.load_net_weights`.
    """
    rng = np.random.default_rng(seed)
    chans = [3, 32, 32, 32, 64, 64, 64, 64, 64, 64]
    layers = []
    for i in range(9):
        cin, cout = chans[i], chans[i + 1]
        fan_in = 25 * cin
        layers.append({
            "kernel": (rng.standard_normal((5, 5, cin, cout)) * np.sqrt(2.0 / fan_in)).astype(np.float32),
            "bias": (rng.standard_normal(cout) * 0.05).astype(np.float32)})
        layers.append({
            "gamma": rng.uniform(0.5, 1.5, cout).astype(np.float32),
            "beta": (rng.standard_normal(cout) * 0.1).astype(np.float32),
            "moving_mean": rng.standard_normal(cout).astype(np.float32),
            "moving_variance": rng.uniform(0.5, 2.0, cout).astype(np.float32)})
    return layers
