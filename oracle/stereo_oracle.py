"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's inference hot path.

Nothing under `oracle/` is part of the product: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import it, and
only as the checker or the timed CPU baseline.  The shipped library never
routes through this code.

The reference (TensorFlow 2 / Keras) cannot be imported in this environment, so
its algorithm is restated op for op with numpy + torch-CPU.  Every function
cites the reference lines it follows.

Parity status:
  * weights        - pinned by the shipped checkpoint (tests/test_checkpoint.py, K1)
  * 3-px metric    - pinned by the shipped PNGs and README numbers (K2, K3)
  * trunk features, cost volumes, WTA maps - PARITY UNPINNED: the reference
    ships no cost volumes and no input images, so these functions are a
    restatement of model.py / inference.py checked only for self-consistency
    (K7: WTA never selects an invalid slot on the shipped maps).
"""
from __future__ import annotations

import numpy as np

BN_EPS = 1e-3   # Keras BatchNormalization default epsilon (models/model.py:8 uses defaults)


def normalize_and_pad(img_u8: np.ndarray, patch_size: int) -> np.ndarray:
    """/root/reference/inference.py:109-133.

    `(x - x.mean()) / x.std()` over all pixels and channels in float64
    (numpy population std), then zero-padded into a float32
    [1, H+p-1, W+p-1, 3] canvas at offset (p-1)//2.
    """
    x = (img_u8 - img_u8.mean()) / img_u8.std()
    h, w = x.shape[:2]
    out = np.zeros([1, h + patch_size - 1, w + patch_size - 1, 3], dtype=np.float32)
    r0 = int((patch_size - 1) // 2)
    c0 = int((patch_size - 1) // 2)
    out[0, r0:r0 + h, c0:c0 + w] = x.reshape(h, w, 3)
    return out


def _net_spec(net: str):
    """(kernel, [relu flags]) per /root/reference/models/model.py.

    base_model_ws_9 (:30-41): 4x Conv2D 3x3 valid, relu on 1-3, none on 4.
    base_model (:5-28): 9x Conv2D 5x5 valid, relu on 1-8, none on 9; the first
    conv8/bn8 pair (:21-22) is overwritten at :23-24 and is not in the graph.
    """
    if net == "win9_dep9":
        return 3, [True, True, True, False]
    if net == "win37_dep9":
        return 5, [True] * 8 + [False]
    raise ValueError(f"unknown net_type {net!r}")


def trunk_forward(x_nhwc: np.ndarray, layers, net: str = "win9_dep9",
                  dtype=np.float64, threads: int | None = None) -> np.ndarray:
    """`model(x, training=False)` of /root/reference/inference.py:136-137.

    Each stage is Conv2D(valid, bias, [relu]) followed by BatchNormalization in
    inference mode: gamma * (x - moving_mean) / sqrt(moving_variance + 1e-3) + beta
    (models/model.py:32-39; conv -> activation -> BN order).
    `layers` is the list produced by `checkpoint.load_net_weights`.
    dtype=float64 is the ground truth; float32 is the timed CPU baseline.
    """
    import torch
    import torch.nn.functional as F

    if threads:
        torch.set_num_threads(threads)
    _, relus = _net_spec(net)
    if len(layers) != 2 * len(relus):
        raise ValueError(f"{net}: expected {2 * len(relus)} weight groups, got {len(layers)}")
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    with torch.no_grad():
        a = torch.from_numpy(np.ascontiguousarray(x_nhwc)).to(tdt).permute(0, 3, 1, 2)
        for i, relu in enumerate(relus):
            conv, bn = layers[2 * i], layers[2 * i + 1]
            wgt = torch.from_numpy(conv["kernel"]).to(tdt).permute(3, 2, 0, 1).contiguous()
            a = F.conv2d(a, wgt, torch.from_numpy(conv["bias"]).to(tdt))
            if relu:
                a = torch.relu(a)
            g = torch.from_numpy(bn["gamma"]).to(tdt)[None, :, None, None]
            b = torch.from_numpy(bn["beta"]).to(tdt)[None, :, None, None]
            m = torch.from_numpy(bn["moving_mean"]).to(tdt)[None, :, None, None]
            v = torch.from_numpy(bn["moving_variance"]).to(tdt)[None, :, None, None]
            a = (a - m) * (g / torch.sqrt(v + BN_EPS)) + b
        return a.permute(0, 2, 3, 1).contiguous().numpy()


def cost_volumes(lmap: np.ndarray, rmap: np.ndarray, disp_range: int,
                 accumulate=np.float64):
    """/root/reference/inference.py:139-150, same loop structure.

    lmap/rmap: [1,H,W,C].  Returns (left, right) volumes [H,W,D] float64 whose
    unwritten slots stay exactly 0.0 (np.zeros at :140-141).
      left [y, x, d] = <L[y,x], R[y,x-d]>   for x >= d
      right[y, x, d] = <L[y,x+d], R[y,x]>   for x <  W-d
    accumulate=float32 reproduces the reference's fp32 reduce_sum; float64 is
    the ground truth the GPU kernels are gated against.
    """
    map_width = lmap.shape[2]
    left = np.zeros((lmap.shape[1], lmap.shape[2], disp_range))
    right = np.zeros((rmap.shape[1], rmap.shape[2], disp_range))
    lm = lmap.astype(accumulate, copy=False)
    rm = rmap.astype(accumulate, copy=False)
    for loc in range(disp_range):
        x_off = -loc
        if loc >= map_width:
            break  # empty slices in the reference: nothing is written
        l = lm[:, :, max(0, -x_off):map_width, :]
        r = rm[:, :, 0:min(map_width, map_width + x_off), :]
        inner = np.sum(l * r, axis=3, dtype=accumulate)
        left[:, max(0, -x_off):map_width, loc] = inner[0]
        right[:, 0:min(map_width, map_width + x_off), loc] = inner[0]
    return left, right


def cost_volumes_banded(lmap: np.ndarray, rmap: np.ndarray, disp_range: int, dtype=np.float32):
    """The same two volumes the "fair" way for a CPU: per image row one W x W product
    S = L[y] @ R[y].T through BLAS, of which only the band 0 <= xl - xr < D is kept,
      left [y, x, d] = S[x, x - d],   right[y, x, d] = S[x + d, x].
    Not the reference's loop structure (inference.py:139-150 runs D element-wise multiply-reduce passes): it is
    timed beside `cost_volumes` in bench.py and labelled as such.  Unwritten slots stay 0.0."""
    h, w = lmap.shape[1], lmap.shape[2]
    left = np.zeros((h, w, disp_range), dtype)
    right = np.zeros((h, w, disp_range), dtype)
    d = np.arange(min(disp_range, w))
    x = np.arange(w)
    xl, dd = np.meshgrid(x, d, indexing="ij")              # left:  S[x, x - d] where x >= d
    ok_l = xl >= dd
    ok_r = xl + dd < w                                     # right: S[x + d, x] where x + d < W
    lm = lmap[0].astype(dtype, copy=False)
    rm = rmap[0].astype(dtype, copy=False)
    for y in range(h):
        S = lm[y] @ rm[y].T
        left[y][:, :len(d)][ok_l] = S[xl[ok_l], (xl - dd)[ok_l]]
        right[y][:, :len(d)][ok_r] = S[(xl + dd)[ok_r], xl[ok_r]]
    return left, right


def wta(volume: np.ndarray) -> np.ndarray:
    """/root/reference/inference.py:169-183: argmax over all D slots (zero-filled
    ones included), first maximum wins, cast to uint8."""
    return np.argmax(volume, axis=2).astype(np.uint8)


def three_pixel_error(pred: np.ndarray, gt: np.ndarray) -> float:
    """/root/reference/inference.py:185-196 and cal.py:17-28."""
    valid = (gt != 0).astype("float")
    masked = pred * valid
    n_valid = valid.sum()
    n_err = (np.abs(masked - gt) > 3).sum()
    return float(n_err / n_valid)


def bin_name(side: str, file_id: int, h: int, w: int, d: int) -> str:
    """File name of inference.py:217 / :228, parsed back by post.lua:81-87."""
    return "%s_%06d_10_%d_%d_%d.bin" % (side, file_id, h, w, d)


def volume_bytes(volume: np.ndarray) -> bytes:
    """inference.py:210-220: astype(float32), C order [H][W][D], raw bytes."""
    return bytearray(volume.astype(np.float32))


def infer_pair(left_u8, right_u8, layers, net="win9_dep9", patch_size=9,
               disp_range=201, dtype=np.float64, banded=False):
    """One iteration of the per-pair loop, inference.py:99-183 (`banded`: the cost volumes through
    `cost_volumes_banded` instead of the reference's D-loop)."""
    l = normalize_and_pad(left_u8, patch_size)
    r = normalize_and_pad(right_u8, patch_size)
    lmap = trunk_forward(l, layers, net, dtype)
    rmap = trunk_forward(r, layers, net, dtype)
    lv, rv = (cost_volumes_banded(lmap, rmap, disp_range, dtype) if banded
              else cost_volumes(lmap, rmap, disp_range, accumulate=dtype))
    return {"lmap": lmap, "rmap": rmap, "left_cost": lv, "right_cost": rv,
            "left_disp": wta(lv), "right_disp": wta(rv)}
