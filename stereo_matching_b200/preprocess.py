"""Patch-location files for training, without MATLAB/octave: restates
/root/reference/preprocess/kitti2015_gene_loc_1.m (+ disp_read.m, preprocess.m).

    python -m stereo_matching_b200.preprocess --data_root ./kitti2015/training --save_dir preprocess/debug_15_ws_9 \
        --tr_num 160 --val_num 40 --psz 4 --half_range 100 --seed 123

writes `tr_<tr_num>_<psz>_<half_range>.bin`, `val_<val_num>_<psz>_<half_range>.bin` (float32, five values per
location: file id, type (1 = horizontal), left x, left y, right x; 1-based as MATLAB writes them, in MATLAB's
column-major `find` order) and `myPerm.bin` (float32 file ids, training images first) - the three files
data_handler.py:31-40 reads.

The permutation: the .m file draws `randperm(200)-1` with its `rng(seed)` line commented out, so the reference's
own split is not reproducible; here it is `numpy.random.default_rng(seed).permutation(n_files)`, or the ids of
an existing `myPerm.bin` passed with `--perm` (to regenerate location files for the reference's split).
"""
from __future__ import annotations

import argparse
import os

import numpy as np


def disp_read(path: str) -> np.ndarray:
    """disp_read.m: 16-bit PNG / 256, 0 -> -1 (invalid).  float64 [h, w]."""
    import cv2
    raw = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if raw is None:
        raise FileNotFoundError(path)
    d = raw.astype(np.float64) / 256.0
    d[raw == 0] = -1.0
    return d


def _matlab_round(x: np.ndarray) -> np.ndarray:
    return np.sign(x) * np.floor(np.abs(x) + 0.5)       # MATLAB rounds halves away from zero


def locations(disp: np.ndarray, file_id: int, psz: int, half_range: int) -> np.ndarray:
    """kitti2015_gene_loc_1.m:52-68 for one image -> float32 [n][5], 1-based coordinates."""
    h, w = disp.shape
    c0, r0 = np.nonzero(disp.T != -1)                    # column-major order of MATLAB's find
    x, y = c0 + 1.0, r0 + 1.0
    xr = _matlab_round(x - disp[r0, c0])
    ll = (x + psz <= w) & (x - psz > 0) & (y + psz <= h) & (y - psz > 0)
    rr1 = (xr - half_range - psz > 0) & (xr + half_range + psz <= w) & (y - psz > 0) & (y + psz <= h)
    ok = ll & rr1                                        # the vertical type (rr_type2) is commented out upstream
    n = int(ok.sum())
    out = np.empty((n, 5), np.float32)
    out[:, 0], out[:, 1], out[:, 2], out[:, 3], out[:, 4] = file_id, 1, x[ok], y[ok], xr[ok]
    return out


def gene_loc(tr_num: int, val_num: int, psz: int, half_range: int, save_dir: str, seed: int, data_root: str,
             perm=None, n_files: int = 200, noc_occ: str = "noc"):
    """Returns (number of training locations, number of validation locations)."""
    os.makedirs(save_dir, exist_ok=True)
    if perm is None:
        perm = np.random.default_rng(seed).permutation(n_files)
    perm = np.asarray(perm).astype(np.int64)
    counts = []
    for name, ids in (("tr", perm[:tr_num]), ("val", perm[tr_num:tr_num + val_num])):
        parts = [locations(disp_read("%s/disp_%s_0/%06d_10.png" % (data_root, noc_occ, i)), int(i), psz, half_range)
                 for i in ids]
        allloc = np.concatenate(parts) if parts else np.empty((0, 5), np.float32)
        allloc.astype("<f4").tofile("%s/%s_%d_%d_%d.bin" % (save_dir, name, len(ids), psz, half_range))
        counts.append(len(allloc))
    perm.astype("<f4").tofile(os.path.join(save_dir, "myPerm.bin"))
    return tuple(counts)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--data_root", default="./kitti2015/training")
    ap.add_argument("--save_dir", default="preprocess/debug_15_ws_9")
    ap.add_argument("--tr_num", type=int, default=160)
    ap.add_argument("--val_num", type=int, default=40)
    ap.add_argument("--psz", type=int, default=4, help="half patch: 4 for win9_dep9, 18 for win37_dep9")
    ap.add_argument("--half_range", type=int, default=100)
    ap.add_argument("--seed", type=int, default=123)
    ap.add_argument("--n_files", type=int, default=200)
    ap.add_argument("--perm", default=None, help="existing myPerm.bin to take the file order from")
    a = ap.parse_args(argv)
    perm = np.fromfile(a.perm, "<f4") if a.perm else None
    tr, val = gene_loc(a.tr_num, a.val_num, a.psz, a.half_range, a.save_dir, a.seed, a.data_root, perm, a.n_files)
    print("training locations: %d, validation locations: %d -> %s" % (tr, val, a.save_dir))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
