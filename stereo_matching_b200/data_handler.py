"""Drop-in for /root/reference/data_handler.py: the KITTI patch sampler `train.py` feeds the trainer with.

Same class name, constructor arguments, attributes (`tr_loc`, `val_loc`, `tr_perm`, `val_perm`, `tr_ptr`,
`curr_epoch`, `img_id_arr`) and methods (`next_batch`, `evaluate`) as the reference.  What it reads
(data_handler.py:31-40,61-69):

    <util_root>/myPerm.bin                        float32: KITTI file ids, training images first
    <util_root>/tr_<n>_<half_patch>_<half_range>.bin   float32 [N][5]: id, type, x, y, right x (1-based)
    <util_root>/val_<n>_<half_patch>_<half_range>.bin
    <data_root>/image_2|image_3/%06d_10.png       (kitti2015, RGB) or image_0|image_1 (kitti2012, grey)

Differences from the reference, all deliberate:
  * images are read with cv2 (scipy.misc.imread is gone from scipy); channel order is converted to RGB;
  * patches are float32 (the reference fills float64 arrays that Keras casts to float32 on entry);
  * the batch is gathered with one fancy-index per image side instead of a Python loop per patch;
  * the reference's epoch wrap-around raises TypeError (`str + int` in its print, data_handler.py:131); here
    the wrap just happens: the sample after the last one is sample 0, `curr_epoch` goes up by one;
  * `val.npy` is written to `val_npy` (default 'val.npy' in the working directory, as data_handler.py:114).
Shuffling uses `np.random.shuffle` on the global generator like the reference, so `np.random.seed(123)` before
construction (train.py:43) gives the same permutations.
"""
from __future__ import annotations

import os
import sys

import numpy as np

_DIST = (0.05, 0.2, 0.5, 0.2, 0.05)        # data_handler.py:81


def read_image(path: str, num_channels: int) -> np.ndarray:
    """uint8 [h, w, num_channels]; RGB order for 3 channels (what scipy.misc.imread returned)."""
    import cv2
    img = cv2.imread(path, cv2.IMREAD_COLOR if num_channels == 3 else cv2.IMREAD_GRAYSCALE)
    if img is None:
        raise FileNotFoundError(path)
    if num_channels == 3:
        return img[:, :, ::-1].copy()
    return img[:, :, None]


def soft_label_rows(rows: int, disp_range: int) -> np.ndarray:
    """data_handler.py:78-85: the 5-tap target centred on disp_range // 2, one row per sample."""
    t = np.zeros((rows, disp_range), np.float32)
    half = len(_DIST) // 2
    for k, wgt in enumerate(_DIST):
        t[:, disp_range // 2 - half + k] = wgt
    return t


class Data_handler:
    def __init__(self, data_version, data_root, util_root, num_tr_img, num_val_img, num_val_loc, batch_size,
                 patch_size, disp_range, val_npy: str = "val.npy"):
        if data_version == "kitti2015":
            self.num_channels, dirs = 3, ("image_2", "image_3")
        elif data_version == "kitti2012":
            self.num_channels, dirs = 1, ("image_0", "image_1")
        else:
            sys.exit("data_version should be either kitti2012 or kitti2015")
        self.batch_size, self.patch_size, self.disp_range = batch_size, patch_size, disp_range
        self.half_patch, self.half_range = patch_size // 2, disp_range // 2
        self.tr_ptr = 0
        self.curr_epoch = 0

        self.file_ids = np.fromfile(os.path.join(util_root, "myPerm.bin"), "<f4")
        name = "%s/%s_%d_%d_%d.bin"
        self.tr_loc = np.fromfile(name % (util_root, "tr", num_tr_img, self.half_patch, self.half_range),
                                  "<f4").reshape(-1, 5).astype(int)
        if num_val_img == 0:
            self.val_loc = self.tr_loc                      # same array, as in the reference (:36)
            print("validate on training set..")
        else:
            self.val_loc = np.fromfile(name % (util_root, "val", num_val_img, self.half_patch, self.half_range),
                                       "<f4").reshape(-1, 5).astype(int)
        print("#training locations: %d -- #valuation locations: %d" % (self.tr_loc.shape[0], self.val_loc.shape[0]))
        # 1-based (MATLAB) -> 0-based; when val_loc IS tr_loc the reference subtracts twice (:45-47), kept
        for i in range(2, 5):
            self.tr_loc[:, i] -= 1
            self.val_loc[:, i] -= 1

        self.tr_perm = np.arange(self.tr_loc.shape[0])
        self.val_perm = np.arange(self.val_loc.shape[0])
        np.random.shuffle(self.tr_perm)
        np.random.shuffle(self.val_perm)

        self.ldata, self.rdata = {}, {}
        for i in range(num_tr_img + num_val_img):
            fn = int(self.file_ids[i])
            for store, d in ((self.ldata, dirs[0]), (self.rdata, dirs[1])):
                img = read_image("%s/%s/%06d_10.png" % (data_root, d, fn), self.num_channels).astype(np.float64)
                store[fn] = ((img - img.mean()) / img.std()).astype(np.float32)      # :68-69

        self.batch_label = soft_label_rows(batch_size, disp_range)
        self.val_label = soft_label_rows(num_val_loc, disp_range)
        self.val_left, self.val_right = self._gather(self.val_loc[self.val_perm[:num_val_loc]], val=True)
        self.img_id_arr = np.unique(self.val_loc[self.val_perm[:num_val_loc], 0])   # :92,110
        if val_npy:
            with open(val_npy, "wb") as f:
                np.save(f, self.img_id_arr)
        print("validation created: num(%d)" % num_val_loc)

    def _gather(self, locs: np.ndarray, val: bool = False):
        """locs [n][5] (id, type, x, y, right x; 0-based) -> left [n,p,p,c], right [n,p,p+D-1,c].  `val`: the
        validation set keeps the left patch of a vertical location untransposed (data_handler.py:98 vs :140)."""
        p, hp, hr, c = self.patch_size, self.half_patch, self.half_range, self.num_channels
        wr = p + self.disp_range - 1
        left = np.empty((len(locs), p, p, c), np.float32)
        right = np.empty((len(locs), p, wr, c), np.float32)
        dy = np.arange(-hp, hp + 1)
        for img_id in np.unique(locs[:, 0]):
            sel = np.nonzero(locs[:, 0] == img_id)[0]
            limg, rimg = self.ldata[int(img_id)], self.rdata[int(img_id)]
            for kind in (1, 2):
                s = sel[locs[sel, 1] == kind]
                if not len(s):
                    continue
                x, y, xr = locs[s, 2], locs[s, 3], locs[s, 4]
                span_x, span_y = (hp + hr, hp) if kind == 1 else (hp, hp + hr)
                h_img, w_img = limg.shape[:2]
                if (min(x.min(), y.min()) < hp or x.max() + hp >= w_img or y.max() + span_y >= h_img or
                        y.min() < span_y or xr.min() < span_x or xr.max() + span_x >= w_img):
                    # the reference fails here too (a short slice does not fit the batch array)
                    raise ValueError("patch location outside image %d" % int(img_id))
                lp = limg[(y[:, None] + dy)[:, :, None], (x[:, None] + dy)[:, None, :]]
                if kind == 1:        # horizontal: right strip along x (:98-99,137-139)
                    dx = np.arange(-hp - hr, hp + hr + 1)
                    left[s] = lp
                    right[s] = rimg[(y[:, None] + dy)[:, :, None], (xr[:, None] + dx)[:, None, :]]
                else:                # vertical: strip along y, both transposed (:100-101,140-142)
                    dv = np.arange(-hp - hr, hp + hr + 1)
                    left[s] = lp if val else lp.transpose(0, 2, 1, 3)
                    right[s] = rimg[(y[:, None] + dv)[:, :, None], (xr[:, None] + dy)[:, None, :]].transpose(0, 2, 1, 3)
        return left, right

    def next_batch(self):
        # data_handler.py:121-133: samples tr_ptr, tr_ptr + 1, ...; past the last one it starts again at sample 0
        n, start = self.tr_perm.shape[0], self.tr_ptr
        end = start + self.batch_size
        idx = (start + np.arange(self.batch_size)) % n
        wraps = (end - 1) // n - (max(start, 1) - 1) // n
        self.curr_epoch += wraps
        self.tr_ptr = end - n * wraps
        self.batch_left, self.batch_right = self._gather(self.tr_loc[self.tr_perm[idx]])
        return self.batch_left, self.batch_right, self.batch_label

    def evaluate(self):
        return self.val_left, self.val_right, self.val_label
