"""Patch-level evaluation - host mirror of /root/reference/evaluate.py.

Validation patches through the trunk in inference mode (`model(x, training=False)`, evaluate.py:80-81), inner
product of the left feature with the D right features (evaluate.py:11-17,84), arg max, and the share of patches
whose prediction is within 3 of the label (evaluate.py:86-98).  The trunk runs in the CUDA library; the
inner product over [N, D, 64] features is one numpy einsum.  Like the reference it evaluates the `val_*.bin`
patches through `Data_handler.evaluate()`; `--synthetic N` draws the validation patches from N generated pairs
instead (KITTI is not shipped).
"""
import argparse
import sys

import numpy as np

from . import train
from .models.model import base_model, base_model_ws_9


def map_inner_product(lmap: np.ndarray, rmap: np.ndarray) -> np.ndarray:
    """evaluate.py:11-17: [N,1,1,C] x [N,1,D,C] -> [N, D]."""
    return np.einsum("nc,ndc->nd", lmap[:, 0, 0, :], rmap[:, 0, :, :])


def evaluate(model, lpatch, rpatch, patch_targets, eval_size: int = 200, verbose: bool = True) -> float:
    labels = np.argmax(patch_targets, axis=1)                              # evaluate.py:78
    predicted = np.empty(len(lpatch), np.int64)
    for i in range(0, len(lpatch), eval_size):                             # batches only bound the memory in use
        lf = model(lpatch[i:i + eval_size], training=False)
        rf = model(rpatch[i:i + eval_size], training=False)
        predicted[i:i + eval_size] = np.argmax(map_inner_product(lf, rf), axis=1)
    acc_count = 0
    for i in range(0, len(lpatch), eval_size):                             # evaluate.py:92-96
        acc_count += int(np.sum(np.abs(predicted[i:i + eval_size] - labels[i:i + eval_size]) <= 3))
        if verbose:
            print("iter. %d finished, with %d correct (3-pixel error)" % (i + 1, acc_count))
    acc = acc_count / len(lpatch) * 100
    if verbose:
        print("accuracy: %.3f" % acc)
    return acc


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--batch_size", type=int, default=128)
    ap.add_argument("--model_dir", default="checkpoint")
    ap.add_argument("--data_version", default="kitti2015")
    ap.add_argument("--data_root", default="./kitti2015/training")
    ap.add_argument("--util_root", default="./preprocess/debug_15")
    ap.add_argument("--net_type", default="win37_dep9", choices=["win37_dep9", "win9_dep9"])
    ap.add_argument("--eval_size", type=int, default=200)
    ap.add_argument("--patch_size", type=int, default=37)
    ap.add_argument("--num_val_loc", type=int, default=1000)
    ap.add_argument("--disp_range", type=int, default=201)
    ap.add_argument("--num_tr_img", type=int, default=160)
    ap.add_argument("--num_val_img", type=int, default=40)
    ap.add_argument("--synthetic", type=int, default=0,
                    help="0 (default): the val_*.bin patches through data_handler.Data_handler, like the reference; "
                         "N > 0: validation patches drawn from N generated pairs")
    args = ap.parse_args(argv)
    model = (base_model if args.net_type == "win37_dep9" else base_model_ws_9)((None, None, 3))
    model.load_checkpoint(args.model_dir)
    if args.synthetic > 0:
        print("NOTICE: evaluating on %d generated pairs, not on KITTI (--synthetic)" % args.synthetic)
        data = train.SyntheticPatches(args.synthetic, args.patch_size, args.disp_range, args.num_val_loc, seed=321)
        lpatch, rpatch, targets = data.next_batch()
    else:
        from .data_handler import Data_handler
        np.random.seed(123)                                            # evaluate.py:42
        dhandler = Data_handler(args.data_version, args.data_root, args.util_root, args.num_tr_img, args.num_val_img,
                                args.num_val_loc, args.batch_size, args.patch_size, args.disp_range)
        lpatch, rpatch, targets = dhandler.evaluate()                  # evaluate.py:77
    evaluate(model, lpatch, rpatch, targets, args.eval_size)
    return 0


if __name__ == "__main__":
    sys.exit(main())
