"""Drop-in for /root/reference/inference.py: same flags, same output files.

    python -m stereo_matching_b200.inference --model_dir new_checkpoint --out_dir results \
        --data_root ./kitti2015/training --disp_range 201

writes, like inference.py:202-231,
    <out>/left_disp/left_%06d_10.png   <out>/right_disp/right_%06d_10.png   <out>/gt/true_%06d_10.png
    <out>/left_cost/left_%06d_10_H_W_D.bin   <out>/right_cost/right_%06d_10_H_W_D.bin   (float32 [H][W][D])
and prints the per-image and mean 3-pixel error.  KITTI and `val.npy` are not
shipped with the reference, so `--synthetic N` runs N generated pairs instead.
The reference never creates the sub-directories; this script does.
"""
from __future__ import annotations

import argparse
import os

import numpy as np

from . import _lib, shard, synth
from .metric import three_pixel_error
from .models.model import base_model, base_model_ws_9


def bin_name(side: str, file_id: int, h: int, w: int, d: int) -> str:
    # inference.py:217,228 - parsed back by post.lua:81-87
    return "%s_%06d_10_%d_%d_%d.bin" % (side, file_id, h, w, d)


def build_parser():
    p = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    # flag names and defaults of inference.py:39-58 (dead flags kept for compatibility)
    p.add_argument("--batch_size", type=int, default=128)
    p.add_argument("--num_iter", type=int, default=40000)
    p.add_argument("--model_dir", default="new_checkpoint")
    p.add_argument("--out_dir", default="results")
    p.add_argument("--data_version", default="kitti2015")
    p.add_argument("--data_root", default="./kitti2015/training")
    p.add_argument("--util_root", default="./preprocess/debug_15")
    p.add_argument("--net_type", default="win9_dep9")
    p.add_argument("--eval_size", type=int, default=200)
    p.add_argument("--num_tr_img", type=int, default=160)
    p.add_argument("--num_val_img", type=int, default=40)
    p.add_argument("--patch_size", type=int, default=9)
    p.add_argument("--num_val_loc", type=int, default=10000)
    p.add_argument("--disp_range", type=int, default=201)
    p.add_argument("--num_imgs", type=int, default=5)
    p.add_argument("--start_id", type=int, default=0)
    # additions
    p.add_argument("--synthetic", type=int, default=0, help="run this many generated pairs instead of KITTI")
    p.add_argument("--height", type=int, default=375)
    p.add_argument("--width", type=int, default=1242)
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--no_bin", action="store_true", help="skip the .bin cost volumes")
    p.add_argument("--val_ids", default="val.npy")
    p.add_argument("--precision", default="bf16x3", choices=sorted(PRECISIONS),
                   help="bf16x3 (default): tcgen05 kernels, split-bf16 operands, fp32 accumulate, within the 1e-4 "
                        "cost-volume gate; fp32: CUDA-core kernels; bf16: single pass, reports its own tolerance")
    return p


PRECISIONS = {"fp32": _lib.PREC_FP32, "bf16x3": _lib.PREC_BF16X3, "bf16": _lib.PREC_BF16}


def _world():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, None) for a plain `python -m` run."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1, None
    return int(os.environ["RANK"]), world, int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))


def _read_kitti(flags, file_id):
    """(left, right) uint8 [h,w,3] and the uint8 ground truth (or None).

    kitti2015: image_2 / image_3 in RGB order (scipy.misc.imread, inference.py:102-103), GT = cv2 GRAYSCALE of the
    16-bit disp_noc_0 PNG, i.e. its high byte (inference.py:104).
    kitti2012: grey image_0 / image_1 (inference.py:105-107).  The reference normalises the grey image and
    assigns it into its 3-channel padded array (inference.py:129-130), which broadcasts the one channel into
    all three; replicating the uint8 grey values gives the same mean, std and tower input.  The reference never
    reads a ground truth in this mode (its metric line then fails on an undefined name); here `disp_noc/` is
    used when it exists and the error is reported as nan when it does not."""
    import cv2
    if flags.data_version == "kitti2015":
        l = cv2.imread("%s/image_2/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_COLOR)
        r = cv2.imread("%s/image_3/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_COLOR)
        gt = cv2.imread("%s/disp_noc_0/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_GRAYSCALE)
        if l is None or r is None or gt is None:
            raise FileNotFoundError("KITTI files for id %06d not found under %s" % (file_id, flags.data_root))
        return l[:, :, ::-1].copy(), r[:, :, ::-1].copy(), gt
    l = cv2.imread("%s/image_0/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_GRAYSCALE)
    r = cv2.imread("%s/image_1/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_GRAYSCALE)
    if l is None or r is None:
        raise FileNotFoundError("KITTI 2012 files for id %06d not found under %s" % (file_id, flags.data_root))
    gt = cv2.imread("%s/disp_noc/%06d_10.png" % (flags.data_root, file_id), cv2.IMREAD_GRAYSCALE)
    return np.repeat(l[:, :, None], 3, axis=2), np.repeat(r[:, :, None], 3, axis=2), gt


def run(flags) -> float:
    import cv2
    if flags.data_version not in ("kitti2015", "kitti2012"):
        raise SystemExit("data_version should be either kitti2012 or kitti2015")
    if flags.disp_range > 256:
        raise SystemExit("--disp_range %d: disparity maps are bytes (the reference writes uint8 PNGs), at most 256"
                         % flags.disp_range)
    # launched with torchrun (one process per GPU): the pair list is sharded round-robin over the ranks
    # (shard.py; the reference's loop, inference.py:99, carries no state between pairs), results meet on rank 0
    rank, world, local_rank = _world()
    if world > 1:
        import torch.distributed as dist
        import torch
        if not dist.is_initialized():
            dist.init_process_group("gloo")          # only the per-pair errors travel; no data-path collective
        # one GPU per rank; ranks beyond the visible GPUs share them (a 2-rank run on a 1-GPU box still works)
        flags.device = local_rank % max(1, torch.cuda.device_count())
    precision = PRECISIONS[flags.precision]
    if flags.net_type == "win37_dep9":
        model = base_model((None, None, 3), device=flags.device, precision=precision)
    else:
        model = base_model_ws_9((None, None, 3), device=flags.device, precision=precision)
    if flags.patch_size != model.shrink + 1:
        raise SystemExit("--patch_size %d does not match net_type %s (receptive field %d)"
                         % (flags.patch_size, flags.net_type, model.shrink + 1))
    if flags.net_type == "win37_dep9" and flags.synthetic and not os.path.isdir(flags.model_dir):
        model.set_layers(synth.make_base_model_weights())
        print("Initializing from seeded synthetic weights (the reference ships none for base_model).")
    else:
        model.load_checkpoint(flags.model_dir)

    if flags.synthetic:
        file_ids = list(range(flags.synthetic))
    else:
        with open(flags.val_ids, "rb") as f:
            file_ids = [int(v) for v in np.load(f)][:flags.num_val_img]
    for sub in ("left_disp", "right_disp", "gt", "left_cost", "right_cost"):
        os.makedirs(os.path.join(flags.out_dir, sub), exist_ok=True)

    ctx = model.context
    all_ids = file_ids
    mine = shard.shard_indices(len(all_ids), rank, world)
    errors = []
    for i, file_id in ((k, all_ids[k]) for k in mine):
        if flags.synthetic:
            linput, rinput, disp_targets = synth.make_pair(1000 + file_id, flags.height, flags.width, flags.disp_range)
        else:
            linput, rinput, disp_targets = _read_kitti(flags, file_id)
        out = ctx.pairs(model.net, linput, rinput, flags.disp_range, model.precision,
                        lr_check=False, volumes=not flags.no_bin)
        left_image_int, right_image_int = out["left_disp"][0], out["right_disp"][0]
        if disp_targets is None:
            error_id = float("nan")
        else:
            error_id = three_pixel_error(left_image_int.astype(np.int64), disp_targets)
        errors.append(error_id)
        print(("%06d" % file_id) + "_10" + "\t" + ("%f" % error_id))
        cv2.imwrite("%s/left_disp/left_%06d_10.png" % (flags.out_dir, file_id), left_image_int)
        cv2.imwrite("%s/right_disp/right_%06d_10.png" % (flags.out_dir, file_id), right_image_int)
        if disp_targets is not None:
            cv2.imwrite("%s/gt/true_%06d_10.png" % (flags.out_dir, file_id), disp_targets)
        if not flags.no_bin:
            h, w = left_image_int.shape
            out["left_cost"][0].tofile(os.path.join(flags.out_dir, "left_cost", bin_name("left", file_id, h, w, flags.disp_range)))
            out["right_cost"][0].tofile(os.path.join(flags.out_dir, "right_cost", bin_name("right", file_id, h, w, flags.disp_range)))
        print("Image %s processed." % (i + 1))
    # inference.py:236 divides by FLAGS.num_val_img; the synthetic mode by the number of generated pairs
    n_items = len(all_ids) if flags.synthetic else flags.num_val_img
    mean = float(sum(shard.gather_results(errors, len(all_ids), rank, world)) / max(n_items, 1))
    if rank == 0:
        print("Mean Error: ", mean)
    return mean


def main(argv=None):
    run(build_parser().parse_args(argv))


if __name__ == "__main__":
    main()
