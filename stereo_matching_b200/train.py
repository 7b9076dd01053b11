"""Patch training - host mirror of /root/reference/train.py (+ data_handler.py).

Same flag names and defaults as the reference (train.py:21-40); the step itself (two tower passes in
training mode, inner product, softmax cross-entropy, gradients, Adam) runs in the CUDA library behind
`stereo_train_*` (csrc/train.cu).  KITTI and the preprocessed location files are not available in this
environment, so `--synthetic N` samples the patches from N synthetic pairs (stereo_matching_b200.synth)
with the validity rule of preprocess/kitti2015_gene_loc_1.m:59-63.

Data-parallel training (`torchrun --nproc-per-node N -m stereo_matching_b200.train ...`): every rank runs
forward/backward on its own batch, the flat gradient array (113 088 floats for win9_dep9) is averaged with
ONE all-reduce (NCCL over NVLink on GPUs) and every rank applies the same Adam step.  As in any data-parallel
BatchNorm without cross-replica sync, the batch statistics are per rank.

Checkpoints are TensorBundles holding what the reference's `tf.train.Checkpoint(step, optimizer, net)` saves,
under the same keys as the shipped ckpt-1000: the object graph TensorFlow's object-based restore walks (for
base_model_ws_9 byte for byte the shipped one), the net's variables, Adam's iteration count, hyper-parameters and
m / v slots, `step`, `save_counter`; at most three are kept (`max_to_keep=3`, train.py:81).  TensorFlow is not
available here, so "the reference restores them" is checked structurally (same keys, dtypes, shapes and graph
as ckpt-1000, tests/test_checkpoint.py), not by running `ckpt.restore`.
"""
import argparse
import os
import sys
import time
from typing import Dict, List, Tuple

import numpy as np

from . import _lib, checkpoint, synth

_SUFFIX = "/.ATTRIBUTES/VARIABLE_VALUE"
NETS = {"win9_dep9": (_lib.NET_WIN9, 3, [3, 64, 64, 64, 64]),
        "win37_dep9": (_lib.NET_WIN37, 5, [3, 32, 32, 32, 64, 64, 64, 64, 64, 64])}


def learning_rate(step: int, base: float = 1e-3) -> float:
    """train.py:125-130: divided by 5 at step 24000 and again every 8000 steps."""
    return base if step < 24000 else base / 5.0 ** (1 + (step - 24000) // 8000)


def soft_labels(batch: int, disp_range: int) -> np.ndarray:
    """data_handler.py:78-85."""
    t = np.zeros((batch, disp_range), np.float32)
    for k, wgt in enumerate([0.05, 0.2, 0.5, 0.2, 0.05]):
        t[:, disp_range // 2 - 2 + k] = wgt
    return t


# ---- flat parameter arrays <-> the checkpoint's layer dicts ------------------------------------------------
def layout(net_type: str):
    _, ks, ch = NETS[net_type]
    out, po, so = [], 0, 0
    for l in range(len(ch) - 1):
        cin, cout = ch[l], ch[l + 1]
        d = {"shape": (ks, ks, cin, cout), "kernel": slice(po, po + ks * ks * cin * cout)}
        po += ks * ks * cin * cout
        for name in ("bias", "gamma", "beta"):
            d[name] = slice(po, po + cout)
            po += cout
        d["moving_mean"], d["moving_variance"] = slice(so, so + cout), slice(so + cout, so + 2 * cout)
        so += 2 * cout
        out.append(d)
    return out, po, so


def init_state(net_type: str, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Keras initialisers: Glorot-uniform kernels, zero bias, gamma 1, beta 0, moving mean 0 / variance 1."""
    lay, n_p, n_s = layout(net_type)
    rng = np.random.default_rng(seed)
    p, s = np.zeros(n_p, np.float32), np.zeros(n_s, np.float32)
    for d in lay:
        kh, kw, cin, cout = d["shape"]
        lim = np.sqrt(6.0 / (kh * kw * (cin + cout)))
        p[d["kernel"]] = rng.uniform(-lim, lim, kh * kw * cin * cout).astype(np.float32)
        p[d["gamma"]] = 1.0
        s[d["moving_variance"]] = 1.0
    return p, s


def state_to_layers(net_type: str, params: np.ndarray, stats: np.ndarray) -> List[Dict[str, np.ndarray]]:
    """[conv, bn, conv, bn, ...] dicts as `checkpoint.load_net_weights` returns them."""
    layers = []
    for d in layout(net_type)[0]:
        layers.append({"kernel": params[d["kernel"]].reshape(d["shape"]).copy(), "bias": params[d["bias"]].copy()})
        layers.append({"gamma": params[d["gamma"]].copy(), "beta": params[d["beta"]].copy(),
                       "moving_mean": stats[d["moving_mean"]].copy(), "moving_variance": stats[d["moving_variance"]].copy()})
    return layers


def layers_to_state(net_type: str, layers: List[Dict[str, np.ndarray]]) -> Tuple[np.ndarray, np.ndarray]:
    lay, n_p, n_s = layout(net_type)
    p, s = np.zeros(n_p, np.float32), np.zeros(n_s, np.float32)
    for i, d in enumerate(lay):
        conv, bn = layers[2 * i], layers[2 * i + 1]
        p[d["kernel"]] = np.asarray(conv["kernel"], np.float32).reshape(-1)
        p[d["bias"]] = conv["bias"]
        p[d["gamma"]], p[d["beta"]] = bn["gamma"], bn["beta"]
        s[d["moving_mean"]], s[d["moving_variance"]] = bn["moving_mean"], bn["moving_variance"]
    return p, s


def _keras_names(net_type: str):
    """Keras' automatic layer names.  base_model builds a Conv2D + BatchNormalization pair that never enters the
    graph (models/model.py:21-22), so its eighth block is conv2d_8 / batch_normalization_8."""
    n = len(NETS[net_type][2]) - 1
    idx = list(range(n)) if net_type == "win9_dep9" else [i for i in range(n + 1) if i != 7]
    return (["conv2d" if i == 0 else "conv2d_%d" % i for i in idx],
            ["batch_normalization" if i == 0 else "batch_normalization_%d" % i for i in idx])


def save_checkpoint(model_dir: str, net_type: str, step: int, params: np.ndarray, stats: np.ndarray,
                    adam_m: np.ndarray = None, adam_v: np.ndarray = None, lr: float = None,
                    max_to_keep: int = 3) -> str:
    """`manager.save()` of train.py:82,121: `ckpt-<n>` + the `checkpoint` state file, at most `max_to_keep`
    checkpoints kept (tf.train.CheckpointManager(..., max_to_keep=3)).

    The bundle holds what `tf.train.Checkpoint(step, optimizer, net)` saves, under the same keys as the shipped
    ckpt-1000: the object graph (restore is object-based and needs it), `step` (int32), `save_counter`, the net's
    variables, Adam's `iter`, hyper-parameters and the `m` / `v` slot of every trainable variable (zeros when the
    caller has none, e.g. weights exported without an optimizer state)."""
    os.makedirs(model_dir, exist_ok=True)
    lay = layout(net_type)[0]
    counter = step // 100 if step % 100 == 0 else step
    m = np.zeros_like(params) if adam_m is None else np.asarray(adam_m, np.float32)
    v = np.zeros_like(params) if adam_v is None else np.asarray(adam_v, np.float32)
    conv_names, bn_names = _keras_names(net_type)
    tensors = {checkpoint.OBJECT_GRAPH_KEY: checkpoint.object_graph(len(lay), conv_names, bn_names),
               "step" + _SUFFIX: np.asarray(step, np.int32),
               "save_counter" + _SUFFIX: np.asarray(counter, np.int64),
               "optimizer/iter" + _SUFFIX: np.asarray(step, np.int64),
               "optimizer/beta_1" + _SUFFIX: np.asarray(0.9, np.float32),
               "optimizer/beta_2" + _SUFFIX: np.asarray(0.999, np.float32),
               "optimizer/decay" + _SUFFIX: np.asarray(0.0, np.float32),
               "optimizer/learning_rate" + _SUFFIX: np.asarray(learning_rate(step) if lr is None else lr, np.float32)}
    for k, layer in enumerate(state_to_layers(net_type, params, stats)):
        for name, arr in layer.items():
            tensors[f"net/layer_with_weights-{k}/{name}{_SUFFIX}"] = arr
    for k, d in enumerate(lay):
        for layer_k, name in ((2 * k, "kernel"), (2 * k, "bias"), (2 * k + 1, "gamma"), (2 * k + 1, "beta")):
            shape = d["shape"] if name == "kernel" else (-1,)
            stem = f"net/layer_with_weights-{layer_k}/{name}/.OPTIMIZER_SLOT/optimizer/"
            tensors[stem + "m" + _SUFFIX] = m[d[name]].reshape(shape)
            tensors[stem + "v" + _SUFFIX] = v[d[name]].reshape(shape)
    name = f"ckpt-{counter}"
    checkpoint.write_bundle(os.path.join(model_dir, name), tensors)
    # the state file lists every kept checkpoint, oldest first; older ones are deleted (max_to_keep)
    state = os.path.join(model_dir, "checkpoint")
    kept = []
    if os.path.exists(state):
        with open(state) as f:
            kept = [ln.split(":", 1)[1].strip().strip('"') for ln in f if ln.startswith("all_model_checkpoint_paths:")]
    kept = [k for k in kept if k != name] + [name]
    while max_to_keep and len(kept) > max_to_keep:
        old = kept.pop(0)
        for ext in (".index", ".data-00000-of-00001"):
            try:
                os.remove(os.path.join(model_dir, old + ext))
            except OSError:
                pass
    with open(state, "w") as f:
        f.write(f'model_checkpoint_path: "{name}"\n')
        for k in kept:
            f.write(f'all_model_checkpoint_paths: "{k}"\n')
    return os.path.join(model_dir, name)


def load_checkpoint_state(prefix: str, net_type: str) -> Dict[str, np.ndarray]:
    """What `ckpt.restore(...)` of train.py:84 brings back: params, stats, Adam slots (None when the bundle has
    none) and the step."""
    params, stats = layers_to_state(net_type, checkpoint.load_net_weights(prefix))
    b = checkpoint.read_bundle(prefix, keys_filter=lambda k: "OPTIMIZER_SLOT" in k or k.startswith("step/"))
    lay = layout(net_type)[0]
    m, v = np.zeros_like(params), np.zeros_like(params)
    found = 0
    for k, d in enumerate(lay):
        for layer_k, name in ((2 * k, "kernel"), (2 * k, "bias"), (2 * k + 1, "gamma"), (2 * k + 1, "beta")):
            stem = f"net/layer_with_weights-{layer_k}/{name}/.OPTIMIZER_SLOT/optimizer/"
            if stem + "m" + _SUFFIX in b and stem + "v" + _SUFFIX in b:
                m[d[name]] = np.asarray(b[stem + "m" + _SUFFIX], np.float32).reshape(-1)
                v[d[name]] = np.asarray(b[stem + "v" + _SUFFIX], np.float32).reshape(-1)
                found += 1
    step = int(b.get("step" + _SUFFIX, 0))
    have = found == 4 * len(lay)
    return {"params": params, "stats": stats, "adam_m": m if have else None, "adam_v": v if have else None, "step": step}


# ---- patches -----------------------------------------------------------------------------------------------
class SyntheticPatches:
    """data_handler.py on synthetic pairs: per-image normalisation (`:68-69`), left patch p x p around
    (x, y), right patch p x (p + D - 1) around (x - d, y) with the ground-truth disparity d, horizontal
    locations only (loc_type 1).  A location is valid if the ground truth is known there and both patches lie
    inside the images (kitti2015_gene_loc_1.m:59-63)."""

    def __init__(self, n_pairs: int, patch: int, disp_range: int, batch: int, seed: int = 123,
                 height: int = 375, width: int = 1242):
        self.patch, self.D, self.batch = patch, disp_range, batch
        self.half_patch, self.half_range = patch // 2, disp_range // 2
        self.rng = np.random.default_rng(seed)
        self.left, self.right, locs = [], [], []
        for i in range(n_pairs):
            l, r, gt = synth.make_pair(5000 + i, height, width, min(disp_range, 201))
            l, r = l.astype(np.float64), r.astype(np.float64)
            self.left.append(((l - l.mean()) / l.std()).astype(np.float32))
            self.right.append(((r - r.mean()) / r.std()).astype(np.float32))
            ys, xs = np.nonzero(gt)
            d = gt[ys, xs].astype(np.int64)
            hp, hr = self.half_patch, self.half_range
            ok = ((ys >= hp) & (ys < height - hp) & (xs >= hp) & (xs < width - hp) &
                  (xs - d - hp - hr >= 0) & (xs - d + hp + hr < width))
            locs.append(np.stack([np.full(ok.sum(), i), xs[ok], ys[ok], xs[ok] - d[ok]], axis=1))
        self.locs = np.concatenate(locs)
        if len(self.locs) < batch:
            raise ValueError("not enough valid patch locations")
        self.perm = self.rng.permutation(len(self.locs))
        self.ptr = 0
        self.labels = soft_labels(batch, disp_range)

    def next_batch(self):
        p, hp, hr = self.patch, self.half_patch, self.half_range
        bl = np.empty((self.batch, p, p, 3), np.float32)
        br = np.empty((self.batch, p, p + self.D - 1, 3), np.float32)
        for k in range(self.batch):
            if self.ptr >= len(self.perm):
                self.perm, self.ptr = self.rng.permutation(len(self.locs)), 0
            img, x, y, xr = self.locs[self.perm[self.ptr]]
            self.ptr += 1
            bl[k] = self.left[img][y - hp:y + hp + 1, x - hp:x + hp + 1]
            br[k] = self.right[img][y - hp:y + hp + 1, xr - hp - hr:xr + hp + hr + 1]
        return bl, br, self.labels


# ---- data-parallel plumbing --------------------------------------------------------------------------------
def average_gradients(grads, world_size: int, group=None):
    """In-place mean over ranks of a flat gradient tensor (torch, on the device of the backend)."""
    import torch.distributed as dist
    if world_size > 1:
        dist.all_reduce(grads, op=dist.ReduceOp.SUM, group=group)
        grads /= world_size
    return grads


class DeviceGradView:
    """The trainer's device gradient array as a torch tensor (no copy), for torch.distributed."""

    def __init__(self, trainer: "_lib.Trainer", device: int):
        import torch

        class _Cai:                                   # __cuda_array_interface__ v2
            pass
        holder = _Cai()
        holder.__cuda_array_interface__ = {"shape": (trainer.n_params,), "typestr": "<f4",
                                           "data": (trainer.grads_dev(), False), "version": 2}
        self.tensor = torch.as_tensor(holder, device=f"cuda:{device}")


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--batch_size", type=int, default=128)
    ap.add_argument("--num_iter", type=int, default=1000)
    ap.add_argument("--max_epoch", type=int, default=120)
    ap.add_argument("--model_dir", default="new_checkpoint")
    ap.add_argument("--data_version", default="kitti2015")
    ap.add_argument("--data_root", default="./kitti2015/training")
    ap.add_argument("--util_root", default="./preprocess/debug_15_ws_9")
    ap.add_argument("--net_type", default="win9_dep9", choices=sorted(NETS))
    ap.add_argument("--patch_size", type=int, default=9)
    ap.add_argument("--disp_range", type=int, default=201)
    ap.add_argument("--eval_size", type=int, default=200)
    ap.add_argument("--num_tr_img", type=int, default=160)
    ap.add_argument("--num_val_img", type=int, default=40)
    ap.add_argument("--num_val_loc", type=int, default=5000)
    ap.add_argument("--phase", default="train")
    ap.add_argument("--synthetic", type=int, default=0,
                    help="0 (default): KITTI through data_handler.Data_handler, like the reference; N > 0: sample the "
                         "patches from N generated pairs instead")
    ap.add_argument("--seed", type=int, default=123)
    args = ap.parse_args(argv)
    if args.data_version != "kitti2015":
        # the reference builds the towers with 3 input channels whatever the data version (train.py:67-70), so its
        # own kitti2012 training stops at the first tower call; nothing to mirror
        ap.error("training takes the 3-channel kitti2015 data (the reference's towers are (None, None, 3), train.py:67-70)")

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl")
    net, _, _ = NETS[args.net_type]
    ctx = _lib.Context(local_rank)
    tr = _lib.Trainer(ctx, net, args.batch_size, args.disp_range)
    if tr.patch != args.patch_size:
        ap.error(f"--patch_size {args.patch_size} does not match {args.net_type} (receptive field {tr.patch})")
    latest = checkpoint.latest_checkpoint(args.model_dir) if os.path.isdir(args.model_dir) else None
    step = 0
    if latest:
        st = load_checkpoint_state(latest, args.net_type)
        step = st["step"]
        tr.set_state(st["params"], st["stats"], st["adam_m"], st["adam_v"], step=step)
        if rank == 0:
            print(f"Restored from {latest}")
    else:
        params, stats = init_state(args.net_type, args.seed)
        tr.set_state(params, stats, step=0)
        if rank == 0:
            print("Initializing from scratch.")
    if args.synthetic > 0:
        data = SyntheticPatches(args.synthetic, args.patch_size, args.disp_range, args.batch_size, seed=args.seed + rank)
    else:
        from .data_handler import Data_handler
        np.random.seed(args.seed + rank)                       # train.py:43 (rank 0 = the reference's order)
        data = Data_handler(args.data_version, args.data_root, args.util_root, args.num_tr_img, args.num_val_img,
                            args.num_val_loc, args.batch_size, args.patch_size, args.disp_range,
                            val_npy="val.npy" if rank == 0 else "")
    gview = DeviceGradView(tr, local_rank).tensor if world > 1 else None

    acc_total = 0
    for epoch in range(args.max_epoch):
        acc = 0
        t0 = time.time()
        for _ in range(args.num_iter):
            left, right, labels = data.next_batch()
            loss, correct, _ = tr.forward_backward(left, right, labels)
            if world > 1:
                average_gradients(gview, world)
                import torch
                torch.cuda.synchronize()
            step += 1
            tr.apply(learning_rate(step))
            acc += correct
            if step % 100 == 0 and rank == 0:
                st = tr.get_state()
                path = save_checkpoint(args.model_dir, args.net_type, step, st["params"], st["stats"],
                                       st["adam_m"], st["adam_v"])
                print(f"Saved checkpoint for step {step}: {path}")
                print("Loss at step: %d: %.6f" % (step, loss))
        if rank == 0:
            print("Epoch %d finished, with accuracy: %f (%.1f steps/s)" %
                  (epoch + 1, acc / (args.batch_size * args.num_iter), args.num_iter / (time.time() - t0)))
        acc_total += acc
    if rank == 0:
        print("Accuracy: ", acc_total / (args.max_epoch * args.num_iter * args.batch_size) * 100)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
