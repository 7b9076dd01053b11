"""Host side of the training path (stereo_matching_b200/train.py): parameter layout, checkpoint round trip in
the reference's TensorBundle naming, patch sampling, learning-rate rule, and the data-parallel gradient
averaging with two gloo ranks on CPU (the N > 1 path; on GPUs the same call runs over NCCL)."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp

from oracle import train_oracle as to
from stereo_matching_b200 import checkpoint, train
from stereo_matching_b200.models.model import fold_batchnorm


def test_layout_matches_oracle_and_round_trips(tmp_path):
    for net in ("win9_dep9", "win37_dep9"):
        lay, n_p, n_s = train.layout(net)
        olay, on_p, on_s = to.param_layout(net)
        assert (n_p, n_s) == (on_p, on_s)
        for a, b in zip(lay, olay):
            assert a["kernel"] == b["kernel"] and a["gamma"] == b["gamma"] and a["moving_variance"] == b["var"]
    p, s = train.init_state("win9_dep9", seed=5)
    layers = train.state_to_layers("win9_dep9", p, s)
    p2, s2 = train.layers_to_state("win9_dep9", layers)
    assert np.array_equal(p, p2) and np.array_equal(s, s2)
    # the inference path accepts what training produces
    packed = fold_batchnorm(layers, 3, [3, 64, 64, 64, 64])
    assert packed.size == 27 * 64 + 64 + 3 * (576 * 64 + 64)
    path = train.save_checkpoint(str(tmp_path), "win9_dep9", 300, p, s)
    assert os.path.basename(path) == "ckpt-3" and checkpoint.latest_checkpoint(str(tmp_path)) == path
    p3, s3 = train.layers_to_state("win9_dep9", checkpoint.load_net_weights(str(tmp_path)))
    assert np.array_equal(p, p3) and np.array_equal(s, s3)


def test_learning_rate_and_labels_agree_with_oracle():
    for step in (1, 23999, 24000, 31999, 32000, 47999, 48000):
        assert np.isclose(train.learning_rate(step), to.learning_rate(step))
    assert np.array_equal(train.soft_labels(4, 201), to.soft_labels(4, 201))


def test_synthetic_patches_are_centred_on_the_ground_truth():
    D, p = 41, 9
    data = train.SyntheticPatches(1, p, D, batch=32, seed=1, height=120, width=400)
    left, right, labels = data.next_batch()
    assert left.shape == (32, p, p, 3) and right.shape == (32, p, p + D - 1, 3) and labels.shape == (32, D)
    # the left patch is the centre window of the right strip (same scene point, images normalised separately)
    centre = right[:, :, D // 2:D // 2 + p, :]
    corr = [np.corrcoef(left[k].ravel(), centre[k].ravel())[0, 1] for k in range(32)]
    assert np.median(corr) > 0.9
    assert labels[0].argmax() == D // 2


def _dp_worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    g = torch.tensor(rng.standard_normal(113088).astype(np.float32))
    train.average_gradients(g, world)
    np.save(os.path.join(out_dir, f"avg{rank}.npy"), g.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_average(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_dp_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a0, a1 = np.load(tmp_path / "avg0.npy"), np.load(tmp_path / "avg1.npy")
    ref = (np.random.default_rng(100).standard_normal(113088).astype(np.float32) +
           np.random.default_rng(101).standard_normal(113088).astype(np.float32)) / 2
    assert np.array_equal(a0, a1) and np.allclose(a0, ref, rtol=1e-6, atol=1e-7)
