"""The N>1 path on CPU: world_size-2 gloo processes shard a list of pairs, each
runs the (CPU oracle) per-pair work on its shard, and the gathered result equals
the single-process result."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from stereo_matching_b200.shard import shard_indices


def test_shard_indices_cover_everything_once():
    for n in (0, 1, 5, 64, 65):
        for world in (1, 2, 3, 8):
            seen = sorted(i for r in range(world) for i in shard_indices(n, r, world))
            assert seen == list(range(n))
            sizes = [len(shard_indices(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_indices(4, 2, 2)


def _worker(rank, world, port, n_pairs, out_dir):
    import torch.distributed as dist
    from oracle import stereo_oracle as so
    from stereo_matching_b200 import synth
    from stereo_matching_b200.shard import gather_results, mean_error
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard_indices(n_pairs, rank, world)
    errs, maps = [], []
    for i in mine:
        l, r, gt = synth.make_pair(2000 + i, 12, 40, 8)
        # a cheap stand-in for the per-pair GPU work: WTA over raw-pixel correlation
        lv, rv = so.cost_volumes(l[None].astype(np.float64), r[None].astype(np.float64), 8)
        d = so.wta(lv)
        maps.append(d)
        errs.append(so.three_pixel_error(d, gt))
    all_maps = gather_results(maps, n_pairs, rank, world)
    m = mean_error(errs, n_pairs, rank, world)
    if rank == 0:
        np.savez(os.path.join(out_dir, "gathered.npz"), maps=np.stack(all_maps), mean=m)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_run_matches_single_process(tmp_path):
    from oracle import stereo_oracle as so
    from stereo_matching_b200 import synth
    n_pairs = 5
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, n_pairs, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "gathered.npz")
    ref_maps, ref_err = [], []
    for i in range(n_pairs):
        l, r, gt = synth.make_pair(2000 + i, 12, 40, 8)
        lv, _ = so.cost_volumes(l[None].astype(np.float64), r[None].astype(np.float64), 8)
        d = so.wta(lv)
        ref_maps.append(d)
        ref_err.append(so.three_pixel_error(d, gt))
    assert np.array_equal(got["maps"], np.stack(ref_maps))
    assert float(got["mean"]) == pytest.approx(sum(ref_err) / n_pairs, abs=1e-12)
