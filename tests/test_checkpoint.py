"""K1: the TensorBundle reader against the checkpoint the reference ships
(/root/reference/new_checkpoint, copied to tests/golden/new_checkpoint)."""
import os

import numpy as np
import pytest

from stereo_matching_b200 import checkpoint as ck


def _prefix(golden_dir):
    return os.path.join(golden_dir, "new_checkpoint", "ckpt-1000")


def test_latest_checkpoint(golden_dir):
    d = os.path.join(golden_dir, "new_checkpoint")
    assert ck.latest_checkpoint(d) == os.path.join(d, "ckpt-1000")
    assert ck.latest_checkpoint(golden_dir) is None


def test_index_entries(golden_dir):
    e = ck.read_index(_prefix(golden_dir) + ".index")
    assert len(e) == 64          # 65 SSTable entries including the "" header key
    k0 = e["net/layer_with_weights-0/kernel/.ATTRIBUTES/VARIABLE_VALUE"]
    assert k0.shape == (3, 3, 3, 64) and k0.offset == 36 and k0.size == 6912 and k0.dtype == 1
    assert e["net/layer_with_weights-2/kernel/.ATTRIBUTES/VARIABLE_VALUE"].offset == 8228
    assert e["net/layer_with_weights-7/moving_variance/.ATTRIBUTES/VARIABLE_VALUE"].offset == 454180
    assert e["step/.ATTRIBUTES/VARIABLE_VALUE"].offset == 0


def test_scalars_and_ranges(golden_dir):
    t = ck.read_bundle(_prefix(golden_dir))
    assert int(t["step/.ATTRIBUTES/VARIABLE_VALUE"]) == 100000
    assert int(t["save_counter/.ATTRIBUTES/VARIABLE_VALUE"]) == 1000
    assert int(t["optimizer/iter/.ATTRIBUTES/VARIABLE_VALUE"]) == 99999
    assert float(t["optimizer/learning_rate/.ATTRIBUTES/VARIABLE_VALUE"]) == pytest.approx(1.024e-10, rel=1e-6)
    assert float(t["optimizer/beta_1/.ATTRIBUTES/VARIABLE_VALUE"]) == pytest.approx(0.9)
    assert float(t["optimizer/beta_2/.ATTRIBUTES/VARIABLE_VALUE"]) == pytest.approx(0.999)
    k = t["net/layer_with_weights-0/kernel/.ATTRIBUTES/VARIABLE_VALUE"]
    assert k.min() == pytest.approx(-0.97165, abs=1e-5) and k.max() == pytest.approx(0.53002, abs=1e-5)


def test_crc_of_first_layer(golden_dir):
    ck.read_bundle(_prefix(golden_dir), verify_crc=True,
                   keys_filter=lambda k: "layer_with_weights-0/" in k and "OPTIMIZER" not in k)


def test_net_weights_layout(ws9_layers):
    assert len(ws9_layers) == 8
    shapes = [(3, 3, 3, 64)] + [(3, 3, 64, 64)] * 3
    for i in range(4):
        assert ws9_layers[2 * i]["kernel"].shape == shapes[i]
        assert ws9_layers[2 * i]["bias"].shape == (64,)
        assert set(ws9_layers[2 * i + 1]) == {"gamma", "beta", "moving_mean", "moving_variance"}
    assert ws9_layers[1]["gamma"].min() < 0      # BN1 gamma has negative entries


def test_write_read_round_trip(golden_dir, tmp_path):
    t = ck.read_bundle(_prefix(golden_dir))
    ck.write_bundle(str(tmp_path / "rt"), t)
    t2 = ck.read_bundle(str(tmp_path / "rt"), verify_crc=False)
    assert set(t) == set(t2)
    for k in t:
        assert t[k].shape == t2[k].shape and t[k].dtype == t2[k].dtype
        assert np.array_equal(t[k], t2[k])


def test_bad_files(tmp_path):
    p = tmp_path / "bad.index"
    p.write_bytes(b"\x00" * 10)
    with pytest.raises(ck.CheckpointError):
        ck.read_index(str(p))
    p.write_bytes(b"\x00" * 64)
    with pytest.raises(ck.CheckpointError):
        ck.read_index(str(p))
    with pytest.raises(ck.CheckpointError):
        ck.load_net_weights(str(tmp_path))


def test_object_graph_of_ws9_is_byte_identical_to_the_shipped_one(golden_dir):
    """tf.train.Checkpoint.restore is object-based: it walks `_CHECKPOINTABLE_OBJECT_GRAPH`.  The graph generated for
    base_model_ws_9 (four conv/BN blocks) must be the very bytes TensorFlow wrote into ckpt-1000; string-tensor
    encoding and both checksums are verified on the shipped entry."""
    t = ck.read_bundle(_prefix(golden_dir), verify_crc=True, strings=True,
                       keys_filter=lambda k: k == ck.OBJECT_GRAPH_KEY)
    shipped = t[ck.OBJECT_GRAPH_KEY]
    assert isinstance(shipped, bytes) and len(shipped) == 8641
    assert ck.object_graph(4) == shipped


def test_saved_checkpoint_has_the_structure_tf_restores(golden_dir, tmp_path):
    """`train.save_checkpoint` writes every key of ckpt-1000 with the same dtype and shape (object graph, Adam slots,
    iter, hyper-parameters, save_counter, int32 step), keeps at most three checkpoints and lists them in the state
    file, and a resume brings the Adam moments back."""
    from stereo_matching_b200 import train
    rng = np.random.default_rng(0)
    p, s = train.init_state("win9_dep9", seed=5)
    m, v = rng.standard_normal(p.size).astype(np.float32), rng.random(p.size).astype(np.float32)
    for step in (100, 200, 300, 400):
        path = train.save_checkpoint(str(tmp_path), "win9_dep9", step, p, s, m, v)
    ref = ck.read_index(_prefix(golden_dir) + ".index")
    got = ck.read_index(path + ".index")
    assert set(got) == set(ref)
    for k in ref:
        assert (got[k].dtype, got[k].shape) == (ref[k].dtype, ref[k].shape), k
    t = ck.read_bundle(path, verify_crc=True, strings=True)
    assert t[ck.OBJECT_GRAPH_KEY] == ck.object_graph(4)
    assert int(t["step/.ATTRIBUTES/VARIABLE_VALUE"]) == 400 and int(t["save_counter/.ATTRIBUTES/VARIABLE_VALUE"]) == 4
    state = (tmp_path / "checkpoint").read_text().splitlines()
    assert state[0] == 'model_checkpoint_path: "ckpt-4"'
    assert state[1:] == ['all_model_checkpoint_paths: "ckpt-%d"' % i for i in (2, 3, 4)]
    assert not (tmp_path / "ckpt-1.index").exists() and (tmp_path / "ckpt-2.index").exists()
    st = train.load_checkpoint_state(path, "win9_dep9")
    assert st["step"] == 400 and np.array_equal(st["params"], p) and np.array_equal(st["stats"], s)
    assert np.array_equal(st["adam_m"], m) and np.array_equal(st["adam_v"], v)
    # base_model: nine blocks, Keras names skip the dead eighth pair (models/model.py:21-22)
    p37, s37 = train.init_state("win37_dep9", seed=1)
    path37 = train.save_checkpoint(str(tmp_path / "b"), "win37_dep9", 100, p37, s37)
    g = ck.read_bundle(path37, strings=True, keys_filter=lambda k: k == ck.OBJECT_GRAPH_KEY)[ck.OBJECT_GRAPH_KEY]
    assert b"conv2d_9/kernel" in g and b"conv2d_7/kernel" not in g and b"layer_with_weights-17/moving_variance" in g
    # the shipped checkpoint resumes too (its Adam slots are read)
    st = train.load_checkpoint_state(_prefix(golden_dir), "win9_dep9")
    assert st["step"] == 100000 and st["adam_v"] is not None and float(st["adam_v"].max()) > 0
