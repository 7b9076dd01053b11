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
