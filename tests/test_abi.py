"""CPU-side checks of the boundary: the shared library loads and exports every
symbol include/stereo_b200.h declares, and the ctypes table matches the header.
No compute entry point is called (there is no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from stereo_matching_b200 import _lib


def _header_functions():
    text = open(os.path.join(ROOT, "include", "stereo_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(stereo_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = _header_functions()
    for must in ("stereo_create", "stereo_destroy", "stereo_set_weights", "stereo_trunk", "stereo_cost_volume",
                 "stereo_wta", "stereo_corr_wta", "stereo_lr_check", "stereo_pairs", "stereo_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        from stereo_matching_b200 import build
        build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _header_functions():
        assert hasattr(lib, name), name


def test_ctypes_table_covers_the_header():
    assert sorted(_lib.SIGNATURES) == _header_functions()


def test_no_gpu_is_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.StereoError):
        _lib.Context(0)


def test_bn_fold_matches_unfolded_network(ws9_layers):
    import numpy as np
    import torch
    import torch.nn.functional as F
    from oracle import stereo_oracle as so
    from stereo_matching_b200.models.model import fold_batchnorm
    p = fold_batchnorm(ws9_layers, 3, [3, 64, 64, 64, 64])
    assert p.size == 112576
    x = np.random.default_rng(0).standard_normal((1, 16, 22, 3)).astype(np.float32)
    ref = so.trunk_forward(x, ws9_layers, "win9_dep9", np.float64)
    ch = [3, 64, 64, 64, 64]
    off = 0
    a = torch.from_numpy(x).double().permute(0, 3, 1, 2)
    for i in range(4):
        nk = 9 * ch[i] * ch[i + 1]
        k = torch.from_numpy(p[off:off + nk].reshape(3, 3, ch[i], ch[i + 1])).double().permute(3, 2, 0, 1)
        off += nk
        b = torch.from_numpy(p[off:off + ch[i + 1]]).double()
        off += ch[i + 1]
        a = F.conv2d(a, k, b)
        if i < 3:
            a = torch.relu(a)
    out = a.permute(0, 2, 3, 1).numpy()
    assert np.abs(out - ref).max() / np.abs(ref).max() < 1e-6
