"""The allocation-free std entry point, `stereo_volume_std_static`.

Some hosts of this library may not allocate device memory once they are running
(a fixed start-up memory map, or a pool that already owns the whole GPU), so this
one call keeps its partial sums in an array built into the library.  The test
drives it through the C ABI with caller-owned device buffers and checks it against
the regular `stereo_volume_std` - twice in a row, the second time on a smaller
input, so that anything the first call left in the shared array would show up.
Collected last on purpose (see the file name).
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_static_scratch_std_matches_regular_entry_point_two_calls_in_a_row():
    import torch
    from stereo_matching_b200 import _lib

    lib = _lib.load()
    ctx = _lib.Context(0)
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(11)
    for shape in [(61, 257, 201), (9, 31, 33)]:
        vol = (rng.standard_normal(shape) * 3.0 + 1.5).astype(np.float32)
        expected = ctx.volume_std(vol)                       # the existing entry point
        d_vol = torch.from_numpy(vol).cuda()
        d_out = torch.full((2,), -1.0, dtype=torch.float32, device="cuda")
        _lib.check(lib.stereo_volume_std_static(d_vol.data_ptr(), vol.size, d_out.data_ptr(), stream))
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        # same tolerance as tests/test_gpu_post.py::test_volume_std: 1 ulp of float32
        np.testing.assert_array_max_ulp(np.float32(got[0]), np.float32(expected), maxulp=1)
        assert abs(got[1] - vol.astype(np.float64).mean()) < 1e-5


def test_static_scratch_std_refuses_oversized_and_bad_input():
    import torch
    from stereo_matching_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.float32, device="cuda")
    too_many = 4096 * 65536 + 1
    assert lib.stereo_volume_std_static(d.data_ptr(), too_many, out.data_ptr(), None) == -5
    assert b"at most" in lib.stereo_last_error()
    assert lib.stereo_volume_std_static(d.data_ptr(), 1, out.data_ptr(), None) == -1
    assert lib.stereo_volume_std_static(None, 16, out.data_ptr(), None) == -1
