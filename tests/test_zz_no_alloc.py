"""pcp_compact_bits_noalloc (no counterpart in the reference): the bitmask -> ascending
index list compaction with its scratch in a file-scope static __device__ array, so that the call allocates nothing.  Checked through
the C ABI against the existing entry point that does the same compaction with a context-owned buffer (pcp_cull_pose), exactly
(integer indices), on two calls in a row (the static scratch keeps the first call's contents), plus the size limit.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_scene

pytestmark = pytest.mark.gpu


def _noalloc(gc, d_bits, n_bits, d_idx, d_cnt):
    from pointcloudprocessor_b200 import _capi
    rc = gc._L.pcp_compact_bits_noalloc(gc.handle, C.c_void_p(d_bits.data_ptr()), n_bits, C.c_void_p(d_idx.data_ptr()), C.c_void_p(d_cnt.data_ptr()))
    assert rc == _capi.PCP_OK, gc._L.pcp_last_error(gc.handle)
    gc.synchronize()


def test_compact_bits_noalloc_equals_cull(built_lib):
    import torch
    from pointcloudprocessor_b200 import _capi, api
    W, H, n = 640, 480, 150_001  # not a multiple of 32: the last mask word is ragged; 5 tiles of 1024 words
    sc = make_scene(n, 3, W, H)
    cam = api.create_camera("pinhole", sc["intr"], sc["dist"])
    with api.Colorizer(cam, api.default_params(image_size=(W, H)), device=0) as gc:
        gc.set_cloud(sc["cloud"])
        d_idx = torch.full((n,), -7, dtype=torch.int32, device="cuda")
        d_cnt = torch.full((1,), 123456, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for k in (1, 2):  # two calls in a row, on different frames: the second call finds the first call's scratch contents
            want = gc.cull_pose(sc["poses"][k])                       # existing entry point: ascending indices of the visible points
            r = gc.submit_frame(sc["poses"][k], sc["images"][k], want_vis=True)
            assert np.array_equal(np.flatnonzero(r["vis"]).astype(np.int32), want) and len(want) > 500
            d_bits = torch.from_numpy(r["vis_bits"].view(np.int32)).cuda()
            torch.cuda.synchronize()
            _noalloc(gc, d_bits, n, d_idx, d_cnt)
            m = int(d_cnt.cpu()[0])
            assert m == len(want)
            assert np.array_equal(d_idx[:m].cpu().numpy(), want)      # integer indices: exact
        # all bits clear / no bits at all: the count word is still written
        d_zero = torch.zeros(((n + 31) // 32,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        _noalloc(gc, d_zero, n, d_idx, d_cnt)
        assert int(d_cnt.cpu()[0]) == 0
        d_cnt.fill_(99)
        torch.cuda.synchronize()
        _noalloc(gc, d_zero, 0, d_idx, d_cnt)
        assert int(d_cnt.cpu()[0]) == 0
        # beyond the static scratch: rejected with a clear error, nothing launched
        rc = gc._L.pcp_compact_bits_noalloc(gc.handle, C.c_void_p(d_zero.data_ptr()), (1 << 27) + 1, C.c_void_p(d_idx.data_ptr()), C.c_void_p(d_cnt.data_ptr()))
        assert rc == _capi.PCP_EINVAL and b"PCP_NOALLOC_MAX_BITS" in gc._L.pcp_last_error(gc.handle)
