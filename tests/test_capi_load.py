"""CPU-side checks of the boundary: the library loads, exports every symbol include/pcp/pcp.h declares, refuses to run
without a GPU (no CPU fallback) and its host pose algebra equals the oracle's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "pcp", "pcp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pcp_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(built_lib):
    from pointcloudprocessor_b200 import _capi
    L = C.CDLL(built_lib)
    decl = _declared_symbols()
    assert len(decl) >= 25
    for s in decl:
        assert hasattr(L, s), f"{s} declared in pcp.h but not exported"
    assert sorted(_capi.SYMBOLS) == decl
    assert L.pcp_abi_version() == 1


def test_default_params_match_reference(built_lib):
    from pointcloudprocessor_b200 import api
    p = api.default_params()
    assert (p.cull_width, p.cull_height) == (4096, 3000)      # src/PointCloudProcessor.cpp:525
    assert p.downsample == 14 and p.depth_tolerance == 0.05   # view_culling.cpp:63, :157
    assert p.top_m == 5 and p.enable_depth_buffer_culling == 1


def test_cloud_slices_partition_the_map(built_lib):
    """pcp_cloud_slice (host arithmetic of pcp_set_cloud_sharded): the ranks' slices are disjoint, in rank order, cover [0, n), have the
    same length except for the last non-empty one, and bad arguments are refused."""
    from pointcloudprocessor_b200 import api
    for n in (0, 1, 7, 64, 60_001, 100_000_000, 4_294_967_279):
        for world in (1, 2, 3, 4, 8, 64):
            per = -(-n // world)
            nxt = 0
            for r in range(world):
                first, count = api.cloud_slice(n, r, world)
                assert first == min(r * per, n) == nxt and count == min(per, n - first)
                nxt = first + count
            assert nxt == n
    for bad in ((10, -1, 4), (10, 4, 4), (10, 0, 0)):
        with pytest.raises(api.PcpError):
            api.cloud_slice(*bad)


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from pointcloudprocessor_b200 import api
    cam = api.create_camera("pinhole", (100, 100, 50, 50), (0, 0, 0, 0, 0))
    with pytest.raises(api.PcpError, match="no CUDA device"):
        api.Colorizer(cam, api.default_params(image_size=(64, 48)))


def test_create_camera_errors(built_lib):
    from pointcloudprocessor_b200 import api
    with pytest.raises(api.PcpError):
        api.create_camera("fisheye", (1, 1, 0, 0), ())
    with pytest.raises(api.PcpError):
        api.create_camera("pinhole", (1, 1, 0), ())
    cam = api.create_camera("pinhole", (1, 2, 3, 4), (0.1, 0.2))
    assert list(cam.dist) == [0.1, 0.2, 0.0, 0.0, 0.0]


def test_pose_matrices_equal_oracle(built_lib, oracle):
    from pointcloudprocessor_b200 import api, synth
    poses, _ = synth.make_poses(50)
    rng = np.random.default_rng(3)
    for i, p in enumerate(poses):
        q = p.copy()
        if i % 3 == 0:
            q[3:] *= 1.0 + 1e-3 * rng.normal()  # Eigen does not normalise the quaternion
        w_g, c_g = api.pose_matrices(q)
        w_o, c_o = oracle.pose_matrices(q)
        assert np.array_equal(w_g, w_o) and np.array_equal(c_g, c_o)
        T = np.eye(4)
        T[:3, 3] = rng.normal(size=3) * 0.05
        a = rng.normal() * 0.01
        T[:3, :3] = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        w_g, c_g = api.pose_matrices(q, T)
        w_o, c_o = oracle.pose_matrices(q, T)
        assert np.array_equal(w_g, w_o) and np.array_equal(c_g, c_o)
        # sanity: w2c is the inverse of c2w
        assert np.abs(w_g.astype(np.float64) @ c_g.astype(np.float64) - np.eye(4)).max() < 1e-4


def _build_facade(tmp_path, built_lib):
    import subprocess
    exe = str(tmp_path / "facade_smoke")
    libdir = os.path.dirname(built_lib)
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "facade_smoke.cpp"),
                           "-o", exe, "-L" + libdir, "-lpcp_b200", "-Wl,-rpath," + libdir])
    return exe


def test_cpp_facade_compiles_and_fails_loudly_without_gpu(tmp_path, built_lib):
    """include/pcp/pcp.hpp (reference class names over the C ABI) compiles with -Wall -Werror, links, and without a CUDA
    device rethrows the library's error (no CPU fallback)."""
    import subprocess
    import torch
    exe = _build_facade(tmp_path, built_lib)
    r = subprocess.run([exe], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stdout + r.stderr
        assert "facade ok" in r.stdout
    else:
        assert r.returncode == 3 and "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_cpp_facade_on_gpu(tmp_path, built_lib):
    import subprocess
    r = subprocess.run([_build_facade(tmp_path, built_lib)], capture_output=True, text=True)
    assert r.returncode == 0 and "facade ok" in r.stdout, r.stdout + r.stderr
