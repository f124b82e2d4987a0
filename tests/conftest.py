import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built_lib():
    from pointcloudprocessor_b200 import build
    return build.build_library()


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


def make_scene(n, f, W, H, seed=42, masks=False):
    from pointcloudprocessor_b200 import synth
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n, seed)
    poses, ts = synth.make_poses(f)
    images = [synth.make_image(i, W, H) for i in range(f)]
    mks = [synth.make_mask(i, W, H) for i in range(f)] if masks else [None] * f
    return dict(intr=intr, dist=dist, cloud=cloud, poses=poses, images=images, masks=mks, W=W, H=H)
