"""Shared fixtures.  `-m "not gpu"` runs here on CPU; `-m gpu` runs on a B200 through the C ABI."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


@pytest.fixture(scope="session")
def rt():
    """The product package (builds csrc/librt_b200.so with nvcc if it is stale)."""
    from cpuraytracer_b200 import build
    build.build()
    import cpuraytracer_b200
    return cpuraytracer_b200


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    from oracle import orc as o
    o.build()
    o.load()
    return o


def orc_camera(orc, cam):
    return orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)


@pytest.fixture(scope="session")
def scene_factory(rt, orc):
    """name -> (rt Scene, oracle Scene), cached per session."""
    cache = {}

    def get(name):
        if name not in cache:
            s = rt.scenes.NAMED[name]()
            cache[name] = (s, orc.Scene.from_flat(s.export()))
        return cache[name]
    return get


def rel_rmse(g, o):
    """sqrt(mean(((g - o) / (o + eps))^2)), eps = 1e-3 * mean(o); on linear mean radiance."""
    g = np.asarray(g, np.float64)
    o = np.asarray(o, np.float64)
    eps = 1e-3 * o.mean()
    return float(np.sqrt(np.mean(((g - o) / (o + eps)) ** 2)))


def mean_rel_err(g, o):
    return float(abs(np.mean(g, dtype=np.float64) - np.mean(o, dtype=np.float64)) / np.mean(o, dtype=np.float64))
