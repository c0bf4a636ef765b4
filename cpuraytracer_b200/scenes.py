"""Frozen, reproducible versions of SpheresApp::InitScene / InitCamera (spheres-app.cpp:35-130).

The reference draws its scene from std::random_device + ranlux24_base + uniform_real_distribution
(spheres-app.cpp:53-55), which is unreproducible.  Here the same *recipe* is driven by a portable
SplitMix64 stream (u = top 24 bits * 2^-24), so the scene is a pure function of (grid, seed):

  D488  : default scene, 22x22 small spheres + ground + 3 large  = 488 spheres   (seed 1)
  S4096 : scaled scene, first 4092 cells of a 64x64 grid + ground + 3 large = 4096 spheres (seed 2)
"""
from __future__ import annotations

import math

import numpy as np

from .api import (Camera, CheckerTexture, ConstTexture, DielectricOpaque, DielectricTransparent, DirectionalLight,
                  Emissive, Metal, Scene, Sphere, XMCOLOR)

# AppSettings, spheres-app.h:5-13
K_BACKBUFFER_WIDTH = 1280
K_BACKBUFFER_HEIGHT = 720
K_RECURSION_DEPTH = 50
K_VERTICAL_FOV = 25.0
K_APERTURE = 0.4
K_EXPOSURE = -15.0          # m_exposure, spheres-app.cpp:48

_f32 = np.float32
_M64 = (1 << 64) - 1


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & _M64

    def next_u64(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & _M64
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
        return z ^ (z >> 31)

    def uniform(self) -> np.float32:
        """uniformDist(generator): float in [0, 1) with 24 random bits."""
        return _f32((self.next_u64() >> 40) * (1.0 / 16777216.0))


class Ranlux24Base:
    """std::ranlux24_base = subtract_with_carry_engine<uint_fast32_t, 24, 10, 24> ([rand.predef]; its output is fixed by the
    C++ standard, so this replays the reference's generator, spheres-app.cpp:54) + MSVC's uniform_real_distribution<float>
    (0, 1): one engine draw x, u = float(x) / (float(max - min) + 1) = x * 2^-24 (VS2017 <random> _Nrand; FROM MEMORY of the
    v141 toolset the reference pins, common-lib.vcxproj:23).  Given the seed std::random_device produced in a Windows run,
    build_scene(prng=Ranlux24Base(seed)) regenerates that run's scene."""
    W_MASK, R, S = (1 << 24) - 1, 24, 10

    def __init__(self, seed: int = 19780503):
        x = seed % 2147483563 if seed else 19780503            # linear_congruential_engine<uint_least32_t, 40014, 0, 2147483563>
        if x == 0:
            x = 1
        self.x = []
        for _ in range(self.R):
            x = (x * 40014) % 2147483563
            self.x.append(x & self.W_MASK)
        self.carry = 1 if self.x[-1] == 0 else 0
        self.p = 0

    def next_u24(self) -> int:
        y = self.x[(self.p + self.R - self.S) % self.R] - self.x[self.p] - self.carry
        self.carry = 1 if y < 0 else 0
        y &= self.W_MASK
        self.x[self.p] = y
        self.p = (self.p + 1) % self.R
        return y

    def uniform(self) -> np.float32:
        return _f32(self.next_u24() * (1.0 / 16777216.0))


def default_camera(aspect: float = K_BACKBUFFER_WIDTH / K_BACKBUFFER_HEIGHT, aperture: float = K_APERTURE) -> Camera:
    """SpheresApp::InitCamera, spheres-app.cpp:35-49: focus = |origin - lookAt| (float32 XMVector3Length)."""
    o = np.array([12.0, 2.0, -2.5], _f32)
    la = np.array([0.0, 1.0, 0.0], _f32)
    d = o - la
    focus = np.sqrt(_f32(_f32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]))     # XMVector3Length: sqrt((x2+y2)+z2)
    return Camera(tuple(float(x) for x in o), tuple(float(x) for x in la), K_VERTICAL_FOV, float(_f32(aspect)),
                  float(focus), aperture)


def build_scene(grid_lo: int = -11, grid_hi: int = 11, seed: int = 1, max_small: int | None = None, prng=None) -> Scene:
    """SpheresApp::InitScene, spheres-app.cpp:51-130, with a portable PRNG.  Draw order per cell (a outer,
    b inner): chooseMat, cx, cz, then the material's own draws, left to right as written in the source.
    NB the colour draws `u() * u()` (spheres-app.cpp:74) are unsequenced operands in C++; MSVC x64 evaluates them left to
    right, which is the order used here."""
    rng = prng if prng is not None else SplitMix64(seed)
    u = rng.uniform
    scene = Scene()
    # Floor, :60-61
    floor = CheckerTexture(XMCOLOR(0.9, 0.9, 0.9, 1.0), XMCOLOR(0.2, 0.3, 0.1, 1.0), 2500.0)
    scene.add(Sphere((0.0, -1000.0, 0.0), 1000.0, DielectricOpaque(floor, 16.0)))
    # Random small spheres, :64-105
    count = 0
    for a in range(grid_lo, grid_hi):
        for b in range(grid_lo, grid_hi):
            if max_small is not None and count >= max_small:
                break
            choose = u()
            cx = _f32(a) + _f32(0.9) * u()
            cz = _f32(b) + _f32(0.9) * u()
            center = (float(cx), float(_f32(0.2)), float(cz))
            if choose < _f32(0.8):
                col = XMCOLOR(float(u() * u()), float(u() * u()), float(u() * u()), 1.0)
                smooth = _f32(8.0) * (_f32(4.0) + u())
                scene.add(Sphere(center, float(_f32(0.2)), DielectricOpaque(ConstTexture(col), float(smooth))))
            elif choose < _f32(0.95):
                col = XMCOLOR(float(_f32(0.5) * (_f32(1.0) + u())), float(_f32(0.5) * (_f32(1.0) + u())),
                              float(_f32(0.5) * (_f32(1.0) + u())), 1.0)
                scene.add(Sphere(center, float(_f32(0.2)), Metal(ConstTexture(col), 0.0)))
            else:
                smooth = _f32(8.0) * (_f32(4.0) + u())
                scene.add(Sphere(center, float(_f32(0.2)), DielectricTransparent(float(smooth), 1.5)))
            count += 1
    # Large spheres, :108-114
    scene.add(Sphere((0.0, 1.0, 0.0), 1.0, DielectricTransparent(16.0, 1.5)))
    scene.add(Sphere((-4.0, 1.0, 0.0), 1.0, DielectricOpaque(ConstTexture(XMCOLOR(0.4, 0.2, 0.1, 1.0)), 16.0)))
    scene.add(Sphere((4.0, 1.0, 0.0), 1.0, Metal(ConstTexture(XMCOLOR(0.7, 0.6, 0.5, 1.0)), 0.0)))
    # Sky :120-121, Sun :129
    scene.set_sky(Emissive(8000.0, ConstTexture(XMCOLOR(0.85, 0.91, 0.98, 1.0))))
    scene.add_light(DirectionalLight((1.0, 1.0, 1.0), XMCOLOR(1.0, 0.97, 0.88, 1.0), 40000.0))
    return scene


def scene_d488() -> Scene:
    """Default `src/spheres` scene: 1 ground + 484 small + 3 large = 488 spheres."""
    return build_scene(-11, 11, seed=1)


def scene_s4096() -> Scene:
    """Scaled synthetic scene (BASELINE config 5): 64x64 grid truncated to 4092 small spheres -> N = 4096."""
    return build_scene(-32, 32, seed=2, max_small=4092)


def scene_small(n_side: int = 3, seed: int = 3) -> Scene:
    """Tiny scene for fast parity tests: (2*n_side)^2 small spheres + ground + 3 large."""
    return build_scene(-n_side, n_side, seed=seed)


def scene_ranlux(seed: int) -> Scene:
    """The reference's own default scene for a given std::random_device seed (see Ranlux24Base)."""
    return build_scene(-11, 11, prng=Ranlux24Base(seed))


NAMED = {"d488": scene_d488, "s4096": scene_s4096, "small": scene_small}
