"""Known-answer tests that pin the CPU oracle to the reference source.

The reference ships no tests or golden vectors; every expected value below is derived by hand from the
cited reference lines (paths relative to the reference tree, src/...).
"""
import ctypes as C
import math

import numpy as np
import pytest

from conftest import orc_camera


def _f(n):
    return (C.c_float * n)()


def f32(x):
    return float(np.float32(x))


# ---- Random::HaltonSample, quasi-random.cpp:3-16 ----
def test_halton_known_values(orc):
    L = orc.load()
    for i, e in zip((1, 2, 3, 4), (0.5, 0.25, 0.75, 0.125)):
        assert L.orc_halton(i, 2) == e
    for i, e in zip((1, 2, 3, 5), (1 / 3, 2 / 3, 1 / 9, 7 / 9)):
        assert abs(L.orc_halton(i, 3) - e) < 1e-7
    for b in (2, 3, 4, 5, 7):
        assert L.orc_halton(0, b) == 0.0
    # base 4 is used although not prime (spheres-app.cpp:152): H4(5) = 1/4 + 1/16
    assert L.orc_halton(5, 4) == 0.25 + 1 / 16


def test_halton_matches_float_loop(orc):
    """Bit-exact against a numpy float32 replay of the loop at quasi-random.cpp:8-13, large indices included."""
    L = orc.load()
    rng = np.random.default_rng(0)
    idx = np.concatenate([np.arange(0, 200), rng.integers(1, 1 << 40, 200)])
    for base in (2, 3, 4, 5, 7):
        for i in idx:
            i = int(i)
            r, f, k = np.float32(0), np.float32(1), i
            while k > 0:
                f = np.float32(f / np.float32(base))
                r = np.float32(r + np.float32(f * np.float32(k % base)))
                k //= base
            assert L.orc_halton(i, base) == float(r)


def test_halton_disk_radius_is_linear_not_sqrt(orc):
    """quasi-random.cpp:52-61: theta = 2 pi H_b1, r = H_b2 (centre-weighted lens)."""
    L = orc.load()
    out = _f(2)
    for k in (1, 2, 7, 33, 1000):
        L.orc_halton_disk(k, 4, 5, out)
        r = math.hypot(out[0], out[1])
        assert abs(r - L.orc_halton(k, 5)) < 2e-7
        th = 2 * math.pi * L.orc_halton(k, 4)
        assert abs(out[0] - r * math.cos(th)) < 1e-6 and abs(out[1] - r * math.sin(th)) < 1e-6


def test_hemisphere_is_uniform_solid_angle(orc):
    """quasi-random.cpp:36-50: z = u1, r = sqrt(1 - u1^2): NOT cosine weighted."""
    L = orc.load()
    out = _f(3)
    for u1, u2 in ((0.0, 0.0), (0.25, 0.5), (0.9, 0.125), (1.0, 0.7)):
        L.orc_hemisphere_from_u(u1, u2, out)
        assert out[2] == f32(u1)
        assert abs(out[0] ** 2 + out[1] ** 2 + out[2] ** 2 - 1.0) < 1e-6
        assert abs(out[0] - math.sqrt(1 - u1 * u1) * math.cos(2 * math.pi * u2)) < 1e-6
    L.orc_halton_hemisphere(7, 5, 7, out)
    assert out[2] == L.orc_halton(7, 5)


def test_pinned_sincos_accuracy(orc):
    L = orc.load()
    s, c = C.c_float(), C.c_float()
    th = np.linspace(0, 2 * math.pi, 20001, dtype=np.float32)
    err = 0.0
    for t in th:
        L.orc_sincos(float(t), C.byref(s), C.byref(c))
        err = max(err, abs(s.value - math.sin(float(t))), abs(c.value - math.cos(float(t))))
    assert err < 3e-7


# ---- Sphere::Intersect, ray-tracing.cpp:42-84 ----
def _sphere(orc, c, r):
    a = np.zeros(1, orc.SPHERE_DTYPE)
    a[0] = (c[0], c[1], c[2], r, 0)
    return a


def _hit(orc, sph, o, d):
    L = orc.load()
    t, pos, n, uv = C.c_float(), _f(3), _f(3), _f(2)
    ok = L.orc_sphere_intersect(sph.ctypes.data, orc.f3(*o), orc.f3(*d), C.byref(t), pos, n, uv)
    return ok, t.value, list(pos), list(n), list(uv)


def test_sphere_front_hit(orc):
    ok, t, pos, n, uv = _hit(orc, _sphere(orc, (0, 0, 0), 1.0), (0, 0, -3), (0, 0, 1))
    assert ok and t == 2.0 and pos == [0.0, 0.0, -1.0] and n == [0.0, 0.0, -1.0]
    assert uv == [0.5, 0.0]          # ComputeUV uses x and z (ray-tracing.cpp:36-37)


def test_sphere_inside_uses_far_root_and_outward_normal(orc):
    ok, t, pos, n, _ = _hit(orc, _sphere(orc, (0, 0, 0), 1.0), (0, 0, 0), (0, 0, 1))
    assert ok and t == 1.0 and n == [0.0, 0.0, 1.0]      # never flipped (:62)


def test_sphere_tangent_is_miss_and_bias(orc):
    sph = _sphere(orc, (0, 0, 0), 1.0)
    assert not _hit(orc, sph, (1, 0, -3), (0, 0, 1))[0]                  # disc == 0, strict '>' (:54)
    assert not _hit(orc, sph, (0, 0, -3), (0, 0, -1))[0]                 # both roots negative
    # origin 0.0005 in front of the far surface: near root < 0, far root 0.0005 <= 0.001 bias -> rejected (:58,:71)
    assert not _hit(orc, sph, (0, 0, 0.9995), (0, 0, 1))[0]
    ok, t, *_ = _hit(orc, sph, (0, 0, 0.99), (0, 0, 1))
    assert ok and abs(t - 0.01) < 1e-6


def test_sphere_unnormalised_direction_divides_by_a(orc):
    ok, t, *_ = _hit(orc, _sphere(orc, (0, 0, 0), 1.0), (0, 0, -3), (0, 0, 2))
    assert ok and t == 1.0               # (-b - sqrt(disc)) / a with a = 4 (:56)


# ---- Camera, camera.cpp:3-48 ----
def test_camera_centre_ray_and_basis(orc, rt):
    L = orc.load()
    cam = rt.scenes.default_camera()
    cp = orc_camera(orc, cam)
    oc = orc.OrcCamera()
    L.orc_camera_make(C.byref(cp), C.byref(oc))
    hh = math.tan(math.radians(25.0) / 2)
    assert abs(math.sqrt(sum(v * v for v in oc.y)) - hh) < 1e-6               # m_y = halfHeight * v (:27)
    assert abs(math.sqrt(sum(v * v for v in oc.x)) - hh * 1280 / 720) < 1e-6  # m_x = halfWidth * u (:26)
    o, d = _f(3), _f(3)
    L.orc_camera_get_ray(C.byref(oc), 0.5, 0.5, 0.0, 0.0, o, d)
    w = np.array(cam.look_at) - np.array(cam.origin)
    w /= np.linalg.norm(w)
    assert np.allclose(list(d), w, atol=1e-6) and list(o) == [12.0, 2.0, -2.5]
    # left-handed: +u on screen (ndc.x > 0) is +cross(up, w)
    L.orc_camera_get_ray(C.byref(oc), 1.0, 0.5, 0.0, 0.0, o, d)
    u = np.cross([0, 1, 0], w)
    assert np.dot(list(d), u) > 0
    # v = 0 is the TOP row: ndc.y = 1 - 2v (:35)
    L.orc_camera_get_ray(C.byref(oc), 0.5, 0.0, 0.0, 0.0, o, d)
    assert d[1] > w[1]


def test_camera_focal_plane_invariance(orc, rt):
    """Every lens offset passes through o + focus * normalize(p - o) (camera.cpp:39,47)."""
    L = orc.load()
    cam = rt.scenes.default_camera()
    cp = orc_camera(orc, cam)
    oc = orc.OrcCamera()
    L.orc_camera_make(C.byref(cp), C.byref(oc))
    o0, d0, o, d = _f(3), _f(3), _f(3), _f(3)
    for (u, v) in ((0.5, 0.5), (0.1, 0.9), (0.99, 0.01)):
        L.orc_camera_get_ray(C.byref(oc), u, v, 0.0, 0.0, o0, d0)
        focal = np.array(list(o0)) + cam.focal_length * np.array(list(d0))
        for (lx, ly) in ((1, 0), (0, 1), (-0.3, 0.7)):
            L.orc_camera_get_ray(C.byref(oc), u, v, lx, ly, o, d)
            oo, dd = np.array(list(o)), np.array(list(d))
            s = np.dot(focal - oo, dd)
            assert np.linalg.norm(oo + s * dd - focal) < 1e-4
            # lens offsets scale by m_x, m_y: the lens is an ellipse (:43-45)
            exp = np.array(cam.origin) + 0.5 * cam.aperture * (lx * np.array(list(oc.x)) + ly * np.array(list(oc.y)))
            assert np.allclose(oo, exp, atol=1e-6)


# ---- materials ----
def test_schlick_and_fresnel(orc):
    L = orc.load()
    assert L.orc_schlick(0.04, 1.0) == f32(0.04)                  # (1 - n.v)^5 = 0
    assert abs(L.orc_schlick(0.04, 0.0) - 1.0) < 1e-7
    assert abs(L.orc_schlick(0.04, 0.5) - (0.04 + 0.96 * 0.5 ** 5)) < 1e-7
    assert abs(L.orc_fresnel_term(1.0, 1.5) - 0.04) < 1e-6        # XMFresnelTerm self-checks
    assert abs(L.orc_fresnel_term(0.0, 1.5) - 1.0) < 1e-6
    # exact unpolarised Fresnel for a dielectric, cos = 0.5, n = 1.5
    c, n = 0.5, 1.5
    g = math.sqrt(n * n + c * c - 1)
    ref = 0.5 * (g - c) ** 2 / (g + c) ** 2 * (((c * (g + c) - 1) / (c * (g - c) + 1)) ** 2 + 1)
    assert abs(L.orc_fresnel_term(0.5, 1.5) - ref) < 1e-6


def test_reflect_refract(orc):
    L = orc.load()
    out = _f(3)
    L.orc_reflect(orc.f3(1, -1, 0), orc.f3(0, 1, 0), out)
    assert list(out) == [1.0, 1.0, 0.0]
    i = np.array([math.sin(0.5), -math.cos(0.5), 0.0])
    L.orc_refract(orc.f3(*i), orc.f3(0, 1, 0), 1 / 1.5, out)      # Snell: sin(t) = sin(i) / 1.5
    assert abs(out[0] - math.sin(0.5) / 1.5) < 1e-6 and out[1] < 0
    assert abs(sum(v * v for v in out) - 1) < 1e-6
    i = np.array([math.sin(1.2), -math.cos(1.2), 0.0])
    L.orc_refract(orc.f3(*i), orc.f3(0, 1, 0), 1.5, out)          # total internal reflection -> zero vector
    assert list(out) == [0.0, 0.0, 0.0]


def test_xmcolor_round_trip(orc):
    L = orc.load()
    assert L.orc_xmcolor_quantize(0.0) == 0 and L.orc_xmcolor_quantize(1.0) == 255
    assert L.orc_xmcolor_quantize(-3.0) == 0 and L.orc_xmcolor_quantize(7.0) == 255
    assert L.orc_xmcolor_quantize(0.85) == 217 and L.orc_xmcolor_quantize(0.91) == 232 and L.orc_xmcolor_quantize(0.98) == 250
    assert L.orc_xmcolor_quantize(0.5 / 255) == 0 and L.orc_xmcolor_quantize(1.5 / 255) == 2     # ties to even (cvtps2dq)
    assert L.orc_xmcolor_load(255) == f32(np.float32(255) * np.float32(1 / 255))


def test_checker_texture(orc):
    """texture.cpp:20-33: int(tiling*u) % 2 == int(tiling*v) % 2 ? c0 : c1."""
    L = orc.load()
    t = np.zeros(1, orc.TEXTURE_DTYPE)
    t[0] = (1, 2500.0, (1, 0, 0, 1), (0, 1, 0, 1))
    out = _f(4)
    L.orc_texture_eval(t.ctypes.data, 0.0001, 0.0001, out)
    assert list(out)[:3] == [1.0, 0.0, 0.0]
    L.orc_texture_eval(t.ctypes.data, 0.0005, 0.0001, out)
    assert list(out)[:3] == [0.0, 1.0, 0.0]
    L.orc_texture_eval(t.ctypes.data, 0.0005, 0.0005, out)
    assert list(out)[:3] == [1.0, 0.0, 0.0]


def _one_sphere_scene(orc, kind, color=(0.5, 0.5, 0.5, 1.0), smoothness=16.0, light=True):
    sp = np.zeros(1, orc.SPHERE_DTYPE)
    sp[0] = (0, 0, 0, 1.0, 0)
    ma = np.zeros(1, orc.MATERIAL_DTYPE)
    ma[0] = (kind, 0 if kind != orc.MAT_TRANSPARENT else -1, smoothness, 1.5 if kind == orc.MAT_TRANSPARENT else 0.0)
    te = np.zeros(2, orc.TEXTURE_DTYPE)
    te[0] = (0, 0.0, color, color)
    te[1] = (0, 0.0, (0.5, 0.5, 0.5, 1), (0.5, 0.5, 0.5, 1))
    li = np.zeros(1 if light else 0, orc.LIGHT_DTYPE)
    if light:
        li[0] = ((1, 1, 1), (1, 1, 1, 1), 100.0)
    return orc.Scene(sp, ma, te, li, 10.0, 1)


def test_metal_always_reflects(orc):
    """material.cpp:81-85: the 4-lane compare includes alpha = 1, so any xi reflects."""
    L = orc.load()
    sc = _one_sphere_scene(orc, orc.MAT_METAL, (0.7, 0.6, 0.5, 1.0), 0.0)
    at, oo, od = _f(3), _f(3), _f(3)
    for xi in (0.0, 0.5, 0.999999):
        r = L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, -3), orc.f3(0, 0, 1), xi, 0.3, 0.6, at, oo, od)
        assert r == 1 and list(od) == [0.0, 0.0, -1.0]
        assert np.allclose(list(at), (0.7, 0.6, 0.5))


def test_opaque_fresnel_coin_and_backface(orc):
    L = orc.load()
    sc = _one_sphere_scene(orc, orc.MAT_OPAQUE, (0.2, 0.4, 0.6, 1.0))
    at, oo, od = _f(3), _f(3), _f(3)
    # head-on: F = 0.04; xi below -> mirror with attenuation 1 (material.cpp:29-37)
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, -3), orc.f3(0, 0, 1), 0.03, 0.3, 0.6, at, oo, od) == 1
    assert list(at) == [1.0, 1.0, 1.0] and list(od) == [0.0, 0.0, -1.0]
    # xi above -> diffuse with albedo; direction in the un-normalised basis (material.cpp:50-56)
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, -3), orc.f3(0, 0, 1), 0.5, 0.3, 0.6, at, oo, od) == 1
    assert np.allclose(list(at), (0.2, 0.4, 0.6))
    n = np.array([0, 0, -1.0])
    b1 = np.cross([1, 0, 0], n)
    b2 = np.cross(n, b1)
    r = math.sqrt(1 - 0.3 ** 2)
    d = r * math.cos(2 * math.pi * 0.6) * b1 + r * math.sin(2 * math.pi * 0.6) * b2 + 0.3 * n
    assert np.allclose(list(od), d / np.linalg.norm(d), atol=1e-6)
    # from inside: -d.n <= 0 -> no scatter (material.cpp:22)
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, 0), orc.f3(0, 0, 1), 0.5, 0.3, 0.6, at, oo, od) == 0


def test_glass_refracts_and_reflects(orc):
    L = orc.load()
    sc = _one_sphere_scene(orc, orc.MAT_TRANSPARENT)
    at, oo, od = _f(3), _f(3), _f(3)
    # head-on entering: P = 0.04; xi = 0.5 -> straight through
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, -3), orc.f3(0, 0, 1), 0.5, 0, 0, at, oo, od) == 1
    assert np.allclose(list(od), (0, 0, 1), atol=1e-6) and list(at) == [1.0, 1.0, 1.0]
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(0, 0, -3), orc.f3(0, 0, 1), 0.01, 0, 0, at, oo, od) == 1
    assert np.allclose(list(od), (0, 0, -1), atol=1e-6)
    # exiting at a grazing angle: total internal reflection regardless of xi (material.cpp:146-149)
    o = np.array([0.95, 0, 0.0])
    assert L.orc_scatter(C.byref(sc.c), 0, orc.f3(*o), orc.f3(0, 0, 1), 0.999, 0, 0, at, oo, od) == 1
    assert od[2] < 0.9 and od[0] < 0


def test_direct_light_lambert_blinn_phong(orc):
    """light.cpp:11-42 at the sub-light point of one opaque sphere, viewed along the light direction."""
    L = orc.load()
    sc = _one_sphere_scene(orc, orc.MAT_OPAQUE, (0.2, 0.4, 0.6, 1.0), smoothness=16.0)
    s = 1 / math.sqrt(3)
    view = (5 * s, 5 * s, 5 * s)
    rgb = _f(3)
    idx = L.orc_shade_hit(C.byref(sc.c), orc.f3(*view), orc.f3(*view), orc.f3(-s, -s, -s), orc.ACCEL_BRUTE, rgb)
    assert idx == 0
    # n = L = V = H: Lin = 100, F = 0.04, spec = 0.04 * (16 + 8) / 8 = 0.12
    exp = [100.0 * (a + 0.12) for a in (0.2, 0.4, 0.6)]
    assert np.allclose(list(rgb), exp, rtol=2e-5)
    # the far side is self-shadowed (the shadow ray re-enters the sphere)
    idx = L.orc_shade_hit(C.byref(sc.c), orc.f3(-5 * s, -5 * s, -5 * s), orc.f3(-5 * s, -5 * s, -5 * s), orc.f3(s, s, s), 0, rgb)
    assert idx == 0 and list(rgb) == [0.0, 0.0, 0.0]


# ---- whole path ----
def test_empty_scene_is_sky_and_resolve(orc, rt):
    sc = _one_sphere_scene(orc, orc.MAT_OPAQUE)
    sc.c.n_spheres = 0
    cp = orc_camera(orc, rt.scenes.default_camera())
    hdr, st = orc.render(sc, cp, orc.config(16, 9, spp=3, threads=1))
    exp = np.float32(10.0 * 0.5) * np.float32(2.0 ** -15)
    assert np.all(hdr[..., :3] == np.float32(3) * exp) and np.all(hdr[..., 3] == 3)
    assert st.paths == 16 * 9 * 3 and st.closest_rays == st.paths and st.shadow_rays == 0
    assert np.all(orc.resolve_ldr(np.zeros((2, 2, 4), np.float32), 1) == 0xFF000000)
    # sky-only pixel of the default scene: 8000 * 2^-15 * {217,232,250}/255, ACES + gamma (spheres-app.cpp:186-214)
    px = np.zeros((1, 1, 4), np.float32)
    px[0, 0, :3] = 8000 * 2.0 ** -15 * np.array([217, 232, 250]) / 255
    v = int(orc.resolve_ldr(px, 1)[0, 0])

    def aces(x):
        y = (x * (2.51 * x + 0.03)) / (x * (2.43 * x + 0.59) + 0.14)
        return round(255 * min(max(y, 0), 1) ** (1 / 2.2))
    assert [(v >> 16) & 255, (v >> 8) & 255, v & 255] == [aces(float(c)) for c in px[0, 0, :3]] and (v >> 24) == 255


def test_recursive_equals_iterative_and_bvh_equals_brute(orc, rt, scene_factory):
    _, sc = scene_factory("small")
    cp = orc_camera(orc, rt.scenes.default_camera())
    base, _ = orc.render(sc, cp, orc.config(96, 54, spp=2, max_depth=50))
    rec, _ = orc.render(sc, cp, orc.config(96, 54, spp=2, max_depth=50, integrator=orc.INTEGRATOR_RECURSIVE))
    bvh, _ = orc.render(sc, cp, orc.config(96, 54, spp=2, max_depth=50, accel=orc.ACCEL_BVH))
    assert np.array_equal(base, bvh)                       # same hits => same bytes
    assert np.allclose(base, rec, rtol=2e-5, atol=1e-7)    # (E+S) + a*(...) vs L += T*S: association differs


def test_halton_mode_single_thread_is_deterministic(orc, rt, scene_factory):
    _, sc = scene_factory("small")
    cp = orc_camera(orc, rt.scenes.default_camera())
    cfg = dict(spp=2, max_depth=50, rng_mode=orc.RNG_HALTON, accel=orc.ACCEL_BVH, integrator=orc.INTEGRATOR_RECURSIVE, threads=1)
    a, _ = orc.render(sc, cp, orc.config(64, 36, **cfg))
    b, _ = orc.render(sc, cp, orc.config(64, 36, **cfg))
    assert np.array_equal(a, b)


def test_depth0_is_direct_only(orc, rt, scene_factory):
    """max_depth = 0 == k_recursionDepth 0: one vertex, glass/metal show only their specular lobe."""
    _, sc = scene_factory("small")
    cp = orc_camera(orc, rt.scenes.default_camera())
    _, st = orc.render(sc, cp, orc.config(64, 36, spp=1, max_depth=0))
    assert st.closest_rays == st.paths and st.shadow_rays <= st.paths
