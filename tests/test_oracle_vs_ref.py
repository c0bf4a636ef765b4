"""Pins the oracle restatement (oracle/rt_oracle.cpp) against oracle/_ref: the reference's OWN camera / light / material /
quasi-random / ray-tracing / texture .cpp files, compiled unmodified from /root/reference against the DirectXMath
stand-in (oracle/dxm_shim) by oracle/Makefile.  With ORC_MATH_REFERENCE (powf, libm sin/cos, rcpps -- exactly the calls
the reference makes) the two must agree BIT FOR BIT, function by function and for whole depth-50 renders.

Skipped when oracle/_ref/libref.so is absent (it needs /root/reference to build; the prebuilt file travels to the GPU box).
The committed fixtures in tests/golden/ (made by tests/golden/make_golden.py from this library) keep the pin without it.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import orc_camera


@pytest.fixture(scope="module")
def ref(orc):
    lib = orc.load_ref()
    if lib is None:
        pytest.skip("oracle/_ref/libref.so not built (needs /root/reference)")
    orc.load().orc_set_math_mode(orc.MATH_REFERENCE)
    yield lib
    orc.load().orc_set_math_mode(orc.MATH_PINNED)


def _f(n):
    return (C.c_float * n)()


def _bits(x):
    return np.asarray(list(x), np.float32).view(np.uint32).tolist()


def test_halton_and_samplers(orc, ref):
    L = orc.load()
    rng = np.random.default_rng(2)
    idx = np.concatenate([np.arange(0, 300), rng.integers(1, 1 << 48, 300)])
    a, b = _f(3), _f(3)
    for i in idx:
        i = int(i)
        for base in (2, 3, 4, 5, 7):
            assert L.orc_halton(i, base) == ref.ref_halton(i, base)
        L.orc_halton_disk(i, 4, 5, a); ref.ref_halton_disk(i, 4, 5, b)
        assert _bits(a)[:2] == _bits(b)[:2]
        L.orc_halton_hemisphere(i, 5, 7, a); ref.ref_halton_hemisphere(i, 5, 7, b)
        assert _bits(a) == _bits(b)


def test_xmcolor_and_texture(orc, ref):
    L = orc.load()
    for k in range(256):
        assert L.orc_xmcolor_load(k) == ref.ref_xmcolor_load(k)
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.uniform(-0.2, 1.2, 2000), (np.arange(0, 256) + 0.5) / 255, [0.9, 0.3, 0.1, 0.2, 0.85, 0.91, 0.98]]).astype(np.float32)
    for x in xs:
        assert L.orc_xmcolor_quantize(float(x)) == ref.ref_xmcolor_quantize(float(x)), x
    t = np.zeros(1, orc.TEXTURE_DTYPE)
    q = lambda k: float(np.float32(k) * np.float32(1.0 / 255.0))     # flat scenes hold post-quantisation floats (XMLoadColor)  # noqa: E731
    t[0] = (1, 2500.0, (q(230), q(230), q(230), 1), (q(51), q(76), q(26), 1))
    a, b = _f(4), _f(4)
    for u, v in rng.uniform(0, 1, (500, 2)):
        L.orc_texture_eval(t.ctypes.data, u, v, a); ref.ref_texture_eval(t.ctypes.data, u, v, b)
        assert _bits(a) == _bits(b)


def test_camera_rays(orc, ref, rt):
    L = orc.load()
    rng = np.random.default_rng(4)
    for cam in (rt.scenes.default_camera(), rt.Camera((1, 2, 3), (-4, 0.5, 9), 60.0, 1.0, 5.0, 0.7)):
        cp = orc_camera(orc, cam)
        oc = orc.OrcCamera()
        L.orc_camera_make(C.byref(cp), C.byref(oc))
        o1, d1, o2, d2 = _f(3), _f(3), _f(3), _f(3)
        for u, v, lx, ly in rng.uniform(-1, 1, (400, 4)):
            L.orc_camera_get_ray(C.byref(oc), abs(u), abs(v), lx, ly, o1, d1)
            ref.ref_camera_get_ray(C.byref(cp), abs(u), abs(v), lx, ly, o2, d2)
            assert _bits(o1) == _bits(o2) and _bits(d1) == _bits(d2)


def test_sphere_intersect_random_rays(orc, ref):
    L = orc.load()
    rng = np.random.default_rng(5)
    sph = np.zeros(1, orc.SPHERE_DTYPE)
    hits = 0
    for k in range(4000):
        big = k % 4 == 0
        sph[0] = (0.0, -1000.0, 0.0, 1000.0, 0) if big else (*rng.uniform(-3, 3, 3), rng.uniform(0.1, 1.5), 0)
        o = rng.uniform(-4, 4, 3) if not big else np.array([rng.uniform(-10, 10), rng.uniform(0, 3) * (k % 8 != 0), rng.uniform(-10, 10)])
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        if k % 5 == 0:
            d *= rng.uniform(0.5, 2.0)                   # the test divides by a = d.d: un-normalised directions too
        r1 = [C.c_float(), _f(3), _f(3), _f(2)]
        r2 = [C.c_float(), _f(3), _f(3), _f(2)]
        h1 = L.orc_sphere_intersect(sph.ctypes.data, orc.f3(*o), orc.f3(*d), C.byref(r1[0]), r1[1], r1[2], r1[3])
        h2 = ref.ref_sphere_intersect(sph.ctypes.data, orc.f3(*o), orc.f3(*d), C.byref(r2[0]), r2[1], r2[2], r2[3])
        assert h1 == h2
        if h1:
            hits += 1
            assert r1[0].value == r2[0].value
            for a, b in zip(r1[1:], r2[1:]):
                assert _bits(a) == _bits(b)
    assert hits > 500


def test_fresnel_reflect_refract(orc, ref):
    L = orc.load()
    rng = np.random.default_rng(6)
    a, b = _f(3), _f(3)
    for c in np.concatenate([rng.uniform(0, 1, 2000), [0.0, 1.0]]):
        assert L.orc_fresnel_term(float(c), 1.5) == ref.ref_fresnel_term(float(c), 1.5)
    for k in range(2000):
        i, n = rng.normal(size=3), rng.normal(size=3)
        i /= np.linalg.norm(i); n /= np.linalg.norm(n)
        L.orc_reflect(orc.f3(*i), orc.f3(*n), a); ref.ref_reflect(orc.f3(*i), orc.f3(*n), b)
        assert _bits(a) == _bits(b)
        eta = 1.5 if k % 2 else float(np.float32(1) / np.float32(1.5))
        L.orc_refract(orc.f3(*i), orc.f3(*n), eta, a); ref.ref_refract(orc.f3(*i), orc.f3(*n), eta, b)
        assert _bits(a) == _bits(b)


def test_shade_hit(orc, ref, rt, scene_factory):
    L = orc.load()
    _, osc = scene_factory("d488")
    cam = rt.scenes.default_camera()
    rng = np.random.default_rng(7)
    view = orc.f3(*cam.origin)
    lit = 0
    a, b = _f(3), _f(3)
    for k in range(1500):
        o = np.array(cam.origin) + rng.normal(size=3) * 0.01
        tgt = np.array([rng.uniform(-8, 8), 0.4 * rng.uniform(0, 1), rng.uniform(-8, 8)])
        d = (tgt - o) / np.linalg.norm(tgt - o)
        i1 = L.orc_shade_hit(C.byref(osc.c), view, orc.f3(*o), orc.f3(*d), orc.ACCEL_BRUTE, a)
        i2 = ref.ref_shade_hit(C.byref(osc.c), view, orc.f3(*o), orc.f3(*d), b)
        assert i1 == i2
        if i1 >= 0:
            assert _bits(a) == _bits(b)
            lit += any(x > 0 for x in a)
    assert lit > 300


def test_scatter_sequences_follow_the_atomic_counters(orc, ref, rt, scene_factory):
    """N consecutive Material::Scatter calls on ONE material (reference code, counters advancing as in material.cpp:29,44,82,151)
    against the oracle's scatter fed with the Halton numbers those counters would produce."""
    L = orc.load()
    _, osc = scene_factory("d488")
    flat_m = osc.materials[osc.spheres["material"]]["kind"]
    rng = np.random.default_rng(8)
    n = 200
    for kind in (orc.MAT_OPAQUE, orc.MAT_METAL, orc.MAT_TRANSPARENT):
        sphere = int(np.nonzero(flat_m == kind)[0][3])
        s = osc.spheres[sphere]
        c = np.array([s["cx"], s["cy"], s["cz"]], np.float64)
        o = np.empty((n, 3), np.float32)
        d = np.empty((n, 3), np.float32)
        for k in range(n):
            inside = kind == orc.MAT_TRANSPARENT and k % 2
            org = c + rng.normal(size=3) * (0.3 * s["radius"] if inside else 3.0)
            tgt = c + rng.normal(size=3) * 0.4 * s["radius"]
            o[k] = org
            d[k] = (tgt - org) / np.linalg.norm(tgt - org)
        sc = np.zeros(n, np.int32)
        at, oo, od = (np.zeros((n, 3), np.float32) for _ in range(3))
        ref.ref_scatter_sequence(C.byref(osc.c), sphere, n, o.ctypes.data, d.ctypes.data, sc.ctypes.data, at.ctypes.data, oo.ctypes.data, od.ctypes.data)
        coin = sample = 0
        a, b, e = _f(3), _f(3), _f(3)
        seen = set()
        for k in range(n):
            if sc[k] < 0:
                continue                                  # the ray missed the sphere
            if kind == orc.MAT_TRANSPARENT:
                xi, u1, u2 = L.orc_halton(sample, 7), 0.0, 0.0
                sample += 1
            else:
                # front-face test first (material.cpp:22,74): back-facing hits consume no sample
                r0 = L.orc_scatter(C.byref(osc.c), sphere, orc.f3(*o[k]), orc.f3(*d[k]), -1.0, 0.0, 0.0, a, b, e)   # xi = -1: every lane "reflects"
                if r0 == 0:
                    assert sc[k] == 0
                    continue
                xi = L.orc_halton(coin, 3)
                coin += 1
                u1 = u2 = 0.0
                if kind == orc.MAT_OPAQUE:
                    # the diffuse branch draws HaltonSampleHemisphere(m_sampleIndex++) only when the coin says "diffuse"
                    r1 = L.orc_scatter(C.byref(osc.c), sphere, orc.f3(*o[k]), orc.f3(*d[k]), xi, 0.5, 0.5, a, b, e)
                    if list(a) != [1.0, 1.0, 1.0]:
                        u1, u2 = L.orc_halton(sample, 5), L.orc_halton(sample, 7)
                        sample += 1
            r = L.orc_scatter(C.byref(osc.c), sphere, orc.f3(*o[k]), orc.f3(*d[k]), xi, u1, u2, a, b, e)
            assert r == sc[k]
            if r == 1:
                assert _bits(a) == _bits(at[k]) and _bits(b) == _bits(oo[k]) and _bits(e) == _bits(od[k]), (kind, k)
                seen.add(tuple(a))
        assert coin + sample > n // 2


def test_resolve_ldr(orc, ref):
    rng = np.random.default_rng(9)
    hdr = (rng.uniform(0, 1, (64, 64, 4)) ** 3 * 4).astype(np.float32)
    hdr[0, 0] = 0
    out = np.empty((64, 64), np.uint32)
    ref.ref_resolve_ldr(hdr.ctypes.data, 64 * 64, 3, out.ctypes.data)
    assert np.array_equal(out, orc.resolve_ldr(hdr, 3))


@pytest.mark.parametrize("name,w,h,spp,depth", [("small", 96, 54, 3, 50), ("d488", 128, 72, 2, 50), ("d488", 96, 54, 2, 0), ("d488", 64, 36, 2, 3)])
def test_whole_render_bit_identical(orc, ref, rt, scene_factory, name, w, h, spp, depth):
    """spp x DrawBitmap on the reference's own classes vs the restatement (Halton counters, recursive integrator, one thread
    so the shared counters are consumed in the same order).  BVH and brute force both: same hits, same bytes."""
    _, osc = scene_factory(name)
    cp = orc_camera(orc, rt.scenes.default_camera(aspect=w / h))
    r, rst = orc.ref_render(osc, cp, orc.config(w, h, spp=spp, max_depth=depth, threads=1))
    for accel in (orc.ACCEL_BVH, orc.ACCEL_BRUTE):
        o, ost = orc.render(osc, cp, orc.config(w, h, spp=spp, max_depth=depth, threads=1, rng_mode=orc.RNG_HALTON, accel=accel,
                                                integrator=orc.INTEGRATOR_RECURSIVE, math_mode=orc.MATH_REFERENCE))
        assert (ost.closest_rays, ost.shadow_rays) == (rst.closest_rays, rst.shadow_rays)
        assert np.array_equal(r, o)


def test_metal_alpha_lane_any_of_four(orc, ref):
    """Metal::Scatter reflects when ANY of the four lanes of f0 + (1 - f0)(1 - n.v)^5 exceeds the random number
    (material.cpp:77-85); lane w is the texture's alpha.  With alpha < 1 and a dark reflectance the answer depends on the
    angle and on the Halton coin: tier A (the reference's own Metal::Scatter) against tier B fed with the same coins."""
    L = orc.load()
    q = lambda x: L.orc_xmcolor_load(L.orc_xmcolor_quantize(x))  # noqa: E731  XMCOLOR round trip: k * (1/255.0f)
    tex = np.zeros(1, orc.TEXTURE_DTYPE)
    tex["kind"], tex["c0"], tex["c1"] = orc.MAT_OPAQUE * 0, [[q(0.1), q(0.15), q(0.05), q(0.3)]], [[q(0.1), q(0.15), q(0.05), q(0.3)]]
    mat = np.zeros(1, orc.MATERIAL_DTYPE)
    mat["kind"], mat["texture"], mat["smoothness"] = orc.MAT_METAL, 0, 0.0
    sph = np.zeros(1, orc.SPHERE_DTYPE)
    sph["cx"], sph["cy"], sph["cz"], sph["radius"], sph["material"] = 0.5, 1.0, -0.25, 1.0, 0
    osc = orc.Scene(sph, mat, tex, np.zeros(0, orc.LIGHT_DTYPE), 1.0, 0)
    rng = np.random.default_rng(17)
    n = 400
    c = np.array([0.5, 1.0, -0.25])
    o = (c + rng.normal(size=(n, 3)) * 3.0).astype(np.float32)
    tgt = c + rng.normal(size=(n, 3)) * 0.55
    d = tgt - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    sc = np.zeros(n, np.int32)
    at, oo, od = (np.zeros((n, 3), np.float32) for _ in range(3))
    ref.ref_scatter_sequence(C.byref(osc.c), 0, n, o.ctypes.data, d.ctypes.data, sc.ctypes.data, at.ctypes.data, oo.ctypes.data, od.ctypes.data)
    a, b, e = _f(3), _f(3), _f(3)
    coin, outcomes = 0, {0: 0, 1: 0}
    L.orc_set_math_mode(orc.MATH_REFERENCE)
    try:
        for k in range(n):
            if sc[k] < 0:
                continue
            if L.orc_scatter(C.byref(osc.c), 0, orc.f3(*o[k]), orc.f3(*d[k]), -1.0, 0.0, 0.0, a, b, e) == 0:
                assert sc[k] == 0                                 # back face: no coin consumed (material.cpp:74)
                continue
            xi = L.orc_halton(coin, 3)
            coin += 1
            r = L.orc_scatter(C.byref(osc.c), 0, orc.f3(*o[k]), orc.f3(*d[k]), xi, 0.0, 0.0, a, b, e)
            assert r == sc[k], k
            outcomes[r] += 1
            if r:
                assert _bits(a) == _bits(at[k]) and _bits(e) == _bits(od[k])
    finally:
        L.orc_set_math_mode(orc.MATH_PINNED)
    assert outcomes[0] > 40 and outcomes[1] > 40                 # both answers really occur


@pytest.mark.parametrize("flags,first", [(3, 1), (0, 1), (0, 9), (1, 4), (2, 1000)])
def test_primary_hits_identical(orc, ref, rt, scene_factory, flags, first):
    """The reference's BVH (std::rand axes) and the brute-force order pick the same sphere for every pixel."""
    for name in ("d488", "s4096"):
        _, osc = scene_factory(name)
        cp = orc_camera(orc, rt.scenes.default_camera(aspect=320 / 180))
        ids, mids = orc.ref_primary_hits(osc, cp, orc.config(320, 180, flags=flags, first_sample=first))
        oids, omids = orc.primary_hits(osc, cp, orc.config(320, 180, flags=flags, first_sample=first, math_mode=orc.MATH_REFERENCE))
        assert np.array_equal(ids, oids) and np.array_equal(mids, omids)


def test_pinned_math_deviations_are_small(orc, rt, scene_factory):
    """D1-D3 (pow5 by multiplication, pinned sin/cos, exact 1/ior instead of rcpps): the counter-RNG paths are the same
    except where a 1-ulp change flips a Fresnel coin, so the two images agree closely at equal seeds."""
    _, osc = scene_factory("d488")
    cp = orc_camera(orc, rt.scenes.default_camera())
    a, _ = orc.render(osc, cp, orc.config(160, 90, spp=8, math_mode=orc.MATH_PINNED))
    b, _ = orc.render(osc, cp, orc.config(160, 90, spp=8, math_mode=orc.MATH_REFERENCE))
    same = np.all(np.isclose(a, b, rtol=1e-4, atol=1e-9), axis=-1).mean()
    assert same > 0.97
    assert abs(a[..., :3].mean() - b[..., :3].mean()) / b[..., :3].mean() < 2e-3
