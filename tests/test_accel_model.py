"""CPU model of the two conservative filters in front of the exact ray-sphere test (no GPU): a numpy restatement, in binary32,
of the packed discriminant cull (cpuraytracer_b200/csrc/rt_device.cuh cull_pair / rt_bvh.cuh cull_pair2) and of the padded slab
test of RT_FLAG_ACCEL (rt_bvh.cuh box_ray / box_touched), run over the hierarchy image the library builds on the host
(rt_scene_accel_image).  Property checked: whatever the exact test -- Sphere::Intersect in the reference's operation order,
ray-tracing.cpp:42-84 -- accepts is never filtered out, for rays that start on sphere surfaces, inside spheres, far away and
along the axes.  The GPU suite checks the kernels themselves (tests/test_gpu_parity.py::test_accel_*, test_cull_is_result_neutral);
this file pins the arithmetic they are built from."""
import numpy as np
import pytest

from test_host_api import _accel_sections, _random_flat_scene

f32 = np.float32


def fma(a, b, c):
    """binary32 fused multiply-add (the product of two binary32 numbers is exact in binary64)"""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def exact_accepts(o, d, cx, cy, cz, r):
    """Sphere::Intersect, reference operation order, one binary32 rounding per operation.  o, d: (R,3); spheres: (N,) -> (R,N) bool"""
    ocx, ocy, ocz = o[:, None, 0] - cx[None], o[:, None, 1] - cy[None], o[:, None, 2] - cz[None]
    dx, dy, dz = d[:, None, 0], d[:, None, 1], d[:, None, 2]
    a = ((dx * dx + dy * dy) + dz * dz).astype(f32)
    b = ((ocx * dx + ocy * dy) + ocz * dz).astype(f32)
    c = (((ocx * ocx + ocy * ocy) + ocz * ocz) - (r * r)[None]).astype(f32)
    disc = (b * b - a * c).astype(f32)
    with np.errstate(invalid="ignore"):
        sq = np.sqrt(np.where(disc > 0, disc, f32(0))).astype(f32)
        t0, t1 = ((-b - sq) / a).astype(f32), ((-b + sq) / a).astype(f32)
    return (disc > 0) & ((t0 > f32(0.001)) | (t1 > f32(0.001)))


def cull_keeps(o, d, cx, cy, cz, w):
    """sign of the conservative discriminant, the fma chain of cull_pair2 (one ray)"""
    ox, oy, oz = (o[:, None, k] for k in range(3))
    dx, dy, dz = (d[:, None, k] for k in range(3))
    oo = fma(ox, ox, fma(oy, oy, (oz * oz).astype(f32)))
    noo = fma(f32(2.0 ** -16), oo, -oo)
    od = fma(ox, dx, fma(oy, dy, (oz * dz).astype(f32)))
    nc2 = (w[None] + noo).astype(f32)
    nc2 = fma((oz + oz).astype(f32), cz[None], nc2); nc2 = fma((oy + oy).astype(f32), cy[None], nc2); nc2 = fma((ox + ox).astype(f32), cx[None], nc2)
    b1 = fma(-dz, cz[None], od); b1 = fma(-dy, cy[None], b1); b1 = fma(-dx, cx[None], b1)
    d1 = fma(b1, b1, nc2)
    return ~np.signbit(d1)


def slab_touched(o, d, info, bc, bh):
    """box_ray + box_touched for rays (R,3) against boxes (M,3) centre / half extent -> (R,M) bool"""
    cs = np.asarray(list(info.bound_center), f32)
    dc = (o - cs[None]).astype(f32)
    dmax = (np.sqrt(fma(dc[:, 0], dc[:, 0], fma(dc[:, 1], dc[:, 1], (dc[:, 2] * dc[:, 2]).astype(f32)))).astype(f32) + f32(info.bound_radius)).astype(f32)
    E = (f32(2.0 ** -18) * fma(dmax, dmax, f32(info.r_max) * f32(info.r_max))).astype(f32)
    pad = ((f32(1.5) * np.minimum((E / (f32(2.0) * f32(info.r_min))).astype(f32), np.sqrt(E).astype(f32))).astype(f32)
           + (f32(2.0 ** -19) * (((np.abs(o[:, 0]) + np.abs(o[:, 1])).astype(f32) + np.abs(o[:, 2])).astype(f32) + f32(info.c_max)).astype(f32)).astype(f32)).astype(f32)
    nts = (f32(-(2.0 ** -16)) * dmax).astype(f32)
    dd = np.where(np.abs(d) < f32(1e-18), np.copysign(f32(1e-18), d), d).astype(f32)
    inv = (f32(1.0) / dd).astype(f32)
    ax = np.abs(inv)
    no = (-(o * inv)).astype(f32)
    ex = (pad[:, None] * ax).astype(f32)
    n, f = (no - ex).astype(f32), (no + ex).astype(f32)
    tn = np.full((o.shape[0], bc.shape[0]), -np.inf, f32)
    tf = np.full((o.shape[0], bc.shape[0]), np.inf, f32)
    for k in range(3):
        m_n = fma(bc[None, :, k], inv[:, None, k], n[:, None, k])
        m_f = fma(bc[None, :, k], inv[:, None, k], f[:, None, k])
        tn = np.maximum(tn, fma(-bh[None, :, k], ax[:, None, k], m_n))
        tf = np.minimum(tf, fma(bh[None, :, k], ax[:, None, k], m_f))
    return np.maximum(tn, nts[:, None]) <= tf


def _rays(rng, centres, radii, n_rays):
    """origins on sphere surfaces (as every bounce ray and shadow ray has), inside spheres, inside the field, far away; random and
    axis-aligned directions"""
    n = len(radii)
    pick = rng.integers(0, n, n_rays)
    u = rng.normal(size=(n_rays, 3)); u /= np.linalg.norm(u, axis=1, keepdims=True)
    on = centres[pick] + radii[pick, None] * u
    inside = centres[pick] + 0.5 * radii[pick, None] * u
    lo, hi = centres.min(0), centres.max(0)
    field = rng.uniform(lo - 1, hi + 1, (n_rays, 3))
    far = field * rng.choice([3.0, 30.0, 300.0], (n_rays, 1))
    o = np.concatenate([on, inside, field, far]).astype(f32)
    d = rng.normal(size=(len(o), 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    # the critical family for the pad: rays that graze a sphere at (almost) the point where it touches its own bounding box,
    # travelling inside that face of the box
    m = n_rays
    pk, axis, sign = rng.integers(0, n, m), rng.integers(0, 3, m), rng.choice([-1.0, 1.0], m)
    depth = rng.choice([0.0, 1e-7, 1e-6, 1e-5, 1e-3], m)
    e = np.zeros((m, 3)); e[np.arange(m), axis] = sign
    p = centres[pk] + radii[pk, None] * (1.0 - depth[:, None]) * e
    t = rng.normal(size=(m, 3)); t -= (t * e).sum(1, keepdims=True) * e; t /= np.linalg.norm(t, axis=1, keepdims=True)
    go = (p - rng.uniform(0.5, 50.0, (m, 1)) * t).astype(f32)
    gd = (p - go.astype(np.float64)); gd /= np.linalg.norm(gd, axis=1, keepdims=True)
    o = np.concatenate([o, go]); d = np.concatenate([d, gd])
    n4 = 4 * n_rays                                                  # the first four families get axis-aligned and aimed directions
    ax = rng.integers(0, 3, n4 // 8)
    d[: len(ax)] = 0.0
    d[np.arange(len(ax)), ax] = rng.choice([-1.0, 1.0], len(ax))
    aim = rng.integers(0, n, n4 // 2)                                # half of them aimed at (near) a sphere: plenty of hits and grazes
    tgt = centres[aim] + radii[aim, None] * rng.uniform(-1.05, 1.05, (len(aim), 3))
    v = tgt - o[n4 // 2: n4].astype(np.float64)
    d[n4 // 2: n4] = v / np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-30)
    return o, d.astype(f32)


@pytest.mark.parametrize("which", ["d488", "random600", "s4096"])
def test_filters_never_drop_what_the_exact_test_accepts(rt, which):
    scene = _random_flat_scene(rt, 21, 600, ground=True) if which == "random600" else rt.scenes.NAMED[which]()
    info, img = scene.accel_image()
    cx, cy, cz, w, R, I, box, top = _accel_sections(info, img)
    used = I != 0xFFFF
    rng = np.random.default_rng(3)
    centres = np.stack([cx, cy, cz], 1)[used].astype(np.float64)
    o, d = _rays(rng, centres, R[used].astype(np.float64), 300 if which == "s4096" else 600)
    acc = exact_accepts(o, d, cx, cy, cz, R) & used[None]            # (rays, slots)
    assert acc.sum() > 2 * len(o) // 3                                 # the sample really exercises hits
    keep = cull_keeps(o, d, cx, cy, cz, w)
    assert not np.any(acc & ~keep), "the discriminant cull dropped an accepted sphere"
    # group boxes / L2 boxes of the small spheres
    gs, first = info.group_slots, 32 * info.n_big_groups
    g_of_slot = (np.arange(2 * info.n_pairs) - first) // gs
    gi = np.arange(info.n_groups)
    bc = np.stack([box[gi // 8, a * 8 + gi % 8] for a in range(3)], 1)
    bh = np.stack([box[gi // 8, 24 + a * 8 + gi % 8] for a in range(3)], 1)
    touched = slab_touched(o, d, info, bc, bh)                        # (rays, groups)
    tt = slab_touched(o, d, info, top[:, 0:3], top[:, 3:6])           # (rays, L2 nodes)
    small = np.arange(2 * info.n_pairs) >= first
    rr, ss = np.nonzero(acc & small[None])
    assert len(rr) > 0
    assert np.all(touched[rr, g_of_slot[ss]]), "a group box was skipped although the exact test accepts one of its spheres"
    assert np.all(tt[rr, g_of_slot[ss] // 8]), "an L2 box was skipped although the exact test accepts one of its spheres"
    # the boxes do filter: far fewer than all (ray, group) pairs are touched
    assert touched.mean() < 0.5


# ---------------------------------------------------------------- brute force: the cull in the light-aligned frame (rt_device.cuh, SceneView)
def _rot(m, v):
    """rotate3(): fmaf(m0, x, fmaf(m1, y, m2 * z)) per row, binary32"""
    x, y, z = v[:, 0], v[:, 1], v[:, 2]
    return np.stack([fma(f32(m[3 * k]), x, fma(f32(m[3 * k + 1]), y, (f32(m[3 * k + 2]) * z).astype(f32))) for k in range(3)], 1)


def shadow_cull_keeps(oq, cx, cy, cz, w):
    """cull_pair<RAYS_SHADOW>: -c' as for the closest ray, b2 = o'z - c'z, sign of b2^2 - c'"""
    ox, oy, oz = (oq[:, None, k] for k in range(3))
    oo = fma(ox, ox, fma(oy, oy, (oz * oz).astype(f32)))
    noo = fma(f32(2.0 ** -16), oo, -oo)
    nc2 = (w[None] + noo).astype(f32)
    nc2 = fma((oz + oz).astype(f32), cz[None], nc2); nc2 = fma((oy + oy).astype(f32), cy[None], nc2); nc2 = fma((ox + ox).astype(f32), cx[None], nc2)
    b2 = fma(f32(-1.0), cz[None], oz)
    return ~np.signbit(fma(b2, b2, nc2))


@pytest.mark.parametrize("which,light", [("d488", 0), ("s4096", 0), ("random600", 0), ("three_lights", 1), ("three_lights", 2), ("d488", -1)])
def test_light_aligned_cull_never_drops_what_the_exact_test_accepts(rt, which, light):
    """The brute-force query culls in a rotated frame (third axis = the light direction) with coordinates rounded once to binary32 and
    rays rotated in binary32 on the device; the exact test runs in world space.  Whatever Sphere::Intersect accepts -- for the
    closest ray and for the shadow ray along L -- must survive the cull, and the arrays must be what the header says."""
    if which == "random600":
        scene = _random_flat_scene(rt, 21, 600, ground=True)
        scene.add_light(rt.DirectionalLight((0.3, 0.9, -0.4), rt.XMCOLOR(1.0, 1.0, 1.0), 30000.0))      # the helper's scenes have no light
    elif which == "three_lights":
        scene = rt.scenes.scene_d488()
        scene.add_light(rt.DirectionalLight((-1.0, 0.6, 0.2), rt.XMCOLOR(0.3, 0.5, 1.0), 25000.0))
        scene.add_light(rt.DirectionalLight((1e-4, 0.0, -1.0), rt.XMCOLOR(1.0, 0.4, 0.3), 15000.0))
    else:
        scene = rt.scenes.NAMED[which]()
    info, A, B, G = scene.cull_image(light)
    n = info.n_spheres
    rot = np.asarray(list(info.rot), np.float64)
    R3 = rot.reshape(3, 3)
    assert np.allclose(R3 @ R3.T, np.eye(3), atol=1e-6) and abs(np.linalg.det(R3) - 1.0) < 1e-5      # a rotation
    Ld = np.asarray(list(info.light_dir), np.float64)
    if light >= 0:
        assert np.allclose(R3[2], Ld / np.linalg.norm(Ld), atol=1e-6)                                 # third axis = the light direction
    cxq, cyq, czq = A[:, 0:2].reshape(-1)[:n], A[:, 2:4].reshape(-1)[:n], B[:, 0:2].reshape(-1)[:n]
    w = B[:, 2:4].reshape(-1)[:n]
    sp = scene.export().spheres
    assert np.array_equal(G[:n, 0], sp["cx"]) and np.array_equal(G[:n, 1], sp["cy"]) and np.array_equal(G[:n, 2], sp["cz"]) and np.array_equal(G[:n, 3], sp["radius"])
    cw = np.stack([sp["cx"], sp["cy"], sp["cz"]], 1).astype(np.float64)
    assert np.allclose(np.stack([cxq, cyq, czq], 1), cw @ R3.T, rtol=0, atol=2e-7 * (1.0 + np.abs(cw).max()))   # c' = rot c, one rounding
    r64 = sp["radius"].astype(np.float64)
    cc = (cw * cw).sum(1)
    assert np.allclose(w, r64 * r64 - cc + 2.0 ** -16 * (cc + r64 * r64), rtol=1e-6, atol=1e-6)
    assert np.all(B[:, 2:4].reshape(-1)[n:] == f32(-1.0e30))                                           # padding never passes
    rng = np.random.default_rng(5)
    o, d = _rays(rng, cw, r64, 300 if which == "s4096" else 600)
    cx, cy, cz, rad = (G[:n, k].copy() for k in range(4))
    oq, dq = _rot(info.rot, o), _rot(info.rot, d)
    acc = exact_accepts(o, d, cx, cy, cz, rad)
    assert acc.sum() > 2 * len(o) // 3
    assert not np.any(acc & ~cull_keeps(oq, dq, cxq, cyq, czq, w)), "the rotated-frame cull dropped a sphere the closest-hit test accepts"
    if light >= 0:
        Lf = np.broadcast_to(np.asarray(list(info.light_dir), f32), o.shape).copy()
        accs = exact_accepts(o, Lf, cx, cy, cz, rad)
        assert accs.sum() > len(o) // 10
        assert not np.any(accs & ~shadow_cull_keeps(oq, cxq, cyq, czq, w)), "the rotated-frame cull dropped an occluder"
        # and it does cull: most (ray, sphere) pairs are discarded
        assert shadow_cull_keeps(oq, cxq, cyq, czq, w).mean() < 0.2
