"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/rt_b200.h
declares, the scene recorder flattens the reference constructors exactly, and errors are loud."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, orc_camera


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    return sorted(set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol(rt):
    lib = rt.load()
    names = _declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"librt_b200.so does not export {n}"
    from cpuraytracer_b200 import _lib
    assert sorted(s[0] for s in _lib.SYMBOLS) == names      # the ctypes table covers the header exactly
    assert b"sm_100a" in lib.rt_build_info()


def test_library_is_sm100a_only(rt):
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", rt.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback(rt):
    """Without a CUDA device the compute API must fail loudly (this container has none)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    with pytest.raises(rt.RtError) as e:
        rt.Renderer(width=8, height=8)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under cpuraytracer_b200/ or include/ may reference it."""
    for base in ("cpuraytracer_b200", "include"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    txt = open(os.path.join(dp, fn), errors="replace").read()
                    assert "rt_oracle" not in txt and "liborc" not in txt and "from oracle" not in txt and "import oracle" not in txt, fn


def test_scene_recorder_counts_and_ids(rt):
    s = rt.scenes.scene_d488()
    ns, nm, nt, nl = s.counts()
    assert (ns, nm, nl) == (488, 488, 1)          # each Sphere owns its material (ray-tracing.h:57)
    flat = s.export()
    sp = flat.spheres
    # insertion order of InitScene (spheres-app.cpp:60-114)
    assert tuple(sp[0])[:4] == (0.0, -1000.0, 0.0, 1000.0)
    assert tuple(sp[485])[:4] == (0.0, 1.0, 0.0, 1.0) and flat.materials[sp[485]["material"]]["kind"] == rt.RT_MAT_TRANSPARENT
    assert tuple(sp[486])[:4] == (-4.0, 1.0, 0.0, 1.0) and flat.materials[sp[486]["material"]]["kind"] == rt.RT_MAT_OPAQUE
    assert tuple(sp[487])[:4] == (4.0, 1.0, 0.0, 1.0) and flat.materials[sp[487]["material"]]["kind"] == rt.RT_MAT_METAL
    small = sp[1:485]
    assert np.all(small["radius"] == np.float32(0.2)) and np.all(small["cy"] == np.float32(0.2))
    a = np.repeat(np.arange(-11, 11), 22)
    b = np.tile(np.arange(-11, 11), 22)
    assert np.all((small["cx"] >= a) & (small["cx"] < a + 0.9001)) and np.all((small["cz"] >= b) & (small["cz"] < b + 0.9001))
    kinds = flat.materials[small["material"]]["kind"]
    frac = [np.mean(kinds == k) for k in (0, 1, 2)]
    assert 0.7 < frac[0] < 0.9 and 0.08 < frac[1] < 0.22 and 0.01 < frac[2] < 0.1     # 80 / 15 / 5 % recipe
    assert flat.sky_luminance == 8000.0 and flat.lights[0]["luminance"] == 40000.0
    assert list(flat.lights[0]["dir"]) == [1.0, 1.0, 1.0]      # the ctor argument; consumers normalise (light.cpp:7)


def test_scene_is_reproducible_and_s4096(rt):
    a, b = rt.scenes.scene_d488().export(), rt.scenes.scene_d488().export()
    assert a.spheres.tobytes() == b.spheres.tobytes() and a.textures.tobytes() == b.textures.tobytes()
    assert rt.scenes.scene_s4096().counts()[0] == 4096


def test_colours_are_xmcolor_quantised(rt, orc):
    """XMCOLOR(float x4) -> XMLoadColor round trip (texture.cpp:5,16-17; light.cpp:8): floats are k * (1/255)."""
    flat = rt.scenes.scene_d488().export()
    L = orc.load()
    for arr in (flat.textures["c0"], flat.textures["c1"], flat.lights["color"]):
        k = np.rint(arr.astype(np.float64) * 255).astype(int)
        assert np.array_equal(arr, (k.astype(np.float32) * np.float32(1.0 / 255.0)))
    ground = flat.textures[flat.materials[flat.spheres[0]["material"]]["texture"]]
    assert ground["kind"] == 1 and ground["tiling"] == 2500.0
    exp0 = [L.orc_xmcolor_load(L.orc_xmcolor_quantize(v)) for v in (0.9, 0.9, 0.9, 1.0)]
    exp1 = [L.orc_xmcolor_load(L.orc_xmcolor_quantize(v)) for v in (0.2, 0.3, 0.1, 1.0)]
    assert list(ground["c0"]) == exp0 and list(ground["c1"]) == exp1
    assert [int(round(v * 255)) for v in exp0[:3]] == [230, 230, 230] and [int(round(v * 255)) for v in exp1[:3]] == [51, 76, 26]   # 0.3f*255 rounds to 76.5 in fp32 -> ties-to-even 76
    sky = flat.textures[flat.sky_texture]
    assert [int(round(v * 255)) for v in sky["c0"][:3]] == [217, 232, 250]


def test_camera_basis_matches_oracle(rt, orc):
    for cam in (rt.scenes.default_camera(), rt.Camera((1, 2, 3), (-4, 0.5, 9), 60.0, 1.0, 5.0, 0.1)):
        b = cam.basis()
        oc = orc.OrcCamera()
        cp = orc_camera(orc, cam)
        orc.load().orc_camera_make(C.byref(cp), C.byref(oc))
        for f in ("origin", "x", "y", "image_origin"):
            assert list(getattr(b, f)) == list(getattr(oc, f)), f
        assert b.aperture == oc.aperture and b.focus == oc.focus
    assert abs(rt.scenes.default_camera().focal_length - np.sqrt(151.25)) < 1e-5     # |origin - lookAt|, spheres-app.cpp:45


def test_recorder_error_behaviour(rt):
    s = rt.Scene()
    with pytest.raises(rt.RtError) as e:
        s._lib.rt_scene_add_sphere.restype  # noqa: B018  (symbol exists)
        from cpuraytracer_b200 import _lib
        _lib.check_id(s._lib.rt_scene_add_sphere(s.handle, (C.c_float * 3)(0, 0, 0), 1.0, 5))
    assert e.value.code == 1 and "material id 5" in str(e.value)
    from cpuraytracer_b200 import _lib
    with pytest.raises(rt.RtError):
        _lib.check_id(s._lib.rt_scene_add_material_opaque(s.handle, 3, 16.0))          # texture out of range
    with pytest.raises(rt.RtError):
        _lib.check_id(s._lib.rt_scene_add_material_transparent(s.handle, 16.0, 0.0))   # ior must be > 0
    t = s.add_texture(rt.ConstTexture(rt.XMCOLOR(1, 1, 1)))
    m = _lib.check_id(s._lib.rt_scene_add_material_opaque(s.handle, t, 16.0))
    with pytest.raises(rt.RtError):
        _lib.check_id(s._lib.rt_scene_add_sphere(s.handle, (C.c_float * 3)(0, 0, 0), -1.0, m))   # radius <= 0
    with pytest.raises(rt.RtError):
        _lib.check(s._lib.rt_scene_add_directional_light(s.handle, (C.c_float * 3)(0, 0, 0), (C.c_float * 4)(1, 1, 1, 1), 1.0))
    with pytest.raises(TypeError):
        s.add(rt.Sphere((0, 0, 0), 1.0, rt.Emissive(1.0, rt.ConstTexture(rt.XMCOLOR(1, 1, 1)))))


def test_flat_scene_round_trip(rt, tmp_path):
    flat = rt.scenes.scene_small().export()
    p = str(tmp_path / "scene.npz")
    flat.save(p)
    back = rt.Scene.from_flat(rt.FlatScene.load(p)).export()
    for f in ("spheres", "materials", "textures", "lights"):
        assert getattr(flat, f).tobytes() == getattr(back, f).tobytes()
    assert (flat.sky_luminance, flat.sky_texture) == (back.sky_luminance, back.sky_texture)


def test_image_sinks(rt, tmp_path):
    bgra = np.array([[0xFF102030, 0xFFFFFFFF]], np.uint32)
    p = str(tmp_path / "a.ppm")
    rt.write_ppm(p, bgra)
    raw = open(p, "rb").read()
    assert raw.startswith(b"P6\n2 1\n255\n") and raw[-6:] == bytes([0x10, 0x20, 0x30, 255, 255, 255])
    hdr = np.ones((2, 3, 4), np.float32) * 4
    q = str(tmp_path / "a.pfm")
    rt.write_pfm(q, hdr, 4)
    raw = open(q, "rb").read()
    assert raw.startswith(b"PF\n3 2\n-1.0\n") and np.frombuffer(raw[-72:], "<f4").tolist() == [1.0] * 18


def test_bench_traffic_lookup_and_reference_arm_shape():
    """bench.py's roofline.traffic comes from the committed ncu capture of the bench launch (profiles/traffic.json), keyed by
    workload and query mode; a workload without a capture reports None instead of a guess."""
    import importlib.util
    import json
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    table = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    for key in ("c3:brute", "c3:accel"):
        wl, q = key.split(":")
        assert bench._ncu_traffic(wl, q) == table[key]["bytes"] == sum(k["bytes"] for k in table[key]["per_kernel"].values())
        # traffic well above the algorithmic bytes = wasted re-reads.  The excess here (2.8x, 21 GB/s in all) is half-sector float4 stores
        # of the running / partial sums (profiles/r02/r02o_traffic_variants.txt); round 1 stood at 16x
        assert table[key]["bytes"] < 3.0 * table[key]["algorithmic_bytes"]
    assert bench._ncu_traffic("c4", "brute") is None
    assert bench._ncu_traffic("nope", "brute") is None


# ---------------------------------------------------------------- RT_FLAG_ACCEL: invariants of the hierarchy image (host only)
def _accel_sections(info, img):
    f32 = img.view(np.float32)
    n_slots = 2 * info.n_pairs
    A = f32[: info.off_B // 4][: 4 * info.n_pairs].reshape(-1, 4)
    B = f32[info.off_B // 4: info.off_B // 4 + 4 * info.n_pairs].reshape(-1, 4)
    R = f32[info.off_R // 4: info.off_R // 4 + n_slots]
    I = img[info.off_I: info.off_I + 2 * n_slots].view(np.uint16)
    box = f32[info.off_box // 4: info.off_top // 4].reshape(info.n_l2_nodes, info.node_floats)
    top = f32[info.off_top // 4: info.off_top // 4 + 8 * info.n_l2_nodes].reshape(info.n_l2_nodes, 8)
    slot = np.arange(n_slots)
    cx, cy = A[slot >> 1, slot & 1], A[slot >> 1, 2 + (slot & 1)]
    cz, w = B[slot >> 1, slot & 1], B[slot >> 1, 2 + (slot & 1)]
    return cx, cy, cz, w, R, I, box, top


def _random_flat_scene(rt, seed, n, ground):
    rng = np.random.default_rng(seed)
    s = rt.Scene()
    mat = rt.DielectricOpaque(rt.ConstTexture(rt.XMCOLOR(0.5, 0.5, 0.5)), 16.0)
    if ground:
        s.add(rt.Sphere((0.0, -1000.0, 0.0), 1000.0, mat))
    for _ in range(n):
        r = float(0.05 * 8.0 ** rng.uniform(0, 1))
        c = rng.uniform(-30, 30, 3)
        s.add(rt.Sphere(tuple(float(x) for x in c), r, mat))
    s.set_sky(rt.Emissive(8000.0, rt.ConstTexture(rt.XMCOLOR(0.85, 0.91, 0.98))))
    return s


@pytest.mark.parametrize("which,expect_slots", [("d488", 8), ("s4096", 16), ("random600", 8), ("random6000", 32)])
def test_accel_image_invariants(rt, which, expect_slots):
    """What makes RT_FLAG_ACCEL equal to brute force is decided on the host: every sphere sits in exactly one slot with the
    same cull constants as the brute-force arrays, and every box contains the spheres below it, rounded outward."""
    if which.startswith("random"):
        scene = _random_flat_scene(rt, 5, int(which[6:]), ground=True)
    else:
        scene = rt.scenes.NAMED[which]()
    sp = scene.export().spheres
    centres = np.stack([sp["cx"], sp["cy"], sp["cz"]], 1).astype(np.float64)
    radii = sp["radius"].astype(np.float64)
    n = len(radii)
    info, img = scene.accel_image()
    assert info.group_slots == expect_slots and info.bytes == img.nbytes and info.bytes % 16 == 0 and info.off_B % 256 == 0
    assert info.n_groups <= 256 and info.n_l2_nodes == (info.n_groups + 7) // 8 <= 32
    cx, cy, cz, w, R, I, box, top = _accel_sections(info, img)
    used = I != 0xFFFF
    assert sorted(I[used].tolist()) == list(range(n))                                  # a permutation of the scene
    o = I[used].astype(np.int64)
    assert np.array_equal(np.stack([cx, cy, cz], 1)[used], centres[o].astype(np.float32)) and np.array_equal(R[used], radii[o].astype(np.float32))
    k = 2.0 ** -16
    cc, r2 = (centres[o] ** 2).sum(1), radii[o] ** 2
    assert np.array_equal(w[used], (r2 - cc + k * (cc + r2)).astype(np.float32))       # same -c' term as the brute-force arrays
    assert np.all(w[~used] == np.float32(-1.0e30))                                     # padding can never be a candidate
    n_big_slots = 32 * info.n_big_groups
    big = I[:n_big_slots][I[:n_big_slots] != 0xFFFF]
    small_slots = np.arange(n_big_slots, 2 * info.n_pairs)
    small = I[small_slots]
    assert len(big) + int((small != 0xFFFF).sum()) == n
    if len(big) and (small != 0xFFFF).any():
        assert radii[big].min() > radii[small[small != 0xFFFF]].max()                  # big = the largest radii
    # boxes: group box and node box contain the spheres below them (centre / half extent, rounded outward)
    gs = info.group_slots
    for g in range(info.n_groups):
        j2, kk = divmod(g, 8)
        c = box[j2, [kk, 8 + kk, 16 + kk]].astype(np.float64)
        h = box[j2, [24 + kk, 32 + kk, 40 + kk]].astype(np.float64)
        members = small[g * gs:(g + 1) * gs]
        members = members[members != 0xFFFF].astype(np.int64)
        assert len(members) > 0
        lo, hi = (centres[members] - radii[members, None]).min(0), (centres[members] + radii[members, None]).max(0)
        assert np.all(c - h <= lo) and np.all(c + h >= hi), g
        tc, th = top[j2, 0:3].astype(np.float64), top[j2, 3:6].astype(np.float64)     # the node box must hold the SPHERES below it (each
        assert np.all(tc - th <= lo) and np.all(tc + th >= hi), g                      # level is rounded outward from the exact union on its own)
    for g in range(info.n_groups, 8 * info.n_l2_nodes):                                # children past the last group are empty
        j2, kk = divmod(g, 8)
        assert np.all(box[j2, [24 + kk, 32 + kk, 40 + kk]] == np.float32(-1.0e30))
    sm = small[small != 0xFFFF].astype(np.int64)
    if len(sm):
        bc = np.asarray(list(info.bound_center), np.float64)
        assert np.all(np.linalg.norm(centres[sm] - bc, axis=1) + radii[sm] <= info.bound_radius)
        assert info.r_min == np.float32(radii[sm].min()) and info.r_max == np.float32(radii[sm].max())
        assert info.c_max >= np.abs(np.concatenate([centres[sm] - radii[sm, None], centres[sm] + radii[sm, None]])).max()


def test_accel_image_group_size_override_and_limits(rt):
    scene = rt.scenes.NAMED["d488"]()
    slots = {}
    for g in (3, 4, 5):
        info, _ = scene.accel_image(min_group_slots=1 << g)      # rt_scene_accel_image_slots (no environment variable is read)
        slots[g] = (info.group_slots, info.n_groups)
    assert slots == {3: (8, 61), 4: (16, 31), 5: (32, 16)}
    with pytest.raises(rt.RtError):
        scene.accel_image(min_group_slots=12)
    too_many = _random_flat_scene(rt, 9, 8300, ground=False)                          # > 8192 small spheres: no hierarchy
    with pytest.raises(rt.RtError) as e:
        too_many.accel_image()
    assert e.value.code == 5                                                           # RT_ERR_UNSUPPORTED


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` times the reference's own CPU code (oracle/_ref when it was built, else the oracle port) and
    prints one JSON line with the arm's keys; under torchrun every rank but 0 exits 0 without work."""
    import json
    import subprocess
    import sys
    env = dict(os.environ, RANK="0", WORLD_SIZE="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "path samples/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("c3")
    other = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                           capture_output=True, text=True, timeout=600, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert other.returncode == 0 and other.stdout.strip() == ""


def test_frozen_scene_files_equal_the_recorder(rt, orc):
    """oracle/scenes/*.rtsc (the CPU arm's input) are byte-identical to what the product's recorder exports for the named
    scenes, and the oracle's own default camera equals the product's."""
    for name in ("d488", "s4096", "small"):
        flat = rt.scenes.NAMED[name]().export()
        fz = orc.frozen_scene(name)
        assert fz.spheres.tobytes() == flat.spheres.tobytes() and fz.materials.tobytes() == flat.materials.tobytes()
        assert fz.textures.tobytes() == flat.textures.tobytes() and fz.lights.tobytes() == flat.lights.tobytes()
        assert (fz.c.sky_luminance, fz.c.sky_texture) == (flat.sky_luminance, flat.sky_texture)
    for aspect in (16 / 9, 1280 / 720, 4 / 3):
        cam = rt.scenes.default_camera(aspect=aspect)
        a = orc.default_camera_params(aspect)
        b = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)
        assert bytes(a) == bytes(b)


def test_cpu_arm_does_not_map_the_product_library():
    """bench.py's CPU arm (cpu_baseline / --impl reference) takes its scene from the frozen files and runs oracle code only:
    after a frame, no librt_b200 is mapped into the process and cpuraytracer_b200 was never imported."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); import bench\n"
            "bench.WORKLOADS['t'] = ('small', 64, 36, 50, 1, 'tiny')\n"
            "arm = bench.CpuArm('t'); st = arm.frames(2); assert st.paths == 2 * 64 * 36\n"
            "d = arm.describe(); assert d['counters'].startswith('per-thread') and d['kind'] in ('reference', 'port')\n"
            "maps = open('/proc/self/maps').read()\n"
            "assert 'librt_b200' not in maps, 'product library mapped'\n"
            "assert not any(m.startswith('cpuraytracer_b200') for m in sys.modules), 'product package imported'\n"
            "print('ok', d['kind'], 'libref' in maps)\n") % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.startswith("ok"), out.stderr[-2000:]


def test_import_rejects_what_the_recorder_rejects(rt):
    """rt_scene_import (scene files, Scene.from_flat) applies the recorder's checks: finite geometry, radius > 0, ior > 0,
    known texture kinds, finite colours / tiling / luminance."""
    base = rt.scenes.scene_small().export()
    import copy

    def broken(edit):
        f = copy.deepcopy(base)
        edit(f)
        with pytest.raises(rt.RtError) as e:
            rt.Scene.from_flat(f)
        assert e.value.code == 1

    glass = int(np.nonzero(base.materials["kind"] == 2)[0][0])
    broken(lambda f: f.materials["ior"].__setitem__(glass, 0.0))
    broken(lambda f: f.materials["ior"].__setitem__(glass, np.inf))
    broken(lambda f: f.spheres["cx"].__setitem__(3, np.nan))
    broken(lambda f: f.spheres["radius"].__setitem__(3, np.inf))
    broken(lambda f: f.textures["kind"].__setitem__(0, 7))
    broken(lambda f: f.textures["tiling"].__setitem__(0, np.nan))
    broken(lambda f: f.textures["c0"].__setitem__((1, 2), np.inf))
    broken(lambda f: f.lights["luminance"].__setitem__(0, np.nan))
    broken(lambda f: f.lights["dir"].__setitem__(0, 0.0))
    rt.Scene.from_flat(base)                                     # the unbroken scene still loads
