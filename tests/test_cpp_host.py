"""The C++17 host side above the C ABI: include/rt_b200.hpp mirror classes, include/rt_image_io.hpp sinks,
librt_b200_group.so (sample-range sharding + ncclReduce) and the apps/spheres CLI (headless SpheresApp).

CPU part: builds, exports, scene recipe == the Python recorder byte for byte, sinks, loud failure without a GPU.
GPU part: the CLI renders the same bits as the Python mirror, checkpoints resume exactly, 2 GPUs == 1 GPU.
"""
import ctypes as C
import json
import os
import re
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="session")
def host(rt):
    from cpuraytracer_b200 import build
    group_lib, cli = build.build_host()
    return group_lib, cli


def run_cli(cli, *args, check=True):
    r = subprocess.run([cli, *[str(a) for a in args]], capture_output=True, text=True, timeout=600)
    if check:
        assert r.returncode == 0, r.stderr
    return r


def read_rtsc(path, L):
    b = open(path, "rb").read()
    assert b[:4] == b"RTSC"
    ver, ns, nm, nt, nl = struct.unpack_from("<5i", b, 4)
    sky_lum, sky_tex = struct.unpack_from("<fi", b, 24)
    assert ver == 1
    p = 32
    out = []
    for n, dt in ((ns, L.SPHERE_DTYPE), (nm, L.MATERIAL_DTYPE), (nt, L.TEXTURE_DTYPE), (nl, L.LIGHT_DTYPE)):
        out.append(np.frombuffer(b, dt, n, p))
        p += n * dt.itemsize
    assert p == len(b)
    return out, sky_lum, sky_tex


def read_exr(path):
    """Independent minimal reader of the uncompressed scanline EXR the C++ sink writes."""
    b = open(path, "rb").read()
    assert b[:4] == bytes([0x76, 0x2f, 0x31, 0x01]) and b[4] == 2
    p = 8
    attrs = {}
    while b[p] != 0:
        e = b.index(0, p); name = b[p:e].decode(); p = e + 1
        e = b.index(0, p); typ = b[p:e].decode(); p = e + 1
        (size,) = struct.unpack_from("<i", b, p); p += 4
        attrs[name] = (typ, b[p:p + size]); p += size
    p += 1
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    assert attrs["compression"][1] == b"\0" and attrs["lineOrder"][1] == b"\0"
    names = re.findall(rb"([A-Z])\0\x02\0\0\0", attrs["channels"][1])
    assert names == [b"A", b"B", b"G", b"R"]
    offsets = struct.unpack_from(f"<{h}Q", b, p)
    img = np.zeros((h, w, 4), np.float32)
    for y in range(h):
        yy, nbytes = struct.unpack_from("<2i", b, offsets[y])
        assert yy == y and nbytes == 16 * w
        planes = np.frombuffer(b, "<f4", 4 * w, offsets[y] + 8).reshape(4, w)
        img[y, :, 3], img[y, :, 2], img[y, :, 1], img[y, :, 0] = planes       # A,B,G,R -> (r,g,b,w)
    (n,) = struct.unpack("<i", attrs["rtSampleCount"][1])
    return img, n


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline() == b"PF\n"
        w, h = map(int, f.readline().split())
        assert float(f.readline()) < 0
        return np.frombuffer(f.read(), "<f4").reshape(h, w, 3)[::-1]


def read_ppm(path):
    with open(path, "rb") as f:
        assert f.readline() == b"P6\n"
        w, h = map(int, f.readline().split())
        assert f.readline() == b"255\n"
        return np.frombuffer(f.read(), np.uint8).reshape(h, w, 3)


# ------------------------------------------------------------------------------------------------ CPU
def test_group_library_exports_every_declared_symbol(host):
    group_lib, _ = host
    src = open(os.path.join(ROOT, "include", "rt_b200_group.h")).read()
    names = sorted(set(re.findall(r"RT_API\s+[\w\s\*]+?\b(rt_group_\w+)\s*\(", src)))
    assert len(names) >= 13
    lib = C.CDLL(group_lib)
    for n in names:
        assert hasattr(lib, n), f"librt_b200_group.so does not export {n}"


def test_group_shard_matches_python(host):
    from cpuraytracer_b200.distributed import shard_samples
    lib = C.CDLL(host[0])
    lib.rt_group_shard.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
    for first, spp, world in [(1, 8, 2), (5, 7, 3), (1, 1, 8), (100, 4096, 8), (1, 0, 4)]:
        for rank in range(world):
            f, n = C.c_uint64(), C.c_int()
            assert lib.rt_group_shard(first, spp, rank, world, C.byref(f), C.byref(n)) == 0
            assert (f.value, n.value) == shard_samples(first, spp, rank, world)
    f, n = C.c_uint64(), C.c_int()
    assert lib.rt_group_shard(0, 4, 0, 2, C.byref(f), C.byref(n)) == 1      # first_sample >= 1
    assert lib.rt_group_shard(1, 4, 2, 2, C.byref(f), C.byref(n)) == 1


@pytest.mark.parametrize("name", ["d488", "s4096", "small"])
def test_cpp_init_scene_equals_python_recorder(host, rt, tmp_path, name):
    """InitScene written against rt_b200.hpp records exactly the flat scene cpuraytracer_b200.scenes records."""
    from cpuraytracer_b200 import _lib as L
    path = tmp_path / f"{name}.rtsc"
    run_cli(host[1], "--scene", name, "--dump-scene", path, "--spp", 0)
    (sp, ma, te, li), sky_lum, sky_tex = read_rtsc(path, L)
    flat = rt.scenes.NAMED[name]().export()
    assert sp.tobytes() == flat.spheres.tobytes()
    assert ma.tobytes() == flat.materials.tobytes()
    assert te.tobytes() == flat.textures.tobytes()
    assert li.tobytes() == flat.lights.tobytes()
    assert (sky_lum, sky_tex) == (flat.sky_luminance, flat.sky_texture)


def test_cpp_sinks(host, tmp_path):
    r = run_cli(host[1], "--selftest-sinks", tmp_path)
    assert "sinks ok" in r.stdout
    w, h, n = 7, 5, 3
    x, y = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    hdr = np.stack([0.25 * x, 0.5 * y, 0.125 * (x + y), np.full_like(x, n)], -1).astype(np.float32)
    img, cnt = read_exr(tmp_path / "t.exr")
    assert cnt == n and np.array_equal(img, hdr)
    assert np.array_equal(read_pfm(tmp_path / "t.pfm"), hdr[..., :3] / np.float32(n))
    ppm = read_ppm(tmp_path / "t.ppm")
    assert np.array_equal(ppm[..., 0], (x * 30).astype(np.uint8)) and np.array_equal(ppm[..., 1], (y * 50).astype(np.uint8))
    assert np.array_equal(ppm[..., 2], (x + y).astype(np.uint8))


def test_cli_rejects_bad_arguments(host, tmp_path):
    assert run_cli(host[1], "--bogus", check=False).returncode == 2
    assert run_cli(host[1], "--scene", "nope", "--dump-scene", tmp_path / "x", "--spp", 0, check=False).returncode == 2
    assert run_cli(host[1], "--width", -1, check=False).returncode == 2


def test_cli_has_no_cpu_fallback(host):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = run_cli(host[1], "--spp", 1, "--quiet", check=False)
    assert r.returncode == 2 and "no CPU fallback" in r.stderr


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_cli_render_equals_python_mirror(host, rt, tmp_path):
    w, h, spp = 256, 144, 4
    out = tmp_path / "img"
    r = run_cli(host[1], "--width", w, "--height", h, "--spp", spp, "--scene", "d488", "--out", out, "--format", "ppm,pfm,exr", "--quiet",
                "--primary-hits", tmp_path / "ids.bin", "--jitter-off", "--lens-off")
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    assert stats["paths"] == w * h * spp and stats["spp"] == spp
    flags = rt.RT_FLAG_JITTER_OFF | rt.RT_FLAG_LENS_OFF
    ren = rt.Renderer(0, w, h)
    ren.set_scene(rt.scenes.scene_d488())
    ren.set_camera(rt.scenes.default_camera(aspect=w / h))
    ids, mids = ren.primary_hits(1, flags)
    ren.render(spp, 50, flags)
    hdr = ren.read_hdr_sum()
    ldr = ren.resolve_ldr()
    img, cnt = read_exr(str(out) + ".exr")
    assert cnt == spp and np.array_equal(img, hdr)                      # same kernels, same bits
    assert np.array_equal(read_pfm(str(out) + ".pfm"), hdr[..., :3] / np.float32(spp))
    ppm = read_ppm(str(out) + ".ppm")
    assert np.array_equal(ppm[..., 0], (ldr >> 16) & 0xFF) and np.array_equal(ppm[..., 2], ldr & 0xFF)
    raw = np.fromfile(tmp_path / "ids.bin", np.int32).reshape(2, h, w)
    assert np.array_equal(raw[0], ids) and np.array_equal(raw[1], mids)
    assert stats["closest_rays"] == ren.stats().closest_rays


@pytest.mark.gpu
def test_cli_checkpoint_resume_is_exact(host, tmp_path):
    w, h = 160, 90
    base = ["--width", w, "--height", h, "--scene", "small", "--quiet", "--format", "exr"]
    run_cli(host[1], *base, "--spp", 6, "--out", tmp_path / "straight")
    run_cli(host[1], *base, "--spp", 2, "--out", tmp_path / "part", "--checkpoint", tmp_path / "ck.exr")
    run_cli(host[1], *base, "--spp", 4, "--out", tmp_path / "resumed", "--resume", tmp_path / "ck.exr", "--frames-per-call", 3)
    a, na = read_exr(tmp_path / "straight.exr")
    b, nb = read_exr(tmp_path / "resumed.exr")
    assert na == nb == 6 and np.array_equal(a, b)


@pytest.mark.gpu
def test_cli_two_gpus_equal_one(host, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    w, h, spp = 320, 180, 8
    base = ["--width", w, "--height", h, "--spp", spp, "--scene", "d488", "--quiet", "--format", "exr"]
    run_cli(host[1], *base, "--gpus", 1, "--out", tmp_path / "g1")
    r = run_cli(host[1], *base, "--gpus", 2, "--out", tmp_path / "g2", "--frames-per-call", 4)
    a, _ = read_exr(tmp_path / "g1.exr")
    b, nb = read_exr(tmp_path / "g2.exr")
    assert nb == spp and np.array_equal(a[..., 3], b[..., 3])           # every pixel got all 8 samples
    np.testing.assert_allclose(b[..., :3], a[..., :3], rtol=2e-6, atol=1e-9)   # reduction order differs, nothing else
    assert json.loads(r.stdout.strip().splitlines()[-1])["gpus"] == 2


def test_ranlux24_base_known_answer(rt):
    """[rand.predef]: the 10000th consecutive invocation of a default-constructed ranlux24_base produces 7937952."""
    g = rt.scenes.Ranlux24Base()
    for _ in range(9999):
        g.next_u24()
    assert g.next_u24() == 7937952


@pytest.mark.parametrize("seed", [1, 123456789, 4000000000])
def test_cpp_ranlux_scene_equals_python_replay(host, rt, tmp_path, seed):
    """--ranlux-seed drives InitScene with std::ranlux24_base from <random> (the reference's generator, spheres-app.cpp:54);
    the Python replay of the engine must record the same scene byte for byte."""
    from cpuraytracer_b200 import _lib as L
    path = tmp_path / "ranlux.rtsc"
    run_cli(host[1], "--ranlux-seed", seed, "--dump-scene", path, "--spp", 0)
    (sp, ma, te, li), sky_lum, sky_tex = read_rtsc(path, L)
    flat = rt.scenes.scene_ranlux(seed).export()
    assert len(sp) == 488
    assert sp.tobytes() == flat.spheres.tobytes()
    assert ma.tobytes() == flat.materials.tobytes()
    assert te.tobytes() == flat.textures.tobytes()


@pytest.mark.gpu
def test_cli_accel_flag_renders_the_same_bits(host, tmp_path):
    """apps/spheres --accel (RT_FLAG_ACCEL through rt_group_render) == the default brute-force query, byte for byte."""
    w, h = 200, 120
    base = ["--width", w, "--height", h, "--scene", "d488", "--quiet", "--format", "exr", "--spp", 5]
    run_cli(host[1], *base, "--out", tmp_path / "brute")
    r = run_cli(host[1], *base, "--out", tmp_path / "accel", "--accel")
    a, na = read_exr(tmp_path / "brute.exr")
    b, nb = read_exr(tmp_path / "accel.exr")
    assert na == nb == 5 and np.array_equal(a, b)
    assert json.loads(r.stdout.strip().splitlines()[-1])["paths"] == w * h * 5
