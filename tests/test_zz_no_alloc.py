"""rt_render_static (include/rt_b200.h): the render entry point that allocates nothing -- its scratch memory (per-call Halton
tables, first-hit records) is a file-scope static __device__ array.  No counterpart in the reference.
Checked through the C ABI against the existing entry point rt_render on the same inputs, two calls in a row.
Collected last on purpose (file name): a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 2e-4     # the GPU parity tests' bound (tests/test_gpu_parity.py); the two entry points are expected to agree far better


@pytest.fixture(scope="module")
def gpu(rt):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    return rt


def _close(g, o):
    g = g[..., :3].astype(np.float64)
    o = o[..., :3].astype(np.float64)
    den = np.maximum(np.abs(o), 1e-3 * o.mean())
    return float(np.max(np.abs(g - o) / den))


def _pair(rt, scene, w, h):
    cam = rt.scenes.default_camera(aspect=w / h)
    out = []
    for _ in range(2):
        r = rt.Renderer(device=0, width=w, height=h)
        r.set_scene(scene)
        r.set_camera(cam)
        out.append(r)
    out[0].set_chunk_spp(1 << 20)        # rt_render: one work item per pixel and batch, i.e. the same sequential per-pixel sums
    return out


@pytest.mark.parametrize("name,w,h,depth,flags", [("d488", 200, 117, 50, 0), ("d488", 160, 90, 0, 0), ("s4096", 96, 54, 6, 8)])
def test_static_entry_point_equals_rt_render_on_two_calls_in_a_row(gpu, name, w, h, depth, flags):
    rt = gpu
    scene = rt.scenes.NAMED[name]()
    ref, sta = _pair(rt, scene, w, h)
    lib = rt.load()
    assert lib.rt_render_static_scratch_bytes(w, h, 40) <= rt._lib.RT_STATIC_SCRATCH_BYTES
    total = 0
    for spp in (21, 40):                 # two calls in a row: 21 frames (a ragged last batch), then 40 more continuing the sum
        ref.render(spp, depth, flags)
        sta.reset_stats()
        # straight through the C ABI: int rt_render_static(ctx, first_sample, spp, max_depth, flags)
        rc = lib.rt_render_static(sta._h, C.c_uint64(total + 1), spp, depth, flags)
        assert rc == 0, lib.rt_last_error().decode()
        sta.sample_count += spp
        total += spp
        a, b = ref.read_hdr_sum(), sta.read_hdr_sum()
        assert np.all(b[..., 3] == total) and np.array_equal(a[..., 3], b[..., 3])
        assert _close(b, a) < RTOL
        st = sta.stats()
        assert st.paths == spp * w * h
    # a different, smaller job right after on the same device reuses the array (stale contents of the larger job must not matter)
    ref2, sta2 = _pair(rt, scene, 64, 36)
    ref2.render(5, depth, flags)
    sta2.render_static(5, depth, flags)
    assert _close(sta2.read_hdr_sum(), ref2.read_hdr_sum()) < RTOL
    # RT_FLAG_OVERWRITE starts the sum without reading the buffer
    sta2.write_hdr_sum(np.full((36, 64, 4), 9.0, np.float32), 1)
    sta2.render_static(5, depth, flags | rt.RT_FLAG_OVERWRITE, first_sample=1)
    assert _close(sta2.read_hdr_sum(), ref2.read_hdr_sum()) < RTOL
    for r in (ref, sta, ref2, sta2):
        r.close()


def test_static_entry_point_rejects_what_does_not_fit(gpu):
    rt = gpu
    lib = rt.load()
    assert lib.rt_render_static_scratch_bytes(1920, 1080, 256) <= rt._lib.RT_STATIC_SCRATCH_BYTES < lib.rt_render_static_scratch_bytes(3840, 2160, 1)
    assert lib.rt_render_static_scratch_bytes(0, 10, 1) == 0
    r = rt.Renderer(device=0, width=3840, height=2160)
    r.set_scene(rt.scenes.scene_small())
    r.set_camera(rt.scenes.default_camera(aspect=16 / 9))
    with pytest.raises(rt.RtError) as e:
        r.render_static(1, 2)
    assert e.value.code == 5 and "scratch" in str(e.value)          # RT_ERR_UNSUPPORTED, with the figures
    r.resize(64, 36)
    with pytest.raises(rt.RtError) as e:
        r.render_static(5_000_000, 2)                               # the tables alone exceed the array
    assert e.value.code == 5
    for bad in (rt.RT_FLAG_HALTON, rt.RT_FLAG_WAVEFRONT, rt.RT_FLAG_DEBUG_ONE_KERNEL):
        with pytest.raises(rt.RtError) as e:
            r.render_static(1, 2, bad)
        assert e.value.code == 1
    r.render_static(3, 2)
    assert np.all(r.read_hdr_sum()[..., 3] == 3)
    r.close()
