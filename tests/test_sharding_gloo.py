"""Multi-rank host logic on CPU: world_size-2 (and 3) gloo process groups exercise cpuraytracer_b200.distributed --
disjoint sample ranges per rank + one sum-reduce -- with the CPU oracle standing in for the per-rank renderer
(the CUDA path itself is covered by the `gpu` tests; bench.py --gpus N runs the same code over NCCL)."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_shard_samples_partition():
    from cpuraytracer_b200.distributed import shard_samples
    for first in (1, 17):
        for spp in (0, 1, 7, 8, 1000, 4096):
            for world in (1, 2, 3, 8):
                got = [shard_samples(first, spp, r, world) for r in range(world)]
                assert sum(n for _, n in got) == spp
                nxt = first
                for start, n in got:
                    assert start == nxt and n >= 0
                    nxt += n
                assert max(n for _, n in got) - min(n for _, n in got) <= 1
    with pytest.raises(ValueError):
        shard_samples(0, 4, 0, 2)
    with pytest.raises(ValueError):
        shard_samples(1, 4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, first, spp, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import cpuraytracer_b200 as rt
    from cpuraytracer_b200.distributed import render_sharded_with
    from oracle import orc
    dist.init_process_group("gloo", rank=rank, world_size=world)
    scene = rt.scenes.scene_small()
    osc = orc.Scene.from_flat(scene.export())
    cam = rt.scenes.default_camera(aspect=w / h)
    cp = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)

    def render_fn(my_first, my_spp):
        hdr = np.zeros((h, w, 4), np.float32)
        if my_spp > 0:
            orc.render(osc, cp, orc.config(w, h, spp=my_spp, first_sample=my_first, threads=1), hdr)
        return torch.from_numpy(hdr)

    total = render_sharded_with(render_fn, first, spp)
    if rank == 0:
        np.save(out_path, total.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _worker_multicall(rank, world, port, w, h, calls, out_path, overwrite_capable=False):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    import cpuraytracer_b200 as rt
    from cpuraytracer_b200.distributed import ShardedAccumulator
    from oracle import orc
    dist.init_process_group("gloo", rank=rank, world_size=world)
    scene = rt.scenes.scene_small()
    osc = orc.Scene.from_flat(scene.export())
    cam = rt.scenes.default_camera(aspect=w / h)
    cp = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)
    hdr_np = np.zeros((h, w, 4), np.float32)
    hdr = torch.from_numpy(hdr_np)                       # shares memory: the oracle adds into the tensor gloo reduces

    def render_into(my_first, my_spp, overwrite=False):
        if overwrite:                                    # what RT_FLAG_OVERWRITE does on the device: the call starts the sum
            hdr_np[...] = 0.0
        elif overwrite_capable and rank != 0:
            raise AssertionError("ranks >= 1 must be asked to overwrite")
        orc.render(osc, cp, orc.config(w, h, spp=my_spp, first_sample=my_first, threads=1), hdr_np)

    if overwrite_capable and rank != 0:
        hdr_np[...] = 123.0                              # stale content the protocol must never add to rank 0
    acc = ShardedAccumulator(hdr, render_into, overwrite_capable=overwrite_capable)
    for spp in calls:
        acc.render(spp)
    if rank == 0:
        np.save(out_path, hdr_np)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("overwrite_capable,calls", [(False, (3, 2, 4)), (True, (3, 2, 4)), (True, (1, 1, 3))])
def test_consecutive_sharded_calls_accumulate_once(orc, rt, tmp_path, overwrite_capable, calls):
    """bench.py's loop: several render+reduce calls in a row.  Rank 0 must end with every frame exactly once -- with a renderer
    that is cleared before each call on ranks >= 1, and with one that overwrites (RT_FLAG_OVERWRITE; a rank whose share of a
    call is empty -- 1 frame over 2 ranks -- must still contribute zeros, not its stale buffer)."""
    import torch.multiprocessing as mp
    w, h = 40, 24
    out = str(tmp_path / "sum.npy")
    mp.spawn(_worker_multicall, args=(2, _free_port(), w, h, calls, out, overwrite_capable), nprocs=2, join=True)
    got = np.load(out)
    scene = rt.scenes.scene_small()
    osc = orc.Scene.from_flat(scene.export())
    cam = rt.scenes.default_camera(aspect=w / h)
    cp = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)
    ref, _ = orc.render(osc, cp, orc.config(w, h, spp=sum(calls), first_sample=1, threads=1))
    assert np.all(got[..., 3] == sum(calls))
    assert np.allclose(got, ref, rtol=2e-6, atol=1e-9)


@pytest.mark.parametrize("world,spp", [(2, 6), (3, 7), (2, 1)])
def test_sample_split_plus_reduce_equals_single_rank(orc, rt, tmp_path, world, spp):
    import torch.multiprocessing as mp
    w, h, first = 48, 27, 5
    out = str(tmp_path / "sum.npy")
    mp.spawn(_worker, args=(world, _free_port(), w, h, first, spp, out), nprocs=world, join=True)
    got = np.load(out)
    scene = rt.scenes.scene_small()
    osc = orc.Scene.from_flat(scene.export())
    cam = rt.scenes.default_camera(aspect=w / h)
    cp = orc.camera_params(cam.origin, cam.look_at, cam.vertical_fov, cam.aspect_ratio, cam.focal_length, cam.aperture)
    ref, _ = orc.render(osc, cp, orc.config(w, h, spp=spp, first_sample=first, threads=1))
    assert np.all(got[..., 3] == spp)                         # every pixel received every frame exactly once
    # same sample set, different float summation order: ~1e-6 relative, not bitwise
    assert np.allclose(got, ref, rtol=2e-6, atol=1e-9)
