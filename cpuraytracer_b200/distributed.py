"""Multi-GPU sharding of the render loop: one process per GPU, disjoint sample ranges, one sum-reduce.

The path shards naturally: frame jitter and lens samples are pure functions of the frame
index n (spheres-app.cpp:140,152) and the counter RNG is keyed on (pixel, n, bounce), so rank g of G renders
frames [first + g*spp/G, first + (g+1)*spp/G) for ALL pixels and the union over ranks is exactly the
single-GPU sample set.  The only exchange is one sum-reduce of the W*H float4 running-sum buffers to rank 0
(ncclReduce over NVLink/NVSwitch via torch.distributed; gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple


def shard_samples(first_sample: int, spp: int, rank: int, world: int) -> Tuple[int, int]:
    """(first_sample, spp) of `rank`: contiguous, disjoint, covering [first_sample, first_sample + spp),
    sizes differ by at most one (the first spp % world ranks take the extra frame)."""
    if spp < 0 or world < 1 or not (0 <= rank < world) or first_sample < 1:
        raise ValueError("shard_samples: bad argument")
    q, r = divmod(spp, world)
    mine = q + (1 if rank < r else 0)
    start = first_sample + rank * q + min(rank, r)
    return start, mine


def reduce_hdr(hdr_tensor, dst: int = 0, group=None):
    """Sum-reduce the per-rank running-sum buffers onto `dst` (xyz = radiance sums, w = sample counts).
    The summation order differs from a single-GPU render, so results agree to ~1e-6 relative, not bitwise."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(hdr_tensor, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return hdr_tensor


class ShardedAccumulator:
    """The multi-call protocol of a sharded render, independent of who renders (CUDA library, or the oracle in the gloo
    tests): rank 0's buffer is the job's running sum (m_backbufferHdr); every other rank's buffer holds only the current
    call's share, so the reduce adds only this call's frames to rank 0.  `render_into(first, n, overwrite)`: with
    overwrite=True the renderer STARTS the sum (RT_FLAG_OVERWRITE: no clear, no read of the old content); a renderer that
    cannot overwrite gets a zeroed buffer instead (overwrite_capable=False)."""

    def __init__(self, hdr, render_into: Callable[..., None], group=None, overwrite_capable: bool = False):
        import torch.distributed as dist
        self.hdr, self.render_into, self.group, self.overwrite_capable = hdr, render_into, group, overwrite_capable
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.sample_count = 0

    def clear(self) -> None:
        self.hdr.zero_()
        self.sample_count = 0

    def render(self, spp: int, first_sample: Optional[int] = None, reduce: bool = True):
        first = self.sample_count + 1 if first_sample is None else first_sample
        my_first, my_spp = shard_samples(first, spp, self.rank, self.world)
        fresh = reduce and self.rank != 0 and self.world > 1
        if fresh and not (self.overwrite_capable and my_spp > 0):
            self.hdr.zero_()
        if my_spp > 0:
            if self.overwrite_capable:
                self.render_into(my_first, my_spp, fresh)
            else:
                self.render_into(my_first, my_spp)
        if reduce:
            reduce_hdr(self.hdr, 0, self.group)
        self.sample_count += spp
        return self.hdr


class ShardedRenderer:
    """One rank's share of a multi-GPU render.  The running-sum buffer is a torch CUDA tensor bound into the
    rt_context (rt_hdr_bind_external), so NCCL reduces the very buffer the kernel accumulated into; the kernels
    run on torch's current stream, which orders them before the collective without a host sync."""

    def __init__(self, width: int, height: int, device: Optional[int] = None, group=None):
        import torch
        from .api import Renderer
        self.device = torch.cuda.current_device() if device is None else device
        with torch.cuda.device(self.device):
            self.hdr = torch.zeros((height, width, 4), dtype=torch.float32, device=f"cuda:{self.device}")
            stream = torch.cuda.current_stream(self.device).cuda_stream
        self.renderer = Renderer(device=self.device, width=width, height=height, stream=stream)
        self.renderer.bind_external_hdr(self.hdr.data_ptr())
        self.width, self.height = width, height
        self._depth, self._flags = 50, 0
        self._acc = ShardedAccumulator(self.hdr, self._render_into, group, overwrite_capable=True)
        self.rank, self.world = self._acc.rank, self._acc.world

    def _render_into(self, first: int, n: int, overwrite: bool) -> None:
        from ._lib import RT_FLAG_DEBUG_ONE_KERNEL, RT_FLAG_OVERWRITE, RT_FLAG_WAVEFRONT
        flags = self._flags
        if overwrite and (flags & (RT_FLAG_WAVEFRONT | RT_FLAG_DEBUG_ONE_KERNEL)):
            self.hdr.zero_()                                  # those schedules only add to the running sum
        elif overwrite:
            flags |= RT_FLAG_OVERWRITE
        self.renderer.render(n, self._depth, flags, first_sample=first)

    @property
    def sample_count(self) -> int:
        return self._acc.sample_count

    def set_scene(self, scene) -> None:
        self.renderer.set_scene(scene)

    def set_camera(self, camera) -> None:
        self.renderer.set_camera(camera)

    def clear(self) -> None:
        self._acc.clear()

    def render(self, spp: int, max_depth: int = 50, flags: int = 0, first_sample: Optional[int] = None, reduce: bool = True):
        """Renders this rank's share of frames [first, first + spp) and (optionally) reduces onto rank 0.
        After a reduce, rank 0's buffer holds the whole job's sum; the other ranks' buffers hold their last share."""
        self._depth, self._flags = max_depth, flags
        return self._acc.render(spp, first_sample, reduce)


def render_sharded_with(render_fn: Callable[[int, int], "object"], first_sample: int, spp: int, group=None):
    """Backend-agnostic form used by the CPU (gloo) tests: render_fn(first, spp) -> torch tensor running sum of this
    rank's share; returns the reduced tensor (valid on rank 0)."""
    import torch.distributed as dist
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    my_first, my_spp = shard_samples(first_sample, spp, rank, world)
    t = render_fn(my_first, my_spp)
    return reduce_hdr(t, 0, group)
