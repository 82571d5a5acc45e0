"""Multi-GPU rendering: one process per GPU, samples-per-pixel sharding, one reduce of the film.

The reference's own `Naive` renderer splits spp across its worker threads (renderer.h:352-360) and adds the
per-thread results; this is the same partition across GPUs.  Samples are i.i.d. and the Philox stream is keyed on
the GLOBAL sample index, so rank r of R renders global samples [offset_r, offset_r + n_r) of every pixel and the
sum over ranks is the same estimator for any R (up to fp32 summation order).  No collective on the data path; the
only exchange is ONE sum-reduce of the accumulation buffer (H x W float4) to rank 0 -- NCCL over NVLink on GPUs
(`torch.distributed`, in place on the library's device buffer), gloo in the CPU tests of this logic.
"""
import ctypes as C

import numpy as np


def shard_samples(total_spp, world_size, rank):
    """(first global sample, count) of `rank`: contiguous, sizes differ by at most one, no gaps, no overlap."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world size")
    base, extra = divmod(int(total_spp), world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


class _DevicePtr:
    """Exposes a raw device allocation through __cuda_array_interface__ so torch can wrap it without a copy."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def film_tensor(renderer):
    """The renderer's accumulation buffer (H x W x 4 fp32 sums on its GPU) as a torch tensor sharing the memory."""
    import torch
    ptr, nbytes = renderer.device_buffer()
    t = torch.as_tensor(_DevicePtr(ptr, nbytes // 4), device=f"cuda:{renderer.scene.device}")
    return t.view(renderer.config.height, renderer.config.width, 4)


def reduce_film(film, dst=0, group=None):
    """Sum the per-rank films into rank `dst` (in place).  Works on CUDA tensors (NCCL) and CPU tensors (gloo)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def reduce_counts(values, device, dst=0, group=None):
    """Sum int64 counters (paths, rays by type) into rank `dst`."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([int(v) for v in values], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return [int(x) for x in t.cpu().tolist()]


def render_sharded(scene, config, rank=0, world_size=1, group=None):
    """Render `config.samples` spp of the whole film split over `world_size` ranks.  Returns (renderer, report);
    on rank 0 the renderer's film holds the reduced sums of ALL samples (so tonemap / write_ppm / framebuffer give
    the final image), on other ranks their own share."""
    import torch
    from .renderer import renderer_from_config
    if config.samples < world_size:  # checked before sharding, so that EVERY rank raises (none is left in the reduce)
        raise ValueError("fewer samples than ranks: use fewer GPUs")
    first, count = shard_samples(config.samples, world_size, rank)
    cfg = config.replace(samples=count, sample_offset=config.sample_offset + first)
    r = renderer_from_config(scene, cfg).start_render()
    report = r.finalize()
    if world_size > 1:
        film = film_tensor(r)
        torch.cuda.current_stream(film.device).synchronize()
        reduce_film(film, 0, group)
        torch.cuda.synchronize(film.device)
        keys = ("camera_rays", "rays", "rays_extension", "rays_shadow", "rays_reference")
        tot = reduce_counts([report[k] for k in keys], film.device, 0, group)
        if rank == 0:
            report = dict(report, **dict(zip(keys, tot)))
            r.spp_total = config.samples
    return r, report
