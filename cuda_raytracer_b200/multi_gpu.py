"""Sample-split rendering across the GPUs of one box: one process per GPU, scene replicated, passes partitioned,
accumulation buffers summed with one reduce to rank 0 (the reference itself is single-GPU).

A pass (<= 20 samples per pixel, raytracing.cu:222-254) is the independent unit: its seeds depend only on the number of
samples still to cast after it (raytracing.cu:229,236), so pass k produces the same rays on any GPU count and the
N-GPU image equals the 1-GPU image up to float summation order. `torch.distributed` is plumbing only (NCCL on GPUs,
gloo in the CPU tests); nothing here computes on the CPU.
"""
import numpy as np

PASS_SPP = 20  # MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109


def pass_count(ray_count, pass_spp=PASS_SPP):
    return (ray_count + pass_spp - 1) // pass_spp


def pass_range(total_passes, rank, world):
    """Contiguous block [first, first + count) of passes for `rank`; the first `total % world` ranks get one more."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    base, extra = divmod(total_passes, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


class DeviceFramebuffer:
    """Zero-copy view of the core's device framebuffer for torch (so NCCL reduces it in place)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def reduce_to_root(tensor, dist, dst=0):
    """Sum the per-rank accumulation buffers into rank `dst` (in place). `dist` = torch.distributed or None."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(tensor, dst=dst, op=dist.ReduceOp.SUM)
    return tensor


def render_distributed(renderer, scene, dist, rank, world, device_index, sort=True, pass_spp=0, out=None, **render_kw):
    """Render this rank's share of the frame on its GPU and reduce to rank 0.

    Returns (sums float32[H, W, 3] on rank 0 else None, rt_stats of this rank). The scene must already be uploaded.
    Ranks beyond the number of passes contribute zeros.
    """
    import torch

    spp = pass_spp if pass_spp > 0 else PASS_SPP
    total = pass_count(scene.desc().ray_count, spp)
    first, count = pass_range(total, rank, world)
    n = renderer.width * renderer.height * 3
    ptr = renderer.render_device(sort=sort, pass_spp=pass_spp, pass_first=first if count else 0, pass_count=max(count, 1), **render_kw)
    stats = renderer.last_stats
    fb = torch.as_tensor(DeviceFramebuffer(ptr, n), device=f"cuda:{device_index}")
    if count == 0:
        fb.zero_()
    reduce_to_root(fb, dist if world > 1 else None, 0)
    if rank != 0:
        return None, stats
    if out is None:
        out = np.empty((renderer.height, renderer.width, 3), np.float32)
    torch.from_numpy(out.reshape(-1)).copy_(fb)
    return out, stats
