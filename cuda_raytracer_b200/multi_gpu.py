"""Sample-split rendering across the GPUs of one box: one process per GPU, scene replicated, the frame's ray sequence
partitioned, accumulation buffers summed with one reduce to rank 0 (the reference itself is single-GPU).

A pass (<= 20 samples per pixel, raytracing.cu:222-254) is an independent unit: its seeds depend only on the number of
samples still to cast after it (raytracing.cu:229,236) and, per ray, on the ray's index inside the pass (raytracing.cu:89,
scene.cu:81). So a pass can be cut at any ray: every rank takes an equal share of every pixel's samples (`sample_range`:
whole passes plus, when the sample count does not divide, a few samples of every pixel of one more pass; `ray_range` and
`pass_range` are the contiguous alternatives), ray k is the same ray on any GPU count
and the N-GPU image equals the 1-GPU image up to float summation order. `torch.distributed` is plumbing only (NCCL on
GPUs, gloo in the CPU tests); nothing here computes on the CPU.
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


def ray_range(total_rays, rank, world, align=1):
    """Contiguous share [first, first + count) of the frame's ray sequence (pass 0's rays, then pass 1's, ...) for `rank`:
    equal up to `align` rays (rt_render_opts.ray_first / ray_count). 50 passes on 8 GPUs = 6.25 passes each, where whole
    passes would give 7 / 6 (89 % at best)."""
    if world <= 0 or not (0 <= rank < world) or align <= 0:
        raise ValueError("bad rank / world size")
    units = (total_rays + align - 1) // align
    base, extra = divmod(units, world)
    first = rank * base + min(rank, extra)
    count = base + (1 if rank < extra else 0)
    lo, hi = min(total_rays, first * align), min(total_rays, (first + count) * align)
    return lo, hi - lo


def sample_range(ray_count, rank, world):
    """Share [first, first + count) of every pixel's `ray_count` samples for `rank` (rt_render_opts.sample_first / sample_count).
    Every share covers the whole image, so equal shares cost the same whatever the picture shows; a contiguous range of the ray
    sequence (`ray_range`) covers whole passes plus a band of image rows, which is sky for one rank and geometry for another
    (teapot as shipped on 8 GPUs: 83 % efficiency with ray ranges)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    base, extra = divmod(ray_count, world)
    return rank * base + min(rank, extra), base + (1 if rank < extra else 0)


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
    Ranks beyond the number of rays contribute zeros.
    """
    import torch

    spp = scene.desc().ray_count
    first, count = sample_range(spp, rank, world)
    n = renderer.width * renderer.height * 3
    if world == 1:
        ptr = renderer.render_device(sort=sort, pass_spp=pass_spp, **render_kw)
    else:  # an empty share renders nothing (a range past the end) and leaves a zeroed framebuffer
        ptr = renderer.render_device(sort=sort, pass_spp=pass_spp, sample_first=first if count else spp, sample_count=max(count, 1), **render_kw)
    stats = renderer.last_stats
    fb = torch.as_tensor(DeviceFramebuffer(ptr, n), device=f"cuda:{device_index}")
    reduce_to_root(fb, dist if world > 1 else None, 0)
    # The reduce runs on torch's stream and reads the core's framebuffer in place; the core's next render clears and
    # accumulates into that buffer on ITS OWN stream. Wait for the reduce on every rank before handing control back.
    torch.cuda.current_stream().synchronize()
    if rank != 0:
        return None, stats
    if out is None:
        out = np.empty((renderer.height, renderer.width, 3), np.float32)
    torch.from_numpy(out.reshape(-1)).copy_(fb)
    return out, stats
