"""CPU tests of the multi-GPU host logic with world_size-2 gloo: pass partitioning + reduce to rank 0.

Each rank renders its block of passes with the ORACLE (the checker standing in for a GPU, CPU box only) and the
package's own `pass_range` / `reduce_to_root` combine them; rank 0 compares with the whole-frame render. On the GPU
box the same functions run over NCCL (bench.py, test_gpu_parity.test_pass_split_sums_to_whole)."""
import os
import socket
import sys

import numpy as np
import pytest

from cuda_raytracer_b200 import multi_gpu as mg


def test_pass_range_partitions_every_pass_once():
    for total in (0, 1, 5, 32, 50, 33):
        for world in (1, 2, 3, 4, 8, 64):
            spans = [mg.pass_range(total, r, world) for r in range(world)]
            covered = [p for f, c in spans for p in range(f, f + c)]
            assert covered == list(range(total))
            counts = [c for _, c in spans]
            assert max(counts) - min(counts) <= 1
    assert mg.pass_count(640) == 32 and mg.pass_count(1000) == 50 and mg.pass_count(21) == 2 and mg.pass_count(0) == 0
    with pytest.raises(ValueError):
        mg.pass_range(4, 2, 2)


def test_ray_range_partitions_every_ray_once():
    for total in (0, 1, 19, 1000 * 1000 * 1000, 3840 * 2160 * 640, 12345):
        for world in (1, 2, 3, 8, 64):
            for align in (1, 20, 640):
                spans = [mg.ray_range(total, r, world, align) for r in range(world)]
                pos = 0
                for f, c in spans:
                    assert c >= 0 and (c == 0 or f == pos)
                    pos += c
                assert pos == total
                counts = [c for _, c in spans]
                assert max(counts) - min(counts) <= (1 if align == 1 else 2 * align - 1)  # the last unit may be partial
    # BASELINE configs[3]: teapot as shipped = 50 passes of 20 M rays on 8 GPUs -> 6.25 passes per rank
    assert [mg.ray_range(10 ** 9, r, 8)[1] for r in range(8)] == [125_000_000] * 8
    with pytest.raises(ValueError):
        mg.ray_range(10, 3, 3)


def test_sample_range_partitions_every_sample_once():
    for spp in (0, 1, 7, 20, 640, 1000, 1001):
        for world in (1, 2, 3, 8, 64):
            spans = [mg.sample_range(spp, r, world) for r in range(world)]
            assert [s for f, c in spans for s in range(f, f + c)] == list(range(spp))
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    assert [mg.sample_range(1000, r, 8) for r in range(8)] == [(125 * r, 125) for r in range(8)]  # teapot as shipped on 8 GPUs


def _worker(rank, world, port, out_path):
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (root, os.path.join(root, "oracle"), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    from util import load, oracle_arrays, oracle_render, oracle_render_samples, oracle_render_select

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        scene = load("cornell_plus").set_image(24, 20, 70, 4, 1.0)  # 4 passes: 20, 20, 20, 10
        A = oracle_arrays(scene)
        total = mg.pass_count(A.ray_count)
        first, count = mg.pass_range(total, rank, world)
        part = oracle_render(A, 1, first, count)[0] if count else np.zeros((A.height, A.width, 3), np.float32)
        t = torch.from_numpy(part.reshape(-1).copy())
        mg.reduce_to_root(t, dist, 0)
        # the equal-share split of the ray sequence (4 passes on 3 ranks cut passes in the middle)
        rf, rc = mg.ray_range(A.width * A.height * A.ray_count, rank, world)
        part2, ctr = oracle_render_select(A, 1, ray_first=rf, ray_count=rc)
        t2 = torch.from_numpy(part2.reshape(-1).copy())
        # the product's split: an equal share of every pixel's samples (70 samples on 3 ranks: 24 / 23 / 23)
        sf, sc = mg.sample_range(A.ray_count, rank, world)
        part3, ctr3 = oracle_render_samples(A, sf, sc)
        t3 = torch.from_numpy(part3.reshape(-1).copy())
        live = torch.tensor([ctr.live_ray_bounces, ctr3.live_ray_bounces], dtype=torch.int64)
        mg.reduce_to_root(t2, dist, 0)
        mg.reduce_to_root(t3, dist, 0)
        dist.reduce(live, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            whole, cw = oracle_render(A, 1)
            np.savez(out_path, combined=t.numpy().reshape(whole.shape), combined_rays=t2.numpy().reshape(whole.shape),
                     combined_samples=t3.numpy().reshape(whole.shape), whole=whole,
                     live=np.array([int(live[0]), int(live[1]), cw.live_ray_bounces]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_split_equals_whole_frame(tmp_path, world):
    import torch.multiprocessing as mp

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "r0.npz")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    z = np.load(out)
    assert np.allclose(z["combined"], z["whole"], rtol=1e-5, atol=1e-5 * float(z["whole"].mean()))
    assert np.allclose(z["combined_rays"], z["whole"], rtol=1e-5, atol=1e-5 * float(z["whole"].mean()))
    assert np.allclose(z["combined_samples"], z["whole"], rtol=1e-5, atol=1e-5 * float(z["whole"].mean()))
    assert z["live"][0] == z["live"][2] and z["live"][1] == z["live"][2]
    assert z["whole"].mean() > 0
