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


def _worker(rank, world, port, out_path):
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (root, os.path.join(root, "oracle"), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    from util import load, oracle_arrays, oracle_render

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        scene = load("cornell_plus").set_image(24, 20, 70, 4, 1.0)  # 4 passes: 20, 20, 20, 10
        A = oracle_arrays(scene)
        total = mg.pass_count(A.ray_count)
        first, count = mg.pass_range(total, rank, world)
        part = oracle_render(A, 1, first, count)[0] if count else np.zeros((A.height, A.width, 3), np.float32)
        t = torch.from_numpy(part.reshape(-1).copy())
        mg.reduce_to_root(t, dist, 0)
        if rank == 0:
            whole = oracle_render(A, 1)[0]
            np.savez(out_path, combined=t.numpy().reshape(whole.shape), whole=whole)
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
    assert z["whole"].mean() > 0
