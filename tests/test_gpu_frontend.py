"""GPU tests of the whole program: the drop-in command line and north_star's second criterion — rendered images at
matched spp within a stated per-pixel RMSE of the REFERENCE's GPU output (the reference program built for sm_100a in
oracle/_ref, run on this box).

Bit parity of images is impossible by design: with sorting the reference seeds each ray from its SORTED slot through a
degenerate key. The bound is therefore calibrated in the test itself: the RMSE between two
reference runs with different random streams (`sort` vs `no_sort`) is the Monte-Carlo noise floor at that spp, and the
drop-in must be within 1.5x of it on the tone-mapped 8-bit image (plus a mean-brightness check of 1 %)."""
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

from util import ROOT, SCENES, read_png_rgb8

pytestmark = pytest.mark.gpu
EXE = os.path.join(ROOT, "cuda_raytracer_b200", "bin", "raytracing")
REF = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref")


def _scene_copy(tmp_path, name, w, h, spp, bounces):
    """The scene with its `image` line replaced; assets are reached through symlinks (paths are CWD-relative)."""
    for sub in ("teapot", "lamp"):
        if not os.path.exists(tmp_path / sub):
            os.symlink(os.path.join(SCENES, sub), tmp_path / sub)
    src = open(os.path.join(SCENES, name + ".scene")).read()
    exposure = re.search(r"(?m)^image\s+\S+\s+\S+\s+\S+\s+\S+\s+(\S+)", src).group(1)
    (tmp_path / (name + ".scene")).write_text(re.sub(r"(?m)^image .*$", f"image {w} {h} {spp} {bounces} {exposure}", src))
    return name + ".scene"


def _run(exe, tmp_path, scene, *flags, png="raytracing.png"):
    r = subprocess.run([exe, scene, *flags], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    img = read_png_rgb8(str(tmp_path / png))
    os.remove(tmp_path / png)
    return img, r.stdout


def test_command_line_writes_the_reference_outputs(tmp_path):
    scene = _scene_copy(tmp_path, "teapot", 120, 90, 40, 6)
    img, out = _run(EXE, tmp_path, scene)
    assert img.shape == (90, 120, 3) and img.mean() > 5
    for line in ("Loaded environment map with size 1024,1024", "Triangle count: 126050", "Node count: 81311", "BVH Took", "GPU Took"):
        assert line in out, out
    img2, out2 = _run(EXE, tmp_path, scene, "no_sort", "no_bvh")
    assert "Node count: 1\n" in out2
    diff = np.sqrt(np.mean((img.astype(np.float64) - img2.astype(np.float64)) ** 2))
    assert diff < 6.0  # same streams, same geometry: only bloom-threshold flips and float order differ
    img4, out4 = _run(EXE, tmp_path, scene, "host_bvh")  # tree built on the host as the reference does (the default builds it
    assert "Node count: 81311" in out4 and "BVH Took" in out4  # on the device): same node count, same picture
    assert np.sqrt(np.mean((img.astype(np.float64) - img4.astype(np.float64)) ** 2)) < 6.0
    img3, out3 = _run(EXE, tmp_path, scene, "stats", "out=custom.png", png="custom.png")
    assert img3.shape == img.shape and '"mrays_per_s"' in out3 and not os.path.exists(tmp_path / "raytracing.png")


def _gpu_count():
    import cuda_raytracer_b200 as rt
    return rt.device_count()


@pytest.mark.parametrize("gpus", [2, 4, 8])
def test_gpus_word_writes_the_same_png(tmp_path, gpus):
    """`raytracing <scene> gpus=N` (rt_multi_*: ray-sequence split + one ncclReduce into device 0) writes the picture of
    `gpus=1`: the rays are the same rays, only the float summation order differs (8-bit RMSE <= 0.5)."""
    if _gpu_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")
    scene = _scene_copy(tmp_path, "lamp", 320, 180, 70, 6)  # 4 passes (20, 20, 20, 10): no GPU count divides the work evenly
    one, out1 = _run(EXE, tmp_path, scene, "no_sort", "stats")
    many, outn = _run(EXE, tmp_path, scene, "no_sort", "stats", f"gpus={gpus}")
    assert f'"gpus": {gpus}' in outn
    live = [int(re.search(r'"live_ray_bounces": (\d+)', o).group(1)) for o in (out1, outn)]
    assert live[0] == live[1]
    rmse = np.sqrt(np.mean((one.astype(np.float64) - many.astype(np.float64)) ** 2))
    assert rmse <= 0.5, rmse


@pytest.mark.parametrize("name,w,h,spp", [("cornell", 160, 160, 200), ("cornell_plus", 160, 160, 200), ("spheres", 200, 150, 200),
                                          ("teapot", 200, 150, 100), ("glass_teapot", 200, 150, 100), ("lamp", 200, 150, 100)])
def test_image_rmse_vs_reference_gpu_binary(tmp_path, name, w, h, spp):
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/raytracing_ref was not built (needs /root/reference at build time)")
    scene = _scene_copy(tmp_path, name, w, h, spp, 10)
    ref_sort, out = _run(REF, tmp_path, scene)
    assert "GPU Took" in out
    ref_nosort, _ = _run(REF, tmp_path, scene, "no_sort")
    ours, _ = _run(EXE, tmp_path, scene)
    ours_nosort, _ = _run(EXE, tmp_path, scene, "no_sort")

    def rmse(a, b):
        return float(np.sqrt(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)))

    floor = rmse(ref_sort, ref_nosort)  # two reference runs with different streams = noise at this spp
    bound = 1.5 * floor + 0.5
    for img in (ours, ours_nosort):
        assert img.shape == ref_sort.shape
        assert rmse(img, ref_sort) <= bound, f"{name}: RMSE {rmse(img, ref_sort):.2f} vs reference, noise floor {floor:.2f}"
        assert rmse(img, ref_nosort) <= bound
        assert abs(img.mean() - ref_sort.mean()) <= 0.01 * ref_sort.mean() + 0.3
