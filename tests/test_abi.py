"""CPU tests of the drop-in boundary: librtcore.so loads, exports every symbol include/*.h declares (and the ctypes
mirror binds each of them), fails loudly without a GPU, and the product never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import pytest

import cuda_raytracer_b200 as rt
from cuda_raytracer_b200 import core
from util import ROOT, SCENES

HEADERS = [os.path.join(ROOT, "include", h) for h in ("rt_core.h", "rt_scene.h")]


def declared_functions():
    names = []
    for h in HEADERS:
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        text = re.sub(r"typedef\s+struct\s+\w+\s*\{.*?\}\s*\w+\s*;", "", text, flags=re.S)
        text = re.sub(r"typedef\s+enum\s+\w+\s*\{.*?\}\s*\w+\s*;", "", text, flags=re.S)
        names += re.findall(r"\b(rt_\w+)\s*\(", text)
    return sorted(set(names))


def test_every_declared_symbol_is_exported_and_bound():
    names = declared_functions()
    assert len(names) >= 20
    lib = rt.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/ but not exported by librtcore.so"
    assert sorted(core.EXPORTS) == names, "core.EXPORTS must list exactly the functions of include/*.h"
    out = subprocess.run(["nm", "-D", "--defined-only", core.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (rt_\w+)", out))
    assert set(names) <= exported


def test_api_version_matches_header():
    v = int(re.search(r"#define RT_API_VERSION (\d+)", open(HEADERS[0]).read()).group(1))
    assert rt.lib().rt_api_version() == v


def test_struct_layouts_match_header():
    """ctypes mirrors vs the C declarations: field order and count (sizes are checked against a compiled probe)."""
    src = r'''
    #include <stdio.h>
    #include "rt_core.h"
    int main(void) { printf("%zu %zu %zu\n", sizeof(rt_scene_desc), sizeof(rt_render_opts), sizeof(rt_stats)); return 0; }
    '''
    exe = os.path.join(ROOT, "oracle", "_build", "abi_probe")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    r = subprocess.run(["/usr/bin/gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src, text=True,
                       capture_output=True)
    assert r.returncode == 0, r.stderr
    sizes = [int(v) for v in subprocess.run([exe], capture_output=True, text=True).stdout.split()]
    assert sizes == [C.sizeof(core.SceneDesc), C.sizeof(core.RenderOpts), C.sizeof(core.Stats)]


def test_no_cpu_fallback_without_a_device():
    """No CUDA device -> rt_create fails with RT_ERR_NO_DEVICE and a message; nothing renders on the CPU."""
    if rt.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtError) as e:
        rt.Renderer(0)
    assert e.value.status == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(rt.RtError):
        rt.gpu_raytrace(rt.load_scene("cornell.scene", cwd=SCENES))


def test_null_arguments_are_rejected():
    lib = rt.lib()
    assert lib.rt_create(None, 0) == -1
    assert lib.rt_scene_load(None, 1, 0, None) == -1
    assert lib.rt_scene_describe(None, None) == -1
    assert lib.rt_render(None, None, None, None) == -1
    assert lib.rt_postprocess(None, 1, None, None) == -1
    lib.rt_destroy(None)
    lib.rt_scene_free(None)


def test_product_never_references_the_oracle():
    """The oracle is test infrastructure: no file of the package (Python or C++/CUDA) may import, include or link it."""
    pkg = os.path.join(ROOT, "cuda_raytracer_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="replace").read()
                assert "pyoracle" not in text and "rt_oracle" not in text and "oracle/" not in text.replace("anything under ``oracle/``", ""), \
                    f"{os.path.join(base, f)} mentions the oracle"
    out = subprocess.run(["ldd", core.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out


def test_front_end_usage_and_exit_codes(tmp_path):
    """raytracing <scene> [no_sort|cpu|no_gpu|no_bvh] (raytracing.cu:305-342): no argument -> usage, exit 1;
    `no_gpu` without `cpu` -> exit 2; `cpu` is refused (no CPU path in the drop-in) with a message."""
    exe = os.path.join(ROOT, "cuda_raytracer_b200", "bin", "raytracing")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 1 and "sage" in (r.stdout + r.stderr)
    r = subprocess.run([exe, "cornell.scene", "no_gpu"], cwd=SCENES, capture_output=True, text=True)
    assert r.returncode == 2 and "No raytracing hardware specified" in (r.stdout + r.stderr)


def test_index_division_by_multiply_and_shift_is_exact():
    """The kernels map a ray id to its pass, pixel and row with n / d computed as multiply-high + two shifts (FastDiv,
    device_scene.cuh); rt_index_divide runs the same constants and formula on the host. Exact for every 32-bit numerator:
    the divisors the render call uses (samples per pixel, widths, rays per pass up to 2^31), powers of two and their
    neighbours, and random pairs, each at the numerators where a wrong constant shows first (multiples of d and +-1)."""
    import random
    lib = rt.lib()
    rng = random.Random(5)
    divisors = [1, 2, 3, 5, 7, 20, 21, 33, 64, 97, 1000, 1920, 3840, 165_888_000, 2**31 - 1, 2**31, 2**31 + 1, 2**32 - 1]
    divisors += [2**k + o for k in range(1, 32) for o in (-1, 0, 1) if 2**k + o > 0] + [rng.randrange(1, 2**32) for _ in range(200)]
    for d in divisors:
        ns = {0, 1, d - 1, d, d + 1, 2**32 - 1, 2**32 - 2, 2**31, (2**32 - 1) // d * d, max(0, (2**32 - 1) // d * d - 1)}
        ns |= {rng.randrange(0, 2**32) for _ in range(40)} | {min(2**32 - 1, k * d + o) for k in (2, 3, 1000) for o in (-1, 0)}
        for n in (n for n in ns if 0 <= n < 2**32):
            assert lib.rt_index_divide(n, d) == n // d, (n, d)
    assert lib.rt_index_divide(12345, 0) == 12345  # d = 0 is treated as 1
