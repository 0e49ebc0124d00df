"""ctypes binding of librtcore.so: the Python mirror of the reference's render-path interface.

Reference interface mirrored (same names, argument meaning and output convention):

* ``load_scene(path, use_bvh)``   <- ``void load_scene(Scene*, const char*, bool)``  /root/reference/scene.cuh:168
* ``gpu_raytrace(scene, sort)``    <- ``Vec3 *gpu_raytrace(const Scene*, bool sort)`` /root/reference/raytracing.cu:170
  returns the un-normalised radiance sums, ``float32[H, W, 3]``, top row first.
* ``bloom`` / ``tonemap``          <- raytracing.cu:360-392 / :286-303

Everything computes on the GPU through the C ABI (``include/rt_core.h``); there is no CPU fallback and no import of
anything under ``oracle/``. Loading fails loudly if the shared library has not been built.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_LIBRTCORE") or os.path.join(_PKG, "librtcore.so")  # override = A/B builds of the same ABI

RT_OK = 0
NODE_DTYPE = np.dtype([("min", np.float32, 3), ("max", np.float32, 3), ("child1", np.int32), ("child2", np.int32)])


class RtError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rt status {status}: {message}")
        self.status = status


class SceneDesc(C.Structure):
    """rt_scene_desc (include/rt_core.h)."""
    _fields_ = [
        ("spheres", C.c_void_p), ("sphere_count", C.c_int32),
        ("triangles", C.c_void_p), ("triangle_count", C.c_int32),
        ("material_indices", C.c_void_p),
        ("materials", C.c_void_p), ("material_count", C.c_int32),
        ("bvh_nodes", C.c_void_p), ("bvh_node_count", C.c_int32),
        ("environment_map", C.c_void_p), ("env_width", C.c_int32), ("env_height", C.c_int32),
        ("width", C.c_int32), ("height", C.c_int32), ("ray_count", C.c_int32), ("bounces", C.c_int32),
        ("exposure", C.c_float),
        ("camera_position", C.c_float * 3), ("near_plane_top_left", C.c_float * 3),
        ("scaled_right", C.c_float * 3), ("scaled_up", C.c_float * 3),
        ("inv_width", C.c_float), ("inv_height", C.c_float),
        ("min_coord", C.c_float * 3), ("inv_dimensions", C.c_float * 3),
        ("triangles_form", C.c_int32),
    ]


class RenderOpts(C.Structure):
    _fields_ = [("sort", C.c_int32), ("pass_spp", C.c_int32), ("pass_first", C.c_int32), ("pass_count", C.c_int32),
                ("collect_stats", C.c_int32), ("stream_id", C.c_int32), ("sort_bits", C.c_int32), ("pass_group", C.c_int32),
                ("kernel_timing", C.c_int32), ("root_start", C.c_int32), ("single_pipe", C.c_int32), ("reserved", C.c_int32),
                ("ray_first", C.c_int64), ("ray_count", C.c_int64), ("sample_first", C.c_int32), ("sample_count", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("loop_seconds", C.c_double), ("total_seconds", C.c_double), ("primary_rays", C.c_int64),
                ("live_ray_bounces", C.c_int64), ("internal_visits", C.c_int64), ("triangle_tests", C.c_int64),
                ("sorted_pairs", C.c_int64), ("kernel_launches", C.c_int32), ("passes", C.c_int32),
                ("trace_ms", C.c_double), ("shade_ms", C.c_double), ("sort_ms", C.c_double), ("live_per_bounce", C.c_int64 * 32),
                ("trace_ms_per_bounce", C.c_double * 32), ("shade_ms_per_bounce", C.c_double * 32)]

    def as_dict(self):
        arrays = ("live_per_bounce", "trace_ms_per_bounce", "shade_ms_per_bounce")
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in arrays}
        d["live_per_bounce"] = [int(v) for v in self.live_per_bounce]
        d["trace_ms_per_bounce"] = [float(v) for v in self.trace_ms_per_bounce]
        d["shade_ms_per_bounce"] = [float(v) for v in self.shade_ms_per_bounce]
        return d


EXPORTS = [
    # rt_core.h
    "rt_api_version", "rt_device_count", "rt_host_pin", "rt_host_unpin", "rt_create", "rt_set_node_mode", "rt_get_node_mode", "rt_destroy", "rt_last_error", "rt_scene_upload", "rt_scene_set_image",
    "rt_render", "rt_render_device", "rt_postprocess", "rt_framebuffer_upload", "rt_trace_closest", "rt_generate_rays",
    "rt_process_rays", "rt_sort_pairs", "rt_sort_pairs_static", "rt_scene_download_bvh", "rt_scene_bvh_device_ms", "rt_sort_device_ms", "rt_index_divide",
    "rt_probe_copy_gbs",
    "rt_plan_waves",
    "rt_multi_create", "rt_multi_destroy", "rt_multi_device_count", "rt_multi_last_error", "rt_multi_context", "rt_multi_set_node_mode",
    "rt_multi_scene_upload", "rt_multi_render", "rt_multi_reduce_ms", "rt_multi_postprocess",
    # rt_scene.h
    "rt_scene_load", "rt_scene_free", "rt_scene_describe", "rt_scene_set_image_host", "rt_scene_camera_raw",
    "rt_scene_bvh_build_ms", "rt_write_png",
]

_lib = None


def lib():
    """Load librtcore.so (built by ``__graft_entry__.build()`` / ``make -C cuda_raytracer_b200/csrc``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no fallback implementation)")
    L = C.CDLL(LIB_PATH)
    vp, i32, f32p = C.c_void_p, C.c_int32, C.POINTER(C.c_float)
    L.rt_last_error.restype = C.c_char_p
    L.rt_last_error.argtypes = [vp]
    L.rt_create.argtypes = [C.POINTER(vp), C.c_int]
    L.rt_host_pin.argtypes = [vp, C.c_size_t]
    L.rt_host_unpin.argtypes = [vp]
    L.rt_host_unpin.restype = None
    L.rt_set_node_mode.argtypes = [vp, C.c_int]
    L.rt_get_node_mode.argtypes = [vp]
    L.rt_destroy.argtypes = [vp]
    L.rt_destroy.restype = None
    L.rt_scene_upload.argtypes = [vp, C.POINTER(SceneDesc)]
    L.rt_scene_set_image.argtypes = [vp, C.POINTER(SceneDesc)]
    L.rt_render.argtypes = [vp, C.POINTER(RenderOpts), vp, C.POINTER(Stats)]
    L.rt_render_device.argtypes = [vp, C.POINTER(RenderOpts), C.POINTER(vp), C.POINTER(Stats)]
    L.rt_postprocess.argtypes = [vp, C.c_int, vp, vp]
    L.rt_framebuffer_upload.argtypes = [vp, vp]
    L.rt_trace_closest.argtypes = [vp, vp, i32, i32, vp, vp, C.POINTER(Stats)]
    L.rt_generate_rays.argtypes = [vp, i32, i32, i32, i32, vp]
    L.rt_process_rays.argtypes = [vp, vp, vp, i32]
    L.rt_sort_pairs.argtypes = [vp, vp, vp, i32, i32]
    L.rt_sort_pairs_static.argtypes = [vp, vp, vp, vp, vp, i32, i32]
    L.rt_scene_download_bvh.argtypes = [vp, vp, i32, C.POINTER(i32), vp, i32]
    L.rt_scene_bvh_device_ms.argtypes = [vp]
    L.rt_scene_bvh_device_ms.restype = C.c_double
    L.rt_sort_device_ms.argtypes = [vp]
    L.rt_sort_device_ms.restype = C.c_double
    L.rt_probe_copy_gbs.argtypes = [vp, C.c_int64, C.c_int32, C.POINTER(C.c_double)]
    L.rt_plan_waves.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int64]
    L.rt_index_divide.argtypes = [C.c_uint32, C.c_uint32]
    L.rt_index_divide.restype = C.c_uint32
    L.rt_multi_create.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), C.c_int]
    L.rt_multi_destroy.argtypes = [vp]
    L.rt_multi_destroy.restype = None
    L.rt_multi_device_count.argtypes = [vp]
    L.rt_multi_last_error.argtypes = [vp]
    L.rt_multi_last_error.restype = C.c_char_p
    L.rt_multi_context.argtypes = [vp, C.c_int]
    L.rt_multi_context.restype = vp
    L.rt_multi_set_node_mode.argtypes = [vp, C.c_int]
    L.rt_multi_scene_upload.argtypes = [vp, C.POINTER(SceneDesc)]
    L.rt_multi_render.argtypes = [vp, C.POINTER(RenderOpts), vp, C.POINTER(Stats)]
    L.rt_multi_reduce_ms.argtypes = [vp]
    L.rt_multi_reduce_ms.restype = C.c_double
    L.rt_multi_postprocess.argtypes = [vp, C.c_int, vp, vp]
    L.rt_scene_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(vp)]
    L.rt_scene_free.argtypes = [vp]
    L.rt_scene_free.restype = None
    L.rt_scene_describe.argtypes = [vp, C.POINTER(SceneDesc)]
    L.rt_scene_set_image_host.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    L.rt_scene_camera_raw.argtypes = [vp, f32p]
    L.rt_scene_bvh_build_ms.argtypes = [vp]
    L.rt_scene_bvh_build_ms.restype = C.c_double
    L.rt_write_png.argtypes = [C.c_char_p, C.c_int, C.c_int, vp]
    _lib = L
    return L


def build_id():
    """Identifies the build an ncu capture (profiles/*_traffic.json) or a bench line ran. For the in-tree library: "src-" + the
    sha1 the Makefile writes next to it (librtcore.srcid: sources + compile flags) — nvcc's output is not byte-reproducible, so
    two builds of the same tree have different binaries but the same id. For a library selected with RT_LIBRTCORE (A/B
    variants built with extra -D flags), or without that file: "bin-" + the sha1 of the binary."""
    import hashlib
    default = os.path.join(os.path.dirname(os.path.abspath(__file__)), "librtcore.so")
    stamp = os.path.join(os.path.dirname(default), "librtcore.srcid")
    if os.path.abspath(LIB_PATH) == default and os.path.exists(stamp):
        return "src-" + open(stamp).read().strip()
    with open(LIB_PATH, "rb") as f:
        return "bin-" + hashlib.sha1(f.read()).hexdigest()[:16]


def plan_waves(width, height, ray_count, max_group, pipes, pass_spp=0, pass_first=0, pass_count=0, ray_first=0, ray_range=0,
               sample_first=0, sample_count=0):
    """rt_plan_waves (host only): the waves rt_render would form, as a list of (members, partial_member) with members =
    list of (pass, offset, count, sub_first, sub_count). ``ray_range`` is rt_render_opts.ray_count."""
    o = Renderer._opts(False, pass_spp, pass_first, pass_count, False, 0, 0, ray_first=ray_first, ray_count=ray_range,
                       sample_first=sample_first, sample_count=sample_count)
    passes = (ray_count + (pass_spp or 20) - 1) // (pass_spp or 20)
    cap = 7 * (passes + 4) + 64
    out = (C.c_int64 * cap)()
    n = lib().rt_plan_waves(C.byref(o), width, height, ray_count, max_group, pipes, out, cap)
    if n < 0:
        raise RtError(n, "rt_plan_waves: option out of range")
    waves, at = [], 0
    for _ in range(n):
        group, partial = int(out[at]), int(out[at + 1])
        at += 2
        waves.append(([tuple(int(v) for v in out[at + 5 * m: at + 5 * m + 5]) for m in range(group)], partial))
        at += 5 * group
    return waves


_KERNEL_ID = {}


def kernel_id(pattern="trace_rays_kernel|shade_kernel"):
    """sha1 of the SASS (``cuobjdump -sass``) of the loaded library's kernels whose names match ``pattern`` — the identity of the
    machine code an ncu capture measured. It is the same for two builds of the same device code (host-only changes and nvcc's
    non-reproducible symbol ids leave it alone) and changes with any instruction of those kernels. None without cuobjdump."""
    import hashlib
    import re
    import shutil
    import subprocess
    if pattern in _KERNEL_ID:
        return _KERNEL_ID[pattern]
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    kid = None
    try:
        out = subprocess.run([tool, "-sass", LIB_PATH], capture_output=True, text=True, timeout=120, check=True).stdout
        keep, picked, rx = False, [], re.compile(pattern)
        for line in out.splitlines():
            if "Function : " in line:
                keep = bool(rx.search(line))
            if keep:
                picked.append(line)
        if picked:
            kid = "sass-" + hashlib.sha1(("\n".join(picked) + "\n").encode()).hexdigest()[:16]
    except (OSError, subprocess.SubprocessError):
        kid = None
    _KERNEL_ID[pattern] = kid
    return kid


def _check(rc, ctx=None):
    if rc != RT_OK:
        msg = lib().rt_last_error(ctx)
        raise RtError(rc, msg.decode() if msg else "")


def _view(ptr, dtype, count):
    if count == 0 or not ptr:
        return np.zeros(0, dtype)
    nbytes = np.dtype(dtype).itemsize * count
    return np.frombuffer((C.c_char * nbytes).from_address(ptr), dtype=dtype, count=count)


class Scene:
    """A host scene (rt_host_scene): what the reference's ``load_scene`` fills into its ``Scene`` struct."""

    def __init__(self, path, use_bvh=True, verbose=False, cwd=None):
        """use_bvh: True / False as the reference's flag; "device" (= 2) defers the BVH build to the GPU; "sbvh" (= 3) builds
        a spatial-split tree on the host (same hits, fewer visited records; not the reference's tree)."""
        if use_bvh == "device":
            use_bvh = 2
        elif use_bvh == "sbvh":
            use_bvh = 3
        self._h = C.c_void_p()
        old = os.getcwd()
        if cwd:
            os.chdir(cwd)  # paths inside scene files are CWD-relative, as in the reference
        try:
            _check(lib().rt_scene_load(os.fsencode(path), int(use_bvh), 1 if verbose else 0, C.byref(self._h)))
        finally:
            os.chdir(old)

    def close(self):
        if self._h:
            lib().rt_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def desc(self):
        d = SceneDesc()
        _check(lib().rt_scene_describe(self._h, C.byref(d)))
        return d

    def set_image(self, width, height, ray_count, bounces, exposure=-1.0):
        _check(lib().rt_scene_set_image_host(self._h, width, height, ray_count, bounces, exposure))
        return self

    @property
    def bvh_build_ms(self):
        return lib().rt_scene_bvh_build_ms(self._h)

    def camera_raw(self):
        out = (C.c_float * 7)()
        _check(lib().rt_scene_camera_raw(self._h, out))
        return np.array(out[:], np.float32)

    def arrays(self):
        """Copies of the reference-layout arrays + scalars as a dict (used by the tests to feed the oracle)."""
        d = self.desc()
        prims = d.sphere_count + d.triangle_count
        cam = self.camera_raw()
        return {
            "spheres": _view(d.spheres, np.float32, 4 * d.sphere_count).copy().reshape(-1, 4),
            "triangles": _view(d.triangles, np.float32, 12 * d.triangle_count).copy().reshape(-1, 12),
            "material_indices": _view(d.material_indices, np.uint16, prims).copy(),
            "materials": _view(d.materials, np.float32, 12 * d.material_count).copy().reshape(-1, 12),
            "bvh": _view(d.bvh_nodes, NODE_DTYPE, d.bvh_node_count).copy(),
            "env": _view(d.environment_map, np.float32, 3 * d.env_width * d.env_height).copy().reshape(-1, 3),
            "env_w": d.env_width, "env_h": d.env_height,
            "triangles_form": d.triangles_form,
            "width": d.width, "height": d.height, "ray_count": d.ray_count, "bounces": d.bounces,
            "exposure": np.float32(d.exposure),
            "camera_position": np.array(d.camera_position[:], np.float32),
            "forward": cam[0:3], "up": cam[3:6], "vertical_fov": cam[6],
            "near_plane_top_left": np.array(d.near_plane_top_left[:], np.float32),
            "scaled_right": np.array(d.scaled_right[:], np.float32),
            "scaled_up": np.array(d.scaled_up[:], np.float32),
            "inv_width": np.float32(d.inv_width), "inv_height": np.float32(d.inv_height),
            "min_coord": np.array(d.min_coord[:], np.float32),
            "inv_dimensions": np.array(d.inv_dimensions[:], np.float32),
        }


def load_scene(path, use_bvh=True, verbose=False, cwd=None):
    """Mirror of the reference's ``load_scene`` (scene.cu:569): parse, load assets, build the BVH."""
    return Scene(path, use_bvh, verbose, cwd)


class Renderer:
    """One render context bound to one CUDA device (rt_ctx)."""

    NODE_MODES = {"f32": 0, "q16": 1, "w4": 2}

    def __init__(self, device=0, nodes=None):
        """nodes: None = the library's default (environment RT_NODES), else "f32" | "q16" | "w4" (rt_set_node_mode)."""
        self._ctx = C.c_void_p()
        _check(lib().rt_create(C.byref(self._ctx), device))
        self.width = self.height = 0
        self.last_stats = None
        if nodes is not None:
            self.set_node_mode(nodes)

    def set_node_mode(self, nodes):
        """Node format of the closest-hit kernel for the NEXT upload: "f32" (64-byte float records), "q16" (32-byte quantised
        records of the same binary tree), "w4" (4-wide quantised nodes)."""
        _check(lib().rt_set_node_mode(self._ctx, self.NODE_MODES[nodes] if isinstance(nodes, str) else int(nodes)), self._ctx)
        return self

    @property
    def node_mode(self):
        m = lib().rt_get_node_mode(self._ctx)
        return {v: k for k, v in self.NODE_MODES.items()}.get(m, m)

    def close(self):
        if self._ctx:
            lib().rt_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, scene):
        d = scene.desc() if isinstance(scene, Scene) else scene
        _check(lib().rt_scene_upload(self._ctx, C.byref(d)), self._ctx)
        self.width, self.height = d.width, d.height
        return self

    def set_image(self, scene):
        d = scene.desc() if isinstance(scene, Scene) else scene
        _check(lib().rt_scene_set_image(self._ctx, C.byref(d)), self._ctx)
        self.width, self.height = d.width, d.height
        return self

    @staticmethod
    def _opts(sort, pass_spp, pass_first, pass_count, collect_stats, stream_id, sort_bits, timing=False, pass_group=0, root_start=False,
              single_pipe=False, ray_first=0, ray_count=0, sample_first=0, sample_count=0):
        return RenderOpts(1 if sort else 0, pass_spp, pass_first, pass_count, 1 if collect_stats else 0, stream_id, sort_bits,
                          pass_group, 1 if timing else 0, 1 if root_start else 0, 1 if single_pipe else 0, 0, ray_first, ray_count,
                          sample_first, sample_count)

    def render(self, sort=True, out=None, pass_spp=0, pass_first=0, pass_count=0, collect_stats=False, stream_id=0, sort_bits=0,
               timing=False, pass_group=0, root_start=False, single_pipe=False, ray_first=0, ray_count=0, sample_first=0, sample_count=0):
        """rt_render: host output (float32[H,W,3] un-normalised sums). ``out`` may be a preallocated (pinned) array.
        ``root_start`` makes primary rays walk from the BVH root instead of their tile's entry records (same results);
        ``ray_first`` / ``ray_count`` (> 0) select a range of the frame's ray sequence instead of whole passes,
        ``sample_first`` / ``sample_count`` (> 0) a range of every pixel's samples (the multi-GPU split)."""
        if out is None:
            out = np.empty((self.height, self.width, 3), np.float32)
        assert out.dtype == np.float32 and out.size == self.width * self.height * 3 and out.flags.c_contiguous
        o = self._opts(sort, pass_spp, pass_first, pass_count, collect_stats, stream_id, sort_bits, timing, pass_group, root_start,
                       single_pipe, ray_first, ray_count, sample_first, sample_count)
        st = Stats()
        _check(lib().rt_render(self._ctx, C.byref(o), out.ctypes.data, C.byref(st)), self._ctx)
        self.last_stats = st
        return out

    def render_device(self, sort=True, pass_spp=0, pass_first=0, pass_count=0, collect_stats=False, stream_id=0, sort_bits=0,
                      timing=False, pass_group=0, root_start=False, single_pipe=False, ray_first=0, ray_count=0, sample_first=0, sample_count=0):
        """rt_render_device: returns the device pointer (int) of the float32[H*W*3] framebuffer."""
        o = self._opts(sort, pass_spp, pass_first, pass_count, collect_stats, stream_id, sort_bits, timing, pass_group, root_start,
                       single_pipe, ray_first, ray_count, sample_first, sample_count)
        st = Stats()
        ptr = C.c_void_p()
        _check(lib().rt_render_device(self._ctx, C.byref(o), C.byref(ptr), C.byref(st)), self._ctx)
        self.last_stats = st
        return ptr.value

    def framebuffer_upload(self, rgb_sum):
        a = np.ascontiguousarray(rgb_sum, np.float32)
        assert a.size == self.width * self.height * 3
        _check(lib().rt_framebuffer_upload(self._ctx, a.ctypes.data), self._ctx)

    def postprocess(self, bloom=True, want_sum=False, want_rgb8=True):
        s = np.empty((self.height, self.width, 3), np.float32) if want_sum else None
        b = np.empty((self.height, self.width, 3), np.uint8) if want_rgb8 else None
        _check(lib().rt_postprocess(self._ctx, 1 if bloom else 0, s.ctypes.data if want_sum else None,
                                    b.ctypes.data if want_rgb8 else None), self._ctx)
        return s, b

    def trace_closest(self, rays6, with_spheres=False, stats=False):
        r = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        n = len(r)
        prim = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        st = Stats()
        _check(lib().rt_trace_closest(self._ctx, r.ctypes.data, n, 1 if with_spheres else 0, prim.ctypes.data, t.ctypes.data,
                                      C.byref(st) if stats else None), self._ctx)
        self.last_stats = st if stats else None
        return prim, t

    def generate_rays(self, rays_per_pixel, first, n, seed):
        out = np.zeros((n, 12), np.float32)
        _check(lib().rt_generate_rays(self._ctx, rays_per_pixel, first, n, seed, out.ctypes.data), self._ctx)
        return out

    def process_rays(self, ray_data12, seeds):
        d = np.ascontiguousarray(ray_data12, np.float32).reshape(-1, 12).copy()
        s = np.ascontiguousarray(seeds, np.uint32)
        assert len(s) == len(d)
        _check(lib().rt_process_rays(self._ctx, d.ctypes.data, s.ctypes.data, len(d)), self._ctx)
        return d

    def download_bvh(self, triangle_count):
        """After uploading a scene loaded with use_bvh="device": (nodes[NODE_DTYPE], order int32[triangle_count])."""
        n = C.c_int32()
        _check(lib().rt_scene_download_bvh(self._ctx, None, 0, C.byref(n), None, 0), self._ctx)
        nodes = np.zeros(n.value, NODE_DTYPE)
        order = np.zeros(triangle_count, np.int32)
        _check(lib().rt_scene_download_bvh(self._ctx, nodes.ctypes.data, n.value, C.byref(n), order.ctypes.data, triangle_count), self._ctx)
        return nodes, order

    def probe_copy_gbs(self, nbytes, iters=20):
        """Copy bandwidth (read + write) of a buffer pair of `nbytes` each: L2-resident when small, HBM when large."""
        g = C.c_double()
        _check(lib().rt_probe_copy_gbs(self._ctx, int(nbytes), int(iters), C.byref(g)), self._ctx)
        return g.value

    @property
    def bvh_device_ms(self):
        return lib().rt_scene_bvh_device_ms(self._ctx)

    @property
    def sort_device_ms(self):
        return lib().rt_sort_device_ms(self._ctx)

    def sort_pairs_static(self, d_keys_in, d_values_in, d_keys_out, d_values_out, n, bits=32):
        """rt_sort_pairs_static: the same sort on DEVICE pointers (ints) with all scratch in the library's static device array:
        no allocation, n <= SORT_STATIC_MAX_PAIRS. Calls on one device must not overlap."""
        _check(lib().rt_sort_pairs_static(self._ctx, int(d_keys_in), int(d_values_in), int(d_keys_out), int(d_values_out), int(n), int(bits)), self._ctx)

    def sort_pairs(self, keys, values, bits=32):
        k = np.ascontiguousarray(keys, np.uint32).copy()
        v = np.ascontiguousarray(values, np.uint32).copy()
        assert len(k) == len(v)
        _check(lib().rt_sort_pairs(self._ctx, k.ctypes.data, v.ctypes.data, len(k), bits), self._ctx)
        return k, v


class MultiRenderer:
    """Several GPUs of one box behind one handle (rt_multi): the frame's rays are split across `devices`, the accumulation
    buffers are summed with one NCCL reduce into devices[0]. Naming one device several times runs the shares on that device
    (no NCCL)."""

    def __init__(self, devices, nodes=None):
        devs = (C.c_int * len(devices))(*devices)
        self._m = C.c_void_p()
        rc = lib().rt_multi_create(C.byref(self._m), devs, len(devices))
        if rc != RT_OK:
            msg = lib().rt_multi_last_error(None)
            raise RtError(rc, msg.decode() if msg else "")
        self.devices = list(devices)
        self.width = self.height = 0
        self.last_stats = None
        if nodes is not None:
            self._check(lib().rt_multi_set_node_mode(self._m, Renderer.NODE_MODES[nodes] if isinstance(nodes, str) else int(nodes)))

    def _check(self, rc):
        if rc != RT_OK:
            msg = lib().rt_multi_last_error(self._m)
            raise RtError(rc, msg.decode() if msg else "")

    def close(self):
        if self._m:
            lib().rt_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, scene):
        d = scene.desc() if isinstance(scene, Scene) else scene
        self._check(lib().rt_multi_scene_upload(self._m, C.byref(d)))
        self.width, self.height = d.width, d.height
        return self

    def render(self, sort=True, out=None, want_host=True, **kw):
        """rt_multi_render: the combined un-normalised sums (float32[H,W,3]); want_host=False leaves them on devices[0]."""
        if want_host and out is None:
            out = np.empty((self.height, self.width, 3), np.float32)
        o = Renderer._opts(sort, kw.pop("pass_spp", 0), kw.pop("pass_first", 0), kw.pop("pass_count", 0), kw.pop("collect_stats", False),
                           kw.pop("stream_id", 0), kw.pop("sort_bits", 0), kw.pop("timing", False), kw.pop("pass_group", 0),
                           kw.pop("root_start", False), kw.pop("single_pipe", False), kw.pop("ray_first", 0), kw.pop("ray_count", 0),
                           kw.pop("sample_first", 0), kw.pop("sample_count", 0))
        assert not kw, f"unknown options {sorted(kw)}"
        st = Stats()
        self._check(lib().rt_multi_render(self._m, C.byref(o), out.ctypes.data if want_host else None, C.byref(st)))
        self.last_stats = st
        return out if want_host else None

    @property
    def reduce_ms(self):
        return lib().rt_multi_reduce_ms(self._m)

    def postprocess(self, bloom=True, want_sum=False, want_rgb8=True):
        s = np.empty((self.height, self.width, 3), np.float32) if want_sum else None
        b = np.empty((self.height, self.width, 3), np.uint8) if want_rgb8 else None
        self._check(lib().rt_multi_postprocess(self._m, 1 if bloom else 0, s.ctypes.data if want_sum else None, b.ctypes.data if want_rgb8 else None))
        return s, b


SORT_STATIC_MAX_PAIRS = 1 << 20  # RT_SORT_STATIC_MAX_PAIRS (include/rt_core.h)


def device_count():
    return lib().rt_device_count()


def gpu_raytrace(scene, sort=True, device=0):
    """Mirror of ``Vec3 *gpu_raytrace(const Scene*, bool sort)`` (raytracing.cu:170): upload, render, download."""
    r = Renderer(device)
    try:
        return r.upload(scene).render(sort=sort)
    finally:
        r.close()


def write_png(path, rgb8):
    a = np.ascontiguousarray(rgb8, np.uint8)
    h, w, _ = a.shape
    _check(lib().rt_write_png(os.fsencode(path), w, h, a.ctypes.data))
