"""oracle/pyoracle.py — TEST INFRASTRUCTURE ONLY: ctypes access to the CPU oracle.

Two libraries:
  * ``oracle/_build/librt_oracle.so`` — the C restatement (``rt_oracle.c``), always buildable (gcc).
  * ``oracle/_ref/libref_shim.so``    — the reference's own code compiled in place (``ref_shim.cu``),
    present only where ``/root/reference`` was available at build time (this container; the built
    file travels to the GPU box).

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs
import this module. Nothing in the product package does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "_build", "librt_oracle.so")
SHIM_SO = os.path.join(HERE, "_ref", "libref_shim.so")
REF_BIN = os.path.join(HERE, "_ref", "raytracing_ref")

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")

NODE_DTYPE = np.dtype([("min", np.float32, 3), ("max", np.float32, 3), ("child1", np.int32), ("child2", np.int32)])
assert NODE_DTYPE.itemsize == 32


class OracleScene(C.Structure):
    _fields_ = [
        ("spheres", C.c_void_p), ("sphere_count", C.c_int32),
        ("triangles", C.c_void_p), ("triangle_count", C.c_int32),
        ("material_indices", C.c_void_p),
        ("materials", C.c_void_p), ("material_count", C.c_int32),
        ("bvh", C.c_void_p), ("bvh_node_count", C.c_int32),
        ("environment_map", C.c_void_p), ("environment_map_width", C.c_int32), ("environment_map_height", C.c_int32),
        ("width", C.c_int32), ("height", C.c_int32), ("ray_count", C.c_int32), ("bounces", C.c_int32),
        ("exposure", C.c_float),
        ("camera_position", C.c_float * 3),
        ("near_plane_top_left", C.c_float * 3),
        ("scaled_right", C.c_float * 3),
        ("scaled_up", C.c_float * 3),
        ("inv_width", C.c_float), ("inv_height", C.c_float),
        ("min_coord", C.c_float * 3),
        ("inv_dimensions", C.c_float * 3),
    ]


class Counters(C.Structure):
    _fields_ = [("live_ray_bounces", C.c_int64), ("internal_visits", C.c_int64),
                ("leaf_visits", C.c_int64), ("triangle_tests", C.c_int64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def build_port(force=False):
    if force or not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(HERE, "rt_oracle.c")):
        subprocess.run(["make", "-C", HERE, "port"], check=True, capture_output=True)
    return PORT_SO


def build_ref():
    """Build oracle/_ref from /root/reference if it is there; returns True if the shim exists."""
    if os.path.isdir("/root/reference") and not (os.path.exists(SHIM_SO) and os.path.exists(REF_BIN)):
        subprocess.run(["make", "-C", HERE, "ref"], check=True, capture_output=True)
    return os.path.exists(SHIM_SO)


_port = None


def port():
    global _port
    if _port is None:
        build_port()
        L = C.CDLL(PORT_SO)
        S = C.POINTER(OracleScene)
        L.oracle_xor_sequence.argtypes = [C.c_uint32, C.c_int, u32p]
        L.oracle_random_floats.argtypes = [C.c_uint32, f32p]
        L.oracle_project.argtypes = [f32p, C.c_int, f32p]
        L.oracle_ray_aabb.argtypes = [f32p, f32p, C.c_float, C.POINTER(C.c_float)]
        L.oracle_ray_aabb.restype = C.c_int
        L.oracle_gpu_gen_seed.argtypes = [C.c_int32, C.c_int32]
        L.oracle_gpu_gen_seed.restype = C.c_uint32
        L.oracle_gpu_proc_seed.argtypes = [C.c_int32, C.c_int32]
        L.oracle_gpu_proc_seed.restype = C.c_uint32
        L.oracle_closest_hit.argtypes = [S, f32p, C.c_int, C.c_int, f32p, i32p, C.POINTER(Counters)]
        L.oracle_closest_hit_ex.argtypes = [S, f32p, C.c_int, C.c_int, C.c_int, f32p, i32p, C.POINTER(Counters)]
        L.oracle_generate.argtypes = [S, f32p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.oracle_process.argtypes = [S, f32p, u32p, C.c_int]
        L.oracle_render.argtypes = [S, C.c_int, C.c_int, C.c_int, f32p, C.POINTER(Counters)]
        L.oracle_render_select.argtypes = [S, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, f32p, C.POINTER(Counters)]
        L.oracle_render_samples.argtypes = [S, C.c_int, C.c_int, C.c_int, f32p, C.POINTER(Counters)]
        L.oracle_sort_key.argtypes = [S, f32p, f32p, f32p]
        L.oracle_sort_key.restype = C.c_uint32
        L.oracle_bloom.argtypes = [f32p, C.c_int, C.c_int, C.c_int]
        L.oracle_tonemap.argtypes = [f32p, C.c_int, C.c_float, C.c_int, u8p]
        L.oracle_max_threads.restype = C.c_int
        _port = L
    return _port


class SceneArrays:
    """Reference-layout host arrays + scalars of one scene; the common currency of the tests.

    Built from the reference shim (``from_ref``), from the product's own host loader
    (``cuda_raytracer_b200.core.load_scene(...).arrays()``) or from numpy by hand.
    """

    def __init__(self, spheres, triangles, material_indices, materials, bvh, env, env_w, env_h, ints, floats):
        self.spheres = np.ascontiguousarray(spheres, np.float32).reshape(-1, 4)
        self.triangles = np.ascontiguousarray(triangles, np.float32).reshape(-1, 12)
        self.material_indices = np.ascontiguousarray(material_indices, np.uint16)
        self.materials = np.ascontiguousarray(materials, np.float32).reshape(-1, 12)
        self.bvh = np.ascontiguousarray(bvh, NODE_DTYPE)
        self.env = np.ascontiguousarray(env, np.float32).reshape(-1, 3)
        self.env_w, self.env_h = int(env_w), int(env_h)
        # ints: width height ray_count bounces ; floats: dict of named vectors / scalars
        self.width, self.height, self.ray_count, self.bounces = (int(v) for v in ints)
        self.f = {k: np.array(v, np.float32) for k, v in floats.items()}

    def oracle_struct(self):
        s = OracleScene()
        s.spheres = self.spheres.ctypes.data; s.sphere_count = len(self.spheres)
        s.triangles = self.triangles.ctypes.data; s.triangle_count = len(self.triangles)
        s.material_indices = self.material_indices.ctypes.data
        s.materials = self.materials.ctypes.data; s.material_count = len(self.materials)
        s.bvh = self.bvh.ctypes.data; s.bvh_node_count = len(self.bvh)
        s.environment_map = self.env.ctypes.data
        s.environment_map_width = self.env_w; s.environment_map_height = self.env_h
        s.width, s.height, s.ray_count, s.bounces = self.width, self.height, self.ray_count, self.bounces
        s.exposure = float(self.f["exposure"])
        for name in ("camera_position", "near_plane_top_left", "scaled_right", "scaled_up", "min_coord", "inv_dimensions"):
            getattr(s, name)[:] = [float(v) for v in self.f[name]]
        s.inv_width = float(self.f["inv_width"]); s.inv_height = float(self.f["inv_height"])
        return s


FLOAT_NAMES = [("camera_position", 3), ("forward", 3), ("up", 3), ("vertical_fov", 1), ("exposure", 1),
               ("min_coord", 3), ("inv_dimensions", 3), ("scaled_right", 3), ("scaled_up", 3),
               ("near_plane_top_left", 3), ("inv_width", 1), ("inv_height", 1)]


def unpack_floats(flat):
    out, k = {}, 0
    for name, n in FLOAT_NAMES:
        out[name] = np.float32(flat[k]) if n == 1 else np.array(flat[k:k + n], np.float32)
        k += n
    return out


_shim = None


def shim():
    """The reference's own functions (oracle/_ref/libref_shim.so); None when it was never built."""
    global _shim
    if _shim is None:
        if not os.path.exists(SHIM_SO):
            return None
        L = C.CDLL(SHIM_SO)
        L.ref_load_scene.argtypes = [C.c_char_p, C.c_int]
        L.ref_load_scene.restype = C.c_void_p
        L.ref_scene_ints.argtypes = [C.c_void_p, i32p]
        L.ref_scene_array.argtypes = [C.c_void_p, C.c_int]
        L.ref_scene_array.restype = C.c_void_p
        L.ref_scene_floats.argtypes = [C.c_void_p, f32p]
        L.ref_set_image.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
        L.ref_xor_sequence.argtypes = [C.c_uint32, C.c_int, u32p]
        L.ref_random_floats.argtypes = [C.c_uint32, f32p]
        L.ref_project.argtypes = [f32p, C.c_int, f32p]
        L.ref_ray_aabb.argtypes = [f32p, f32p, C.c_float, C.POINTER(C.c_float)]
        L.ref_ray_aabb.restype = C.c_int
        L.ref_closest_hit.argtypes = [C.c_void_p, f32p, C.c_int, f32p, i32p]
        L.ref_closest_hit_gpu.argtypes = [C.c_void_p, f32p, C.c_int, f32p, i32p]
        L.ref_closest_hit_gpu.restype = C.c_int
        L.ref_generate.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_process.argtypes = [C.c_void_p, f32p, u32p, C.c_int]
        L.ref_render.argtypes = [C.c_void_p, C.c_int, f32p, C.POINTER(C.c_longlong)]
        L.ref_cub_sort_pairs.argtypes = [u32p, u32p, C.c_int, C.c_int, C.POINTER(C.c_float)]
        L.ref_cub_sort_pairs.restype = C.c_int
        L.ref_gpu_gen_seed.argtypes = [C.c_int, C.c_int]
        L.ref_gpu_gen_seed.restype = C.c_uint32
        L.ref_gpu_proc_seed.argtypes = [C.c_int, C.c_int]
        L.ref_gpu_proc_seed.restype = C.c_uint32
        _shim = L
    return _shim


class RefScene:
    """A scene loaded by the reference's own load_scene (scene.cu:569). cwd must hold the assets."""

    def __init__(self, path, use_bvh=True, cwd=None):
        L = shim()
        assert L is not None, "oracle/_ref/libref_shim.so not built"
        old = os.getcwd()
        if cwd:
            os.chdir(cwd)
        try:
            # the reference prints Triangle count / BVH Took / Node count on stdout
            self.handle = L.ref_load_scene(os.fsencode(path), 1 if use_bvh else 0)
        finally:
            os.chdir(old)

    def set_image(self, width, height, ray_count, bounces, exposure):
        shim().ref_set_image(self.handle, width, height, ray_count, bounces, exposure)
        return self

    def arrays(self):
        L = shim()
        ints = np.zeros(10, np.int32)
        L.ref_scene_ints(self.handle, ints)
        ns, nt, nm, nn, w, h, rc, bo, ew, eh = (int(v) for v in ints)

        def grab(which, dtype, count):
            ptr = L.ref_scene_array(self.handle, which)
            if count == 0 or not ptr:
                return np.zeros(0, dtype)
            nbytes = np.dtype(dtype).itemsize * count
            buf = (C.c_char * nbytes).from_address(ptr)
            return np.frombuffer(buf, dtype=dtype, count=count).copy()

        flat = np.zeros(28, np.float32)
        L.ref_scene_floats(self.handle, flat)
        return SceneArrays(grab(0, np.float32, ns * 4), grab(1, np.float32, nt * 12), grab(2, np.uint16, ns + nt),
                           grab(3, np.float32, nm * 12), grab(4, NODE_DTYPE, nn), grab(5, np.float32, ew * eh * 3),
                           ew, eh, (w, h, rc, bo), unpack_floats(flat))
