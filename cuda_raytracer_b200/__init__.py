"""cuda_raytracer_b200 — B200-native render core behind the reference's render-path interface.

The package holds only what the hot path needs: ``csrc/`` (host scene layer in C++, sm_100a CUDA kernels, the C ABI)
and ``core.py`` (ctypes mirror of ``load_scene`` / ``gpu_raytrace``). See DESIGN.md.
"""
from .core import (MultiRenderer, Renderer, RtError, Scene, device_count, gpu_raytrace, lib, load_scene, write_png)  # noqa: F401

__all__ = ["MultiRenderer", "Renderer", "RtError", "Scene", "device_count", "gpu_raytrace", "lib", "load_scene", "write_png"]
