#!/usr/bin/env python3
"""Synthesise the environment map the bundled teapot/lamp scenes name.

The reference checkout lists `teapot/textures/envmap.pfm` in `.MISSING_LARGE_BLOBS`
(SURVEY.md F6), so `teapot.scene`, `glass_teapot.scene` and `lamp.scene` cannot load
their `sky_map`.  This script writes a deterministic stand-in in exactly the layout the
reference's `load_pfm` reads (`/root/reference/scene.cu:548-567`): three text lines
(`PF`, `<W> <H>`, `-1.0`) followed by W*H little-endian float[3] texels in file order.

The map is square (the reference indexes rows with the *height*), a smooth
sky gradient in roughly 0.3..1.0 plus a small HDR "sun" patch >> 1 so the bloom
threshold triggers.  No random numbers: the file is a pure function of (size,).
"""
import sys
import os
import numpy as np


def make_envmap(size: int = 1024) -> np.ndarray:
    v, u = np.meshgrid(
        (np.arange(size, dtype=np.float64) + 0.5) / size,
        (np.arange(size, dtype=np.float64) + 0.5) / size,
        indexing="ij",
    )
    # centre of the square = +z pole of the equal-area octahedral map, rim = -z pole
    r = np.maximum(np.abs(u - 0.5), np.abs(v - 0.5)) * 2.0
    up = 1.0 - r  # 1 at the pole, 0 at the rim
    horizon = np.exp(-((r - 0.55) ** 2) / 0.02)
    img = np.empty((size, size, 3), dtype=np.float64)
    img[..., 0] = 0.30 + 0.35 * up + 0.30 * horizon
    img[..., 1] = 0.38 + 0.40 * up + 0.22 * horizon
    img[..., 2] = 0.50 + 0.48 * up + 0.08 * horizon
    # low-frequency cloud-like modulation, deterministic
    img *= (0.9 + 0.1 * np.sin(9.0 * u + 2.0 * np.sin(7.0 * v)))[..., None]
    # HDR sun: small disc, peak ~60
    d2 = (u - 0.62) ** 2 + (v - 0.41) ** 2
    sun = 60.0 * np.exp(-d2 / (2 * 0.012 ** 2))
    img[..., 0] += sun
    img[..., 1] += 0.92 * sun
    img[..., 2] += 0.75 * sun
    return img.astype("<f4")


def write_pfm(path: str, img: np.ndarray) -> None:
    h, w, _ = img.shape
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(b"PF\n%d %d\n-1.0\n" % (w, h))
        f.write(img.tobytes())
    os.replace(tmp, path)


def main(argv):
    here = os.path.dirname(os.path.abspath(__file__))
    out = argv[1] if len(argv) > 1 else os.path.join(here, "teapot", "textures", "envmap.pfm")
    size = int(argv[2]) if len(argv) > 2 else 1024
    write_pfm(out, make_envmap(size))
    print(f"wrote {out} ({size}x{size})")


if __name__ == "__main__":
    main(sys.argv)
