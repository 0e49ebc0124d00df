"""CPU tests of the host scene layer (scene parser, PLY / PFM loaders, camera precompute, binned-SAH BVH build, PNG).

The product loader must reproduce the reference's `load_scene` (scene.cu:569-831) byte for byte: same primitive order
after the BVH build (`BvhNode::maybe_split`, scene.cu:866-1000), same node array, same camera block. That is pinned
two ways: against SHA-256 digests + float blocks stored in tests/golden/ref_vectors.npz (made from the reference's own
loader by tests/golden/make_golden.py; works anywhere), and — where oracle/_ref exists — against the reference loader
run live. Error behaviour follows include/rt_scene.h: status codes + message instead of the reference's uncaught
exceptions.
"""
import hashlib
import os
import struct
import zlib

import numpy as np
import pytest

import pyoracle as po
import cuda_raytracer_b200 as rt
from util import ALL_SCENES, GOLDEN, SCENES

G = np.load(GOLDEN + "/ref_vectors.npz")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def flat_floats(a):
    return np.concatenate([np.atleast_1d(a[k]) for k, _ in po.FLOAT_NAMES]).astype(np.float32)


@pytest.mark.parametrize("name", ALL_SCENES)
@pytest.mark.parametrize("use_bvh", [True, False])
def test_loader_matches_reference_digests(name, use_bvh):
    s = rt.load_scene(name + ".scene", use_bvh=use_bvh, cwd=SCENES)
    a = s.arrays()
    tag = f"{name}_{'bvh' if use_bvh else 'nobvh'}"
    counts = G[f"scene_{tag}_counts"]
    assert [len(a["spheres"]), len(a["triangles"]), len(a["materials"]), len(a["bvh"])] == list(counts)
    want = list(G[f"scene_{tag}_sha"])
    got = [sha(a["spheres"]), sha(a["triangles"]), sha(a["material_indices"]), sha(a["materials"]), sha(a["bvh"])]
    assert got == want, "spheres/triangles/material_indices/materials/bvh digests (in that order)"
    if use_bvh:
        assert [a["width"], a["height"], a["ray_count"], a["bounces"]] == list(G[f"scene_{tag}_image"])
        # camera block, scene bounds and image scalars exactly as load_scene leaves them
        assert flat_floats(a).tobytes() == G[f"scene_{tag}_floats"].tobytes()


@pytest.mark.parametrize("name", ALL_SCENES)
def test_loader_matches_reference_live(name):
    if po.shim() is None:
        pytest.skip("oracle/_ref not built (no /root/reference at build time)")
    ref = po.RefScene(name + ".scene", cwd=SCENES).arrays()
    a = rt.load_scene(name + ".scene", cwd=SCENES).arrays()
    for key, want in (("spheres", ref.spheres), ("triangles", ref.triangles), ("material_indices", ref.material_indices),
                      ("materials", ref.materials), ("bvh", ref.bvh), ("env", ref.env)):
        assert np.ascontiguousarray(a[key]).tobytes() == np.ascontiguousarray(want).tobytes(), key
    for k, _ in po.FLOAT_NAMES:
        assert np.asarray(a[k], np.float32).tobytes() == np.asarray(ref.f[k], np.float32).tobytes(), k


def test_reference_primitive_counts():
    """REPORT.pdf Table 1 / SURVEY §4: cornell 32 triangles, cornell_plus 32 + 2 spheres, spheres 4, teapot 126,050;
    lamp is the 48-mesh subset (288,822) because two meshes are absent from the reference checkout (F6)."""
    counts = {n: rt.load_scene(n + ".scene", cwd=SCENES).arrays() for n in ("cornell", "cornell_plus", "spheres", "teapot", "lamp")}
    assert len(counts["cornell"]["triangles"]) == 32 and len(counts["cornell"]["bvh"]) == 21
    assert len(counts["cornell_plus"]["triangles"]) == 32 and len(counts["cornell_plus"]["spheres"]) == 2
    assert len(counts["spheres"]["spheres"]) == 4 and len(counts["spheres"]["triangles"]) == 0
    assert len(counts["teapot"]["triangles"]) == 126050 and len(counts["teapot"]["bvh"]) == 81311
    assert len(counts["lamp"]["triangles"]) == 288822 and len(counts["lamp"]["bvh"]) == 189285


def test_bvh_invariants():
    """Q13: children adjacent, leaves are [child2, child1), every triangle in exactly one leaf, boxes contain them."""
    a = rt.load_scene("teapot.scene", cwd=SCENES).arrays()
    n, tri = a["bvh"], a["triangles"].reshape(-1, 4, 3)
    leaf = n["child2"] <= n["child1"]
    inner = ~leaf
    assert (n["child2"][inner] == n["child1"][inner] + 1).all()
    start, end = n["child2"][leaf], n["child1"][leaf]
    order = np.argsort(start)
    assert start[order][0] == 0 and end[order][-1] == len(tri)
    assert (start[order][1:] == end[order][:-1]).all()
    p1 = tri[:, 0]
    corners = np.stack([p1, p1 + tri[:, 1], p1 + tri[:, 2]], 1)  # render form: p1, p2-p1, p3-p1 (scene.cu:1029-1035)
    for s, e, lo, hi in list(zip(start, end, n["min"][leaf], n["max"][leaf]))[:2000]:
        c = corners[s:e]
        assert (c >= lo - 1e-3).all() and (c <= hi + 1e-3).all()
    nrm = np.cross(tri[:, 2], tri[:, 1])
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)
    ok = np.isfinite(tri[:, 3]).all(axis=1)
    assert np.allclose(tri[ok, 3], nrm[ok], atol=2e-3)  # Q9: normal = normalise(cross(p3-p1, p2-p1))


def _write(tmp_path, name, text, newline="\n"):
    p = tmp_path / name
    p.write_bytes(text.replace("\n", newline).encode())
    return str(p)


MINI = """image 8 6 3 2 0.5
sky 0.1 0.2 0.3
camera position 0 0 -5 forward 0 0 1 up 0 1 0 fov 60

material red diffuse 1 0 0 roughness 0.25
material lamp emit 5 5 5 diffuse 0 0 0 specular 0 0 0
material glass ior 1.5
sphere glass 0 0 0 1
triangle red -1 -1 2 1 -1 2 0 1 2
quad lamp -1 2 -1 1 2 -1 1 2 1 -1 2 1
bogus command is ignored
"""


def test_parser_grammar_and_defaults(tmp_path):
    s = rt.load_scene(_write(tmp_path, "mini.scene", MINI))
    a = s.arrays()
    assert (a["width"], a["height"], a["ray_count"], a["bounces"]) == (8, 6, 3, 2) and a["exposure"] == np.float32(0.5)
    assert a["env_w"] == 1 and a["env_h"] == 1 and np.allclose(a["env"], [[0.1, 0.2, 0.3]])
    assert len(a["spheres"]) == 1 and len(a["triangles"]) == 3  # triangle + quad = (p1,p2,p3),(p1,p3,p4)  (Q15)
    m = a["materials"]  # diffuse[3] metallicity specular[3] roughness emitted[3] ior; defaults diffuse 1 specular 1 rest 0
    assert np.allclose(m[0], [1, 0, 0, 0, 1, 1, 1, 0.25, 0, 0, 0, 0])
    assert np.allclose(m[1], [0, 0, 0, 0, 0, 0, 0, 0, 5, 5, 5, 0])
    assert np.allclose(m[2], [1, 1, 1, 0, 1, 1, 1, 0, 0, 0, 0, 1.5])
    assert a["material_indices"][0] == 2  # the sphere comes first in primitive order
    assert sorted(a["material_indices"][1:]) == [0, 1, 1]
    assert np.allclose(a["camera_position"], [0, 0, -5])
    # precompute_camera_data (scene.cu:62-76): near-plane height 2 tan(fov/2), width by aspect, inv = 1/(size - 1)
    assert np.isclose(np.linalg.norm(a["scaled_up"]), 2 * np.tan(np.radians(60) / 2), rtol=1e-6)
    assert np.isclose(np.linalg.norm(a["scaled_right"]), 2 * np.tan(np.radians(60) / 2) * 8 / 6, rtol=1e-6)
    assert a["inv_width"] == np.float32(1 / 7) and a["inv_height"] == np.float32(1 / 5)


def test_crlf_scene_equals_lf_scene(tmp_path):
    """the reference's parser keeps the trailing CR on names and paths; the drop-in strips it."""
    a = rt.load_scene(_write(tmp_path, "lf.scene", MINI)).arrays()
    b = rt.load_scene(_write(tmp_path, "crlf.scene", MINI, newline="\r\n")).arrays()
    for k in ("spheres", "triangles", "material_indices", "materials", "bvh", "env"):
        assert a[k].tobytes() == b[k].tobytes(), k


def test_set_image_recomputes_camera():
    s = rt.load_scene("cornell.scene", cwd=SCENES)
    a = s.arrays()
    b = s.set_image(200, 100, 7, 3, 2.0).arrays()
    assert (b["width"], b["height"], b["ray_count"], b["bounces"]) == (200, 100, 7, 3) and b["exposure"] == np.float32(2.0)
    assert b["inv_width"] == np.float32(1.0 / 199) and b["inv_height"] == np.float32(1.0 / 99)
    assert np.isclose(np.linalg.norm(b["scaled_right"]) / np.linalg.norm(b["scaled_up"]), 2.0, rtol=1e-6)
    assert a["triangles"].tobytes() == b["triangles"].tobytes()
    c = s.set_image(300, 300, 1, 1).arrays()  # exposure < 0 keeps the current value
    assert c["exposure"] == np.float32(2.0)


def test_error_codes_instead_of_exceptions(tmp_path):
    with pytest.raises(rt.RtError) as e:
        rt.load_scene(str(tmp_path / "does_not_exist.scene"))
    assert e.value.status == -5 and "cannot open scene file" in str(e.value)
    with pytest.raises(rt.RtError) as e:
        rt.load_scene(_write(tmp_path, "undef.scene", "image 4 4 1 1 1\nsky 0 0 0\nsphere nosuch 0 0 0 1\n"))
    assert e.value.status == -6 and "used before definition" in str(e.value)
    with pytest.raises(rt.RtError) as e:
        rt.load_scene(_write(tmp_path, "noply.scene", "material m\nply m missing.ply\n"))
    assert e.value.status == -5 and "PLY" in str(e.value)
    with pytest.raises(rt.RtError) as e:
        rt.load_scene(_write(tmp_path, "nopfm.scene", "sky_map missing.pfm\n"))
    assert e.value.status == -5 and "PFM" in str(e.value)


def test_ply_and_pfm_loaders(tmp_path):
    """Fixed-layout binary PLY (8 floats per vertex, uint8 count + int32 indices per face, fan triangulation,
    scene.cu:491-546) and PFM (3 header lines + raw float3 texels, no flip, scene.cu:548-567)."""
    verts = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0.5, 2, 0]], np.float32)
    v8 = np.zeros((5, 8), np.float32)
    v8[:, :3] = verts
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex 5\n" +
              "".join(f"property float {n}\n" for n in ("x", "y", "z", "nx", "ny", "nz", "u", "v")) +
              "element face 2\nproperty list uint8 int vertex_indices\nend_header\n")
    faces = struct.pack("<B4i", 4, 0, 1, 2, 3) + struct.pack("<B3i", 3, 3, 2, 4)
    (tmp_path / "m.ply").write_bytes(header.encode() + v8.tobytes() + faces)
    env = np.arange(2 * 2 * 3, dtype=np.float32).reshape(4, 3)
    (tmp_path / "e.pfm").write_bytes(b"PF\n2 2\n-1.0\n" + env.tobytes())
    path = _write(tmp_path, "p.scene", "image 4 4 1 1 1\nsky_map e.pfm\nmaterial m\nply m m.ply\n")
    a = rt.load_scene(path, cwd=str(tmp_path)).arrays()
    assert len(a["triangles"]) == 3  # quad fan (0,1,2),(0,2,3) + one triangle
    assert a["env_w"] == 2 and a["env_h"] == 2 and a["env"].tobytes() == env.tobytes()
    tri = a["triangles"].reshape(-1, 4, 3)
    got = sorted(tuple(np.round(np.concatenate([t[0], t[0] + t[1], t[0] + t[2]]), 4)) for t in tri)
    want = sorted(tuple(np.concatenate([verts[i], verts[j], verts[k]])) for i, j, k in ((0, 1, 2), (0, 2, 3), (3, 2, 4)))
    assert np.allclose(got, want)
    (tmp_path / "bad.ply").write_bytes(header.encode() + v8.tobytes()[:-8])
    with pytest.raises(rt.RtError):
        rt.load_scene(_write(tmp_path, "b.scene", "material m\nply m bad.ply\n"), cwd=str(tmp_path))


def test_png_writer_roundtrip(tmp_path):
    """rt_write_png replaces stbi_write_png (raytracing.cu:395): 8-bit RGB, rows top to bottom. Parity is on decoded
    pixels, so decode with zlib and compare."""
    rs = np.random.RandomState(0)
    img = rs.randint(0, 256, size=(37, 53, 3), dtype=np.uint8)
    path = str(tmp_path / "x.png")
    rt.write_png(path, img)
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, ihdr = 8, b"", None
    while pos < len(data):
        n, kind = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(kind + body)
        if kind == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        if kind == b"IDAT":
            idat += body
        pos += 12 + n
    assert ihdr == (53, 37, 8, 2, 0, 0, 0)
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(37, 1 + 53 * 3)
    assert (raw[:, 0] == 0).all()  # filter type 0 on every row
    assert (raw[:, 1:].reshape(37, 53, 3) == img).all()


@pytest.mark.parametrize("name", ["cornell_plus", "teapot"])
def test_spatial_split_tree_gives_the_same_hits(name, monkeypatch):
    """use_bvh="sbvh" (rt_scene_load(use_bvh = 3)) selects the spatial-split builder (bvh_builder.cpp; not the reference's tree):
    triangles that straddle a split plane are referenced from both children. Whatever the tree, closest hits must not
    change: same hit / miss per ray and the same t as on the reference's tree (walked by the oracle), arrays consistent
    (every leaf range inside the triangle array, materials carried along with the repeated triangles)."""
    from util import oracle_arrays, oracle_closest, scene_rays
    monkeypatch.delenv("RT_SBVH", raising=False)
    ref = rt.load_scene(name + ".scene", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    A0 = oracle_arrays(ref)
    sb = rt.load_scene(name + ".scene", use_bvh="sbvh", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    A1 = oracle_arrays(sb)
    assert len(A1.triangles) >= len(A0.triangles) and len(A1.material_indices) == len(A1.spheres) + len(A1.triangles)
    leaves = A1.bvh[A1.bvh["child2"] <= A1.bvh["child1"]] if A1.bvh.dtype.names and "child1" in A1.bvh.dtype.names else None
    if leaves is not None:
        assert (leaves["child2"] >= 0).all() and (leaves["child1"] <= len(A1.triangles)).all()
    rays = scene_rays(A0, 30_000, seed=13)
    p0, t0, c0 = oracle_closest(A0, rays, 1)
    p1, t1, c1 = oracle_closest(A1, rays, 1)
    assert ((p0 >= 0) == (p1 >= 0)).all()
    hit = p0 >= 0
    assert np.allclose(t0[hit], t1[hit], rtol=1e-6, atol=0)
    ns = len(A0.spheres)
    tri = hit & (p0 >= ns)
    same_material = A0.material_indices[p0[tri]] == A1.material_indices[p1[tri]]
    assert same_material.mean() > 0.999  # a different triangle only on exact-t ties between neighbours
