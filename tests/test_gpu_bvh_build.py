"""GPU tests of the device BVH build (csrc/device/bvh_build.cuh).

The reference builds its tree on the host (generate_bvh / maybe_split, scene.cu:866-1036); the product's host builder
is pinned to it byte for byte (tests/test_host_scene.py). The device builder takes the same decisions level by level, so
it must produce the same tree as a SET of nodes — same boxes, same triangle sets per leaf — only the numbering of the
nodes and the order of the (at most four) triangles inside a leaf differ. Closest-hit distances are then bit-identical.
"""
import numpy as np
import pytest

import cuda_raytracer_b200 as rt
from util import SCENES, oracle_arrays, scene_rays

pytestmark = pytest.mark.gpu
MESH_SCENES = ["cornell", "cornell_plus", "teapot", "glass_teapot", "lamp"]


@pytest.fixture(scope="module")
def renderer():
    r = rt.Renderer(0)
    yield r
    r.close()


def _subtree_counts(nodes):
    """Triangles below every node; children always carry larger indices than their parent in both builders."""
    leaf = nodes["child2"] <= nodes["child1"]
    count = np.where(leaf, nodes["child1"] - nodes["child2"], 0).astype(np.int64)
    for i in range(len(nodes) - 1, -1, -1):
        if not leaf[i]:
            count[i] = count[nodes["child1"][i]] + count[nodes["child2"][i]]
    return count, leaf


def _node_signatures(nodes):
    count, leaf = _subtree_counts(nodes)
    rows = np.concatenate([nodes["min"].astype(np.float64), nodes["max"].astype(np.float64), count[:, None].astype(np.float64),
                           leaf[:, None].astype(np.float64)], axis=1)
    return rows[np.lexsort(rows.T[::-1])]


def _leaf_rows(nodes, tri_keys_in_bvh_order):
    """One row per triangle: its 9 defining floats followed by the box of the leaf that holds it, sorted."""
    leaf = np.nonzero(nodes["child2"] <= nodes["child1"])[0]
    start, end = nodes["child2"][leaf], nodes["child1"][leaf]
    owner = np.zeros(len(tri_keys_in_bvh_order), np.int64)
    for s, e, node in zip(start, end, leaf):
        owner[s:e] = node
    rows = np.concatenate([tri_keys_in_bvh_order.astype(np.float64), nodes["min"][owner].astype(np.float64),
                           nodes["max"][owner].astype(np.float64)], axis=1)
    return rows[np.lexsort(rows.T[::-1])]


def _keys_from_build_form(tris12):
    """build form {p1, p2, p3, centroid} -> {p1, p2 - p1, p3 - p1} in float32, the render form's first nine floats."""
    t = tris12.reshape(-1, 4, 3)
    return np.concatenate([t[:, 0], t[:, 1] - t[:, 0], t[:, 2] - t[:, 0]], axis=1).astype(np.float32)


@pytest.mark.parametrize("name", MESH_SCENES)
def test_device_tree_equals_host_tree_as_a_set(renderer, name):
    host = rt.load_scene(name + ".scene", cwd=SCENES)
    dev = rt.load_scene(name + ".scene", use_bvh="device", cwd=SCENES)
    ha, da = host.arrays(), dev.arrays()
    assert da["triangles_form"] == 1 and ha["triangles_form"] == 0 and len(da["bvh"]) == 1
    assert da["bvh"][0]["min"].tobytes() == ha["bvh"][0]["min"].tobytes()  # root box (scene bounds) computed on the host
    assert da["min_coord"].tobytes() == ha["min_coord"].tobytes() and da["inv_dimensions"].tobytes() == ha["inv_dimensions"].tobytes()

    renderer.upload(dev)
    nodes, order = renderer.download_bvh(len(da["triangles"]))
    assert renderer.bvh_device_ms > 0
    assert len(nodes) == len(ha["bvh"]), "node count"
    assert sorted(order.tolist()) == list(range(len(order))), "order must be a permutation of the file-order ids"
    internal = nodes["child2"] > nodes["child1"]
    assert (nodes["child2"][internal] == nodes["child1"][internal] + 1).all()  # adjacent pairs (Q13), the packer relies on it
    assert np.array_equal(_node_signatures(nodes), _node_signatures(ha["bvh"])), "boxes / subtree sizes / leaf flags"
    host_rows = _leaf_rows(ha["bvh"], ha["triangles"][:, :9])
    dev_rows = _leaf_rows(nodes, _keys_from_build_form(da["triangles"])[order])
    assert np.array_equal(host_rows, dev_rows), "every triangle sits in a leaf with the same box as on the host"
    # inside a leaf the device keeps file order (stable partition)
    leaf = np.nonzero(~internal)[0][:5000]
    assert all((np.diff(order[nodes["child2"][i]:nodes["child1"][i]]) > 0).all() for i in leaf)
    print(f"{name}: {len(order)} triangles, {len(nodes)} nodes, device build {renderer.bvh_device_ms:.2f} ms, host build {host.bvh_build_ms:.1f} ms")


@pytest.mark.parametrize("name", MESH_SCENES)
def test_closest_hit_is_bit_identical_on_both_trees(renderer, name):
    host = rt.load_scene(name + ".scene", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    dev = rt.load_scene(name + ".scene", use_bvh="device", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    ha, da = host.arrays(), dev.arrays()
    rays = scene_rays(oracle_arrays(host), 200_000, seed=31)
    renderer.upload(host)
    prim_h, t_h = renderer.trace_closest(rays, with_spheres=True, stats=True)
    visits_h = renderer.last_stats.internal_visits
    renderer.upload(dev)
    prim_d, t_d = renderer.trace_closest(rays, with_spheres=True, stats=True)
    visits_d = renderer.last_stats.internal_visits
    _, order = renderer.download_bvh(len(da["triangles"]))
    assert t_d.tobytes() == t_h.tobytes()
    assert visits_d == visits_h  # same boxes, same near/far roles: the walk itself is the same
    ns = len(ha["spheres"])
    hit = prim_h >= ns
    assert ((prim_d >= ns) == hit).all() and (prim_d[~hit] == prim_h[~hit]).all()
    key_h = ha["triangles"][:, :9][prim_h[hit] - ns]
    key_d = _keys_from_build_form(da["triangles"])[order][prim_d[hit] - ns]
    same = (key_h == key_d).all(axis=1)
    assert same.mean() > 0.9999  # a different triangle only where two of one leaf are hit at exactly the same t
    assert (hit.sum() > 1000)


def test_images_agree_and_empty_scene_builds(renderer):
    for name, w, h in (("lamp", 96, 64), ("spheres", 64, 48)):  # spheres.scene has no triangle at all (Q12)
        host = rt.load_scene(name + ".scene", cwd=SCENES).set_image(w, h, 40, 8, 1.0)
        dev = rt.load_scene(name + ".scene", use_bvh="device", cwd=SCENES).set_image(w, h, 40, 8, 1.0)
        a = renderer.upload(host).render(sort=False)
        n_host = renderer.last_stats.live_ray_bounces
        b = renderer.upload(dev).render(sort=False)
        assert abs(renderer.last_stats.live_ray_bounces - n_host) <= 1e-4 * n_host
        assert np.allclose(a, b, rtol=1e-3, atol=1e-3 * a.mean())
        nodes, _ = renderer.download_bvh(len(dev.arrays()["triangles"]))
        assert len(nodes) == len(host.arrays()["bvh"])
    with pytest.raises(rt.RtError):
        renderer.upload(rt.load_scene("cornell.scene", cwd=SCENES))
        renderer.download_bvh(32)  # host-built scene: nothing to download


def test_oversized_leaves_fall_back_to_the_host_packer(renderer, tmp_path):
    """Twenty coincident quads: no plane through their 40 triangles beats the leaf cost (scene.cu:959), so the tree holds a
    leaf of 40 > 15 triangles. Device packing hands over to the host packer, which chains such leaves."""
    pile = "quad white -100 50 -100 -100 50 100 100 50 100 100 50 -100\n" * 20
    text = ("material light diffuse 0 0 0 specular 0 0 0 emit 20 20 20\nmaterial white diffuse 0.7 0.7 0.7\n"
            "quad light -50 300 -50 -50 300 50 50 300 50 50 300 -50\n" + pile +
            "quad white -500 0 -500 -500 0 500 500 0 500 500 0 -500\nquad white -300 0 200 -300 90 200 300 90 200 300 0 200\n"
            "sky 0.3 0.4 0.6\ncamera position 0 150 -600 forward 0 -0.1 1 up 0 1 0 fov 40\nimage 64 48 40 6 1\n")
    (tmp_path / "pile.scene").write_text(text)
    host = rt.load_scene("pile.scene", cwd=str(tmp_path))
    dev = rt.load_scene("pile.scene", use_bvh="device", cwd=str(tmp_path))
    ha = host.arrays()
    leaf = ha["bvh"]["child2"] <= ha["bvh"]["child1"]
    assert (ha["bvh"]["child1"][leaf] - ha["bvh"]["child2"][leaf]).max() == 40
    rays = scene_rays(oracle_arrays(host), 50_000, seed=7)
    renderer.upload(host)
    prim_h, t_h = renderer.trace_closest(rays, with_spheres=True, stats=True)
    tests_h = renderer.last_stats.triangle_tests
    renderer.upload(dev)
    prim_d, t_d = renderer.trace_closest(rays, with_spheres=True, stats=True)
    assert t_d.tobytes() == t_h.tobytes() and renderer.last_stats.triangle_tests == tests_h
    nodes, order = renderer.download_bvh(len(dev.arrays()["triangles"]))
    assert np.array_equal(_node_signatures(nodes), _node_signatures(ha["bvh"]))
