"""CPU: the oracle against the committed golden fixtures and against itself
(brute force vs its BVH path, crops vs full frames, thread counts)."""
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as O
from simpleraytracer_b200 import api

GOLD = Path(__file__).resolve().parent / "golden"


def load_cornell(name):
    g = np.load(GOLD / name)
    cam, light = api.srt_camera(), api.srt_light()
    cam.origin[:] = g["cam_origin"]; cam.forward[:] = g["cam_forward"]
    cam.right[:] = g["cam_right"]; cam.up[:] = g["cam_up"]
    cam.half_w, cam.half_h = [float(x) for x in g["cam_half"]]
    light.position[:] = g["light_pos"]; light.intensity[:] = g["light_int"]
    w, h, spp, jitter, seed, shade = [int(x) for x in g["params"]]
    p = api.make_params(w, h, spp=spp, jitter=jitter, seed=seed, shade=shade,
                        t_min=float(g["t_min"]), shadow_eps=float(g["shadow_eps"]))
    return g, cam, light, p


@pytest.mark.parametrize("name", ["cornell_8_shade0_spp1.npz", "cornell_64_shade0_spp1.npz",
                                  "cornell_64_shade2_spp4.npz"])
def test_oracle_matches_golden_cornell(name):
    g, cam, light, p = load_cornell(name)
    rgb, hid, ht, counts, _ = O.render(g["vertices"], g["albedo"], cam, light, p)
    assert np.array_equal(hid, g["hit_id"])
    assert np.array_equal(ht, g["hit_t"])
    assert np.array_equal(rgb, g["rgb"])
    assert np.array_equal(counts, g["counts"])
    # the generator still produces the stored scene
    v, a, _, _ = api.gen_cornell_box(p.width, p.height)
    assert np.array_equal(v, g["vertices"]) and np.array_equal(a, g["albedo"])


def test_oracle_matches_golden_mesh():
    g = np.load(GOLD / "mesh_2000_48x27.npz")
    w, h, spp, jitter, seed, shade, mesh_seed, n = [int(x) for x in g["params"]]
    v, a, cam, light = api.gen_mesh(mesh_seed, n, w, h)
    assert float(v.astype(np.float64).sum()) == float(g["vertices_sum"])
    p = api.make_params(w, h, spp=spp, jitter=jitter, seed=seed, shade=shade, background=(0.5, 0.7, 0.9))
    for use_bvh in (False, True):
        rgb, hid, ht, counts, _ = O.render(v, a, cam, light, p, use_bvh=use_bvh)
        assert np.array_equal(hid, g["hit_id"])
        assert np.array_equal(ht, g["hit_t"])
        assert np.array_equal(rgb, g["rgb"])
        assert np.array_equal(counts, g["counts"])


def test_oracle_degenerate_cases():
    g = np.load(GOLD / "degenerate.npz")
    for name in ["shared_edge", "shared_vertex", "parallel", "coplanar_dup", "miss", "behind", "backface"]:
        res = [O.intersect(g[name + "_o"], g[name + "_d"], t) for t in g[name + "_tris"]]
        assert [r[0] for r in res] == list(g[name + "_inside"]), name
        t = np.array([r[1] if r[0] else np.inf for r in res], np.float32)
        assert np.array_equal(t, g[name + "_t"]), name
    # the semantics the fixture freezes
    assert list(g["shared_edge_inside"]) == [True, True] and g["shared_edge_t"][0] == g["shared_edge_t"][1]
    assert list(g["parallel_inside"]) == [False]
    assert list(g["miss_inside"]) == [False, False]
    assert g["behind_t"][0] < 0          # inside but t < t_min: rejected by the closest-hit rule
    assert list(g["backface_inside"]) == [True]  # no culling


def test_oracle_tie_break_lower_id_wins():
    tri = [0, 0, 0, 1, 0, 0, 0, 1, 0]
    v = np.array([tri, tri, tri], np.float32)
    cam = api.srt_camera()
    cam.origin[:] = (0.25, 0.25, 2); cam.forward[:] = (0, 0, -1); cam.right[:] = (1, 0, 0); cam.up[:] = (0, 1, 0)
    cam.half_w = cam.half_h = 0.01
    light = api.srt_light()
    p = api.make_params(2, 2)
    _, hid, ht, _, _ = O.render(v, None, cam, light, p)
    assert (hid == 0).all() and np.isfinite(ht).all()
    _, hid2, _, _, _ = O.render(v[::-1].copy(), None, cam, light, p, use_bvh=True)
    assert (hid2 == 0).all()


def test_oracle_crop_band_and_threads_consistent():
    v, a, cam, light = api.gen_mesh(5, 3000, 40, 30)
    p = api.make_params(40, 30, spp=3, jitter=1, seed=2, shade=2, background=(0.1, 0.2, 0.3))
    full = O.render(v, a, cam, light, p, threads=1)
    multi = O.render(v, a, cam, light, p, threads=4, use_bvh=True)
    for k in range(3):
        assert np.array_equal(full[k], multi[k])
    crop = O.render(v, a, cam, light, p, crop=(7, 23, 5, 19), use_bvh=True)
    for k in range(3):
        assert np.array_equal(crop[k], full[k][5:19, 7:23])
    assert int(full[3][0]) == 40 * 30 * 3 and 0 < int(full[3][1]) <= 40 * 30 * 3


def test_oracle_empty_scene_is_background():
    cam = api.srt_camera(); cam.forward[:] = (0, 0, -1); cam.right[:] = (1, 0, 0); cam.up[:] = (0, 1, 0)
    cam.half_w = cam.half_h = 0.5
    p = api.make_params(5, 3, background=(0.25, 0.5, 0.75), shade=2)
    rgb, hid, ht, counts, _ = O.render(np.zeros((0, 9), np.float32), None, cam, api.srt_light(), p)
    assert (hid == -1).all() and np.isinf(ht).all()
    assert np.allclose(rgb, [0.25, 0.5, 0.75]) and int(counts[1]) == 0
