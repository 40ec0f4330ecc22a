"""GPU parity tests of the render loop: the CUDA path (through the C ABI,
srtRender / srtRenderDevice / srtRenderToImage) against the CPU oracle and the
committed golden fixtures.  Bar (BASELINE.json north_star): hit ids bit-exact,
hit distance and radiance within 1e-5 relative."""
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as O
from simpleraytracer_b200 import api
from test_oracle_golden import load_cornell

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
REL = 1e-5  # tolerance on hit_t and rgb, relative


def assert_parity(gpu, ref, what=""):
    rgb, hid, ht = gpu[:3]
    rrgb, rhid, rht = ref[:3]
    assert np.array_equal(hid, rhid), f"{what}: {np.count_nonzero(hid != rhid)} hit ids differ"
    finite = np.isfinite(rht)
    assert np.array_equal(np.isfinite(ht), finite), what
    np.testing.assert_allclose(ht[finite], rht[finite], rtol=REL, atol=0, err_msg=what)
    np.testing.assert_allclose(rgb, rrgb, rtol=REL, atol=1e-7, err_msg=what)


@pytest.mark.parametrize("name", ["cornell_8_shade0_spp1.npz", "cornell_64_shade0_spp1.npz",
                                  "cornell_64_shade2_spp4.npz"])
@pytest.mark.parametrize("accel", [api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH])
def test_golden_cornell(ctx, name, accel):
    g, cam, light, p = load_cornell(name)
    p.accel = accel
    scene = ctx.create_scene(g["vertices"], g["albedo"])
    rgb, hid, ht, stats = scene.render(cam, light, p)
    assert_parity((rgb, hid, ht), (g["rgb"], g["hit_id"], g["hit_t"]), name)
    # in practice the spec makes the results bit-identical, not merely close
    assert np.array_equal(ht, g["hit_t"]) and np.array_equal(rgb, g["rgb"])
    assert stats.primary_rays == int(g["counts"][0]) and stats.shadow_rays == int(g["counts"][1])
    scene.release()


def test_config2_cornell_512_vs_oracle(ctx):
    """BASELINE config 2: ~12-triangle Cornell box, 512x512, 1 spp, pixel-centre rays."""
    v, a, cam, light = api.gen_cornell_box(512, 512)
    scene = ctx.create_scene(v, a)
    for shade in (api.SRT_SHADE_ALBEDO, api.SRT_SHADE_SHADOW):
        p = api.make_params(512, 512, shade=shade)
        ref = O.render(v, a, cam, light, p)
        for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
            p.accel = accel
            got = scene.render(cam, light, p)
            assert_parity(got, ref, f"cornell512 shade{shade} accel{accel}")
            assert got[3].primary_rays == 512 * 512
            assert got[3].shadow_rays == int(ref[3][1])
    scene.release()


def test_golden_mesh_and_modes(ctx):
    g = np.load(GOLD / "mesh_2000_48x27.npz")
    w, h, spp, jitter, seed, shade, mesh_seed, n = [int(x) for x in g["params"]]
    v, a, cam, light = api.gen_mesh(mesh_seed, n, w, h)
    scene = ctx.create_scene(v, a)
    p = api.make_params(w, h, spp=spp, jitter=jitter, seed=seed, shade=shade, background=(0.5, 0.7, 0.9))
    for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
        p.accel = accel
        got = scene.render(cam, light, p)
        assert_parity(got, (g["rgb"], g["hit_id"], g["hit_t"]), f"mesh accel{accel}")
        assert got[3].shadow_rays == int(g["counts"][1])
    scene.release()


@pytest.mark.parametrize("shade", [api.SRT_SHADE_ALBEDO, api.SRT_SHADE_LAMBERT, api.SRT_SHADE_SHADOW])
def test_mesh_20k_brute_oracle_crop(ctx, shade):
    """Brute-force oracle on a crop of a larger frame (the oracle is O(rays x triangles))."""
    W, H = 320, 180
    v, a, cam, light = api.gen_mesh(2, 20000, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=2, jitter=1, seed=5, shade=shade, background=(0.2, 0.3, 0.4))
    crop = (100, 164, 60, 100)
    ref = O.render(v, a, cam, light, p, crop=crop)
    for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
        p.accel = accel
        rgb, hid, ht, _ = scene.render(cam, light, p)
        x0, x1, y0, y1 = crop
        assert_parity((rgb[y0:y1, x0:x1], hid[y0:y1, x0:x1], ht[y0:y1, x0:x1]), ref, f"accel{accel}")
    scene.release()


def test_bvh_equals_brute_force_100k(ctx):
    """Size-independent property at a BASELINE-sized scene: the BVH only culls, so the
    BVH render equals the brute-force render bit for bit (100K triangles, config 3 scene)."""
    W, H = 480, 270
    v, a, cam, light = api.gen_mesh(1, 100000, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=2, jitter=1, seed=3, shade=api.SRT_SHADE_SHADOW, background=(0.5, 0.7, 0.9))
    p.collect_stats = 1
    p.accel = api.SRT_ACCEL_BRUTE
    brute = scene.render(cam, light, p)
    p.accel = api.SRT_ACCEL_BVH
    bvh = scene.render(cam, light, p)
    assert np.array_equal(brute[1], bvh[1])
    assert np.array_equal(brute[2], bvh[2])
    assert np.array_equal(brute[0], bvh[0])
    assert brute[3].shadow_rays == bvh[3].shadow_rays > 0
    assert bvh[3].tri_tests < brute[3].tri_tests / 100      # the BVH actually culls
    # and the oracle's BVH path agrees on a band of the same frame
    ref = O.render(v, a, cam, light, p, crop=(0, W, 128, 136), use_bvh=True)
    assert_parity((bvh[0][128:136], bvh[1][128:136], bvh[2][128:136]), ref, "oracle band")
    scene.release()


@pytest.mark.parametrize("spp", [8, 16, 40, 64])
def test_sample_parallel_bvh_kernel_vs_oracle(ctx, spp):
    """spp >= 8 selects the sample-parallel BVH kernel (8 / 16 / 32 sample lanes per pixel):
    samples traced by different lanes, radiance still summed in sample order => identical to the
    oracle's sequential loop.  40 spp is ragged against the 32 lanes; W = 61 is ragged against the
    8-pixel tile; the interleaved launch owns every 2nd 8-row block."""
    W, H = 61, 43
    v, a, cam, light = api.gen_mesh(7, 4000, W, H)
    scene = ctx.create_scene(v, a)
    common = dict(spp=spp, jitter=1, seed=9, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                  background=(0.5, 0.7, 0.9))
    p = api.make_params(W, H, **common)
    p.collect_stats = 1
    ref = O.render(v, a, cam, light, p, use_bvh=True)
    got = scene.render(cam, light, p)
    assert_parity(got, ref, f"spp{spp}")
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[2], ref[2])     # bit-identical in practice
    assert got[3].primary_rays == W * H * spp and got[3].shadow_rays == int(ref[3][1])
    half = scene.render(cam, light, api.make_params(W, H, row_block_stride=2, row_block_phase=1, **common))
    from simpleraytracer_b200.sharding import interleaved_rows
    rows = interleaved_rows(H, 2, 1)
    assert np.array_equal(half[0][rows], ref[0][rows]) and np.array_equal(half[1][rows], ref[1][rows])
    scene.release()


def test_row_bands_compose_the_full_frame(ctx):
    """Multi-GPU sharding contract: rendering bands separately equals rendering the frame."""
    W, H = 200, 120
    v, a, cam, light = api.gen_mesh(4, 5000, W, H)
    scene = ctx.create_scene(v, a)
    full_p = api.make_params(W, H, spp=2, jitter=1, seed=1, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH)
    full = scene.render(cam, light, full_p)
    rgb = np.zeros((H, W, 3), np.float32); hid = np.zeros((H, W), np.int32); ht = np.zeros((H, W), np.float32)
    from simpleraytracer_b200.sharding import row_bands
    rays = 0
    for y0, y1 in row_bands(H, 3):
        p = api.make_params(W, H, spp=2, jitter=1, seed=1, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                            y0=y0, y1=y1, buf_y0=y0)
        r, i, t, st = scene.render(cam, light, p)
        rgb[y0:y1], hid[y0:y1], ht[y0:y1] = r[: y1 - y0], i[: y1 - y0], t[: y1 - y0]
        rays += st.primary_rays
    assert rays == W * H * 2
    assert np.array_equal(rgb, full[0]) and np.array_equal(hid, full[1]) and np.array_equal(ht, full[2])
    scene.release()


@pytest.mark.parametrize("accel", [api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH])
def test_interleaved_row_blocks_compose_the_full_frame(ctx, accel):
    """Interleaved multi-GPU sharding (row_block_stride/phase): N launches that each own every
    N-th 8-row block write disjoint rows of ONE device frame and together equal the full render."""
    import torch
    from simpleraytracer_b200.sharding import interleaved_rows
    W, H, N = 150, 77, 3            # ragged: 77 rows = 9 full blocks + 5 rows
    v, a, cam, light = api.gen_mesh(4, 5000, W, H)
    scene = ctx.create_scene(v, a)
    common = dict(spp=2, jitter=1, seed=1, shade=api.SRT_SHADE_SHADOW, accel=accel)
    full = scene.render(cam, light, api.make_params(W, H, **common))
    frame = torch.full((H, W, 3), -7.0, device="cuda")
    stream = torch.cuda.current_stream()
    torch.cuda.synchronize()
    rays = 0
    for r in range(N):
        p = api.make_params(W, H, row_block_stride=N, row_block_phase=r, **common)
        scene.render_device(cam, light, p, frame.data_ptr(), 0, 0, 0)
        rays += scene.stats().primary_rays
        rows = interleaved_rows(H, N, r)
        got = frame.cpu().numpy()
        assert np.array_equal(got[rows], full[0][rows])
    assert rays == W * H * 2
    assert np.array_equal(frame.cpu().numpy(), full[0])
    # host-buffer entry point: only the owned rows are written
    host = scene.render(cam, light, api.make_params(W, H, row_block_stride=N, row_block_phase=1, **common))
    rows = interleaved_rows(H, N, 1)
    other = [y for y in range(H) if y not in set(rows)]
    assert np.array_equal(host[0][rows], full[0][rows]) and np.array_equal(host[1][rows], full[1][rows])
    assert (host[0][other] == 0).all() and (host[1][other] == -2).all()
    p = api.make_params(W, H, y0=4, row_block_stride=2, row_block_phase=0, **common)
    with pytest.raises(api.MLError, match="interleaved"):
        scene.render(cam, light, p)
    scene.release()


@pytest.mark.parametrize("shape", [
    dict(W=64, H=300, y0=0, y1=300, stride=1, spp=8),     # 8 chunks of 40 rows, last one short
    dict(W=96, H=301, y0=13, y1=293, stride=1, spp=1),    # a band that starts off the 8-row grid
    dict(W=64, H=203, y0=0, y1=203, stride=2, spp=8),     # interleaved: chunks are whole 16-row groups
    dict(W=40, H=500, y0=0, y1=500, stride=5, spp=1),     # 40-row groups, 13 of them over 8 chunks
    dict(W=33, H=63, y0=0, y1=63, stride=1, spp=2),       # below the chunking threshold: one piece
])
def test_host_render_pipeline_equals_device_render(ctx, shape):
    """srtRender (host buffers) renders, copies and scatters the band in up to 8 overlapped chunks on
    alternating streams; every owned pixel must equal the one-launch device render, the rest stays untouched."""
    import torch
    from simpleraytracer_b200.sharding import interleaved_rows
    W, H = shape["W"], shape["H"]
    v, a, cam, light = api.gen_mesh(9, 3000, W, H)
    scene = ctx.create_scene(v, a)
    common = dict(spp=shape["spp"], jitter=1 if shape["spp"] > 1 else 0, seed=3, shade=api.SRT_SHADE_SHADOW,
                  accel=api.SRT_ACCEL_BVH)
    frame = torch.zeros((H, W, 3), device="cuda")
    torch.cuda.synchronize()
    scene.render_device(cam, light, api.make_params(W, H, **common), frame.data_ptr(), 0, 0, 0)
    scene.stats()                      # srtRenderDevice is asynchronous on the context's stream: this waits for it
    full = frame.cpu().numpy()
    for phase in range(shape["stride"]):
        p = api.make_params(W, H, y0=shape["y0"], y1=shape["y1"], row_block_stride=shape["stride"],
                            row_block_phase=phase, **common)
        rgb, hid, ht, _ = scene.render(cam, light, p)
        own = [y for y in (interleaved_rows(H, shape["stride"], phase) if shape["stride"] > 1 else range(H))
               if shape["y0"] <= y < shape["y1"]]
        other = [y for y in range(H) if y not in set(own)]
        assert np.array_equal(rgb[own], full[own]), f"phase {phase}"
        assert (rgb[other] == 0).all() and (hid[other] == -2).all()
        assert (hid[own] >= -1).all()
        # the same into page-locked buffers: filled straight from the device, no staging / host scatter
        pin = [torch.zeros(s_, dtype=t_).pin_memory() for s_, t_ in
               (((H, W, 3), torch.float32), ((H, W), torch.int32), ((H, W), torch.float32))]
        pin[1].fill_(-2)
        prgb, phid, pht, _ = scene.render(cam, light, p, buffers=tuple(t.numpy() for t in pin))
        assert np.array_equal(prgb, rgb) and np.array_equal(phid, hid) and np.array_equal(pht, ht), f"phase {phase}"
    scene.release()


def test_degenerate_cases_on_gpu(ctx):
    """Tie-breaks and rejections of the golden single-ray cases, through a 1x1 render."""
    g = np.load(GOLD / "degenerate.npz")
    for name in ["shared_edge", "shared_vertex", "parallel", "coplanar_dup", "miss", "behind", "backface"]:
        tris, o, d = g[name + "_tris"], g[name + "_o"], g[name + "_d"]
        cam = api.srt_camera()
        cam.origin[:] = o; cam.forward[:] = d
        # any orthonormal completion; with half_w = half_h = 0 the pixel-centre ray is exactly `forward`
        right = np.cross(d, [0.0, 1.0, 0.0]); right = right if np.linalg.norm(right) > 0 else np.array([1.0, 0, 0])
        cam.right[:] = right / np.linalg.norm(right)
        cam.up[:] = np.cross(cam.right[:], d)
        cam.half_w = cam.half_h = 0.0
        p = api.make_params(1, 1, t_min=1e-4)
        ref = O.render(tris, None, cam, api.srt_light(), p)
        scene = ctx.create_scene(tris)
        for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
            p.accel = accel
            _, hid, ht, _ = scene.render(cam, api.srt_light(), p)
            assert hid[0, 0] == ref[1][0, 0], name
            assert (ht[0, 0] == ref[2][0, 0]) or (np.isinf(ht[0, 0]) and np.isinf(ref[2][0, 0])), name
        inside, t = g[name + "_inside"], g[name + "_t"]
        cands = [(t[i], i) for i in range(len(t)) if inside[i] and t[i] > 1e-4]
        assert hid[0, 0] == (min(cands)[1] if cands else -1), name
        scene.release()


def test_empty_scene_and_tiny_frames(ctx):
    cam = api.srt_camera(); cam.forward[:] = (0, 0, -1); cam.right[:] = (1, 0, 0); cam.up[:] = (0, 1, 0)
    cam.half_w = cam.half_h = 0.5
    scene = ctx.create_scene(np.zeros((0, 9), np.float32))
    for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
        p = api.make_params(5, 3, background=(0.25, 0.5, 0.75), shade=api.SRT_SHADE_SHADOW, accel=accel)
        rgb, hid, ht, st = scene.render(cam, api.srt_light(), p)
        assert (hid == -1).all() and np.isinf(ht).all() and np.allclose(rgb, [0.25, 0.5, 0.75])
        assert st.primary_rays == 15 and st.shadow_rays == 0
    scene.release()
    # ragged sizes: widths/heights that are not multiples of any tile
    v, a, cam, light = api.gen_cornell_box(37, 19)
    scene = ctx.create_scene(v, a)
    for accel in (api.SRT_ACCEL_BRUTE, api.SRT_ACCEL_BVH):
        p = api.make_params(37, 19, shade=api.SRT_SHADE_SHADOW, accel=accel, spp=3, jitter=1)
        assert_parity(scene.render(cam, light, p), O.render(v, a, cam, light, p), f"ragged accel{accel}")
    scene.release()


def test_render_errors_are_reported(ctx):
    v, a, cam, light = api.gen_cornell_box(16, 16)
    scene = ctx.create_scene(v, a, build_bvh=False)
    p = api.make_params(16, 16, accel=api.SRT_ACCEL_BVH)
    with pytest.raises(api.MLError, match="without a BVH"):
        scene.render(cam, light, p)
    p = api.make_params(16, 16, y0=8, y1=4)
    with pytest.raises(api.MLError, match="Bad row band"):
        scene.render(cam, light, p)
    p = api.make_params(16, 16, spp=0)
    with pytest.raises(api.MLError, match="sample count"):
        scene.render(cam, light, p)
    scene.release()
    bad = np.full((1, 9), np.nan, np.float32)
    with pytest.raises(api.MLError, match="Non-finite"):
        ctx.create_scene(bad)


def test_render_to_image_stays_on_device_and_maps_back(ctx):
    W, H = 96, 64
    v, a, cam, light = api.gen_cornell_box(W, H)
    scene = ctx.create_scene(v, a)
    img = ctx.create_image(W, H, 3)
    p = api.make_params(W, H, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH)
    st = scene.render_to_image(cam, light, p, img)
    assert st.primary_rays == W * H
    ref = O.render(v, a, cam, light, p)
    np.testing.assert_allclose(img.read(), ref[0], rtol=REL, atol=1e-7)
    wrong = ctx.create_image(W, H, 4)
    with pytest.raises(api.MLError, match="does not match"):
        scene.render_to_image(cam, light, p, wrong)
    for o in (wrong, img, scene):
        o.release()


def test_render_device_into_torch_tensor(ctx):
    """srtRenderDevice with caller-owned device memory and stream (torch is plumbing)."""
    import torch
    W, H = 128, 72
    v, a, cam, light = api.gen_mesh(6, 4000, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=2, jitter=1, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH)
    rgb = torch.zeros(H, W, 3, device="cuda"); hid = torch.zeros(H, W, dtype=torch.int32, device="cuda")
    ht = torch.zeros(H, W, device="cuda")
    stream = torch.cuda.current_stream()
    scene.render_device(cam, light, p, rgb.data_ptr(), hid.data_ptr(), ht.data_ptr(), stream.cuda_stream)
    st = scene.stats()
    host = scene.render(cam, light, p)
    assert np.array_equal(rgb.cpu().numpy(), host[0]) and np.array_equal(hid.cpu().numpy(), host[1])
    assert st.primary_rays == W * H * 2 and st.kernel_ms > 0
    scene.release()


def _soup(rng, n, extent):
    """Triangle soup: mostly small triangles, some long slivers, a few huge ones that cross any camera plane."""
    c = rng.uniform(-extent, extent, (n, 1, 3))
    size = np.where(rng.random((n, 1, 1)) < 0.9, 0.04, 0.6) * extent
    v = c + rng.normal(0, 1, (n, 3, 3)) * size
    sl = rng.random(n) < 0.05
    v[sl, 2] = v[sl, 1] + rng.normal(0, 1e-3 * extent, (int(sl.sum()), 3))        # slivers
    v[:4] = rng.uniform(-3 * extent, 3 * extent, (4, 3, 3))                         # huge
    return v.reshape(n, 9).astype(np.float32), rng.random((n, 3)).astype(np.float32)


def _camera(rng, origin, look_at, fov_w, fov_h, shear=0.0):
    cam = api.srt_camera()
    f = np.asarray(look_at, np.float64) - origin
    f /= np.linalg.norm(f)
    r = np.cross(f, [0.0, 1.0, 0.0] if abs(f[1]) < 0.9 else [1.0, 0.0, 0.0]); r /= np.linalg.norm(r)
    u = np.cross(r, f)
    if shear:      # a basis that is neither orthogonal nor unit length is still a valid camera
        r = 1.3 * r + shear * u
        u = 0.8 * u + 0.5 * shear * f
    cam.origin[:] = [float(x) for x in origin]; cam.forward[:] = [float(x) for x in f]
    cam.right[:] = [float(x) for x in r]; cam.up[:] = [float(x) for x in u]
    cam.half_w, cam.half_h = fov_w, fov_h
    return cam


@pytest.mark.parametrize("case", range(8))
def test_screen_and_light_grids_equal_exhaustive_search(ctx, case):
    """spp >= 32 renders primary rays through the per-frame screen grid and shadow rays through the light
    grid (docs/RENDER_SPEC.md section 8).  Both only select candidates, so the frame must equal the
    exhaustive kernel's bit for bit -- for cameras inside the geometry, triangles behind / across the camera
    plane, slivers, huge triangles, sheared camera bases, and lights above, beside and inside the scene."""
    rng = np.random.default_rng(100 + case)
    W, H = 72, 40
    v, a = _soup(rng, 600, 10.0)
    origin = rng.uniform(-8, 8, 3) if case % 2 else rng.uniform(-1, 1, 3) * 25     # inside / outside the soup
    cam = _camera(rng, origin, rng.uniform(-5, 5, 3), 0.3 + 0.6 * rng.random(), 0.2 + 0.4 * rng.random(),
                  shear=0.3 if case in (2, 5) else 0.0)
    light = api.srt_light()
    lp = [(0, 60, 5), (45, 30, -38), (2, 1, -3), (-70, -55, 64)][case % 4]          # (2,1,-3): inside -> BVH shadows
    light.position[:] = [float(x) for x in lp]; light.intensity[:] = [900.0, 800.0, 700.0]
    scene = ctx.create_scene(v, a)
    # 32+ spp: sample-parallel kernel; 1..3 spp: thread-per-pixel kernel (same grids); 9 spp: light grid only
    for spp in (32 + case, 1 + case % 3, 9):
        common = dict(spp=spp, jitter=1, seed=case, shade=api.SRT_SHADE_SHADOW, background=(0.1, 0.2, 0.3))
        brute = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BRUTE, **common))
        grid = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, **common))
        for k, name in enumerate(("rgb", "hit_id", "hit_t")):
            assert np.array_equal(brute[k], grid[k]), f"case {case} spp {spp}: {name} differs from the exhaustive kernel"
        assert brute[3].shadow_rays == grid[3].shadow_rays
    # a band that starts off the 8-row tile grid falls back to the BVH for primary rays: same pixels
    band = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, y0=13, y1=37, buf_y0=13, **common))
    assert np.array_equal(band[0][:24], brute[0][13:37]) and np.array_equal(band[1][:24], brute[1][13:37])
    scene.release()


def _layer(z, n, half, rng, jiggle):
    """An n x n grid of quads (2 triangles each) in the plane z = const (slightly jiggled), spanning [-half, half]^2."""
    xs = np.linspace(-half, half, n + 1)
    tris = []
    for i in range(n):
        for j in range(n):
            c = np.array([[xs[i], xs[j], z], [xs[i + 1], xs[j], z], [xs[i + 1], xs[j + 1], z], [xs[i], xs[j + 1], z]])
            c[:, 2] += rng.normal(0, jiggle, 4)
            tris += [c[[0, 1, 2]], c[[0, 2, 3]]]
    return np.asarray(tris)


@pytest.mark.parametrize("case", range(4))
def test_nearest_first_candidate_walk_equals_exhaustive_search(ctx, case):
    """The sample-parallel kernel walks a pixel's candidates nearest first and stops at the first one that lies wholly
    behind the hit in hand by more than 1/64 (docs/RENDER_SPEC.md section 8).  Layers stacked in depth exercise every side of
    that rule: coincident layers (equal t: the smaller id must win, whichever is staged first), layers 1e-4, 1 % and 2 %
    apart (inside and just outside the margin), far layers (skipped), ids assigned back to front, a sheared camera basis
    (t = depth * |D| must hold for it) and a camera whose rays graze the layers.  Bit-identical to the exhaustive kernel."""
    rng = np.random.default_rng(900 + case)
    W, H = 96, 64
    depths = [10.0, 10.0, 10.001, 10.1, 10.2, 12.0, 20.0, 20.0 * (1 + 1 / 64.0), 40.0]
    layers = [_layer(z, 14 + k, 9.0 + k, rng, 0.0 if k < 2 else [0.0, 2e-4, 0.02, 0.3][case]) for k, z in enumerate(depths)]
    # holes: drop a random third of every layer's triangles so deeper layers show through
    layers = [l[rng.random(len(l)) > 0.33] for l in layers]
    order = list(range(len(layers)))
    if case % 2:
        order.reverse()                                  # far layers get the small ids
    v = np.concatenate([layers[k] for k in order]).reshape(-1, 9).astype(np.float32)
    a = rng.random((len(v), 3)).astype(np.float32)
    origin = np.array([0.3, -0.2, 0.0]) if case < 2 else np.array([14.0, 3.0, 2.0])       # head-on / grazing
    cam = _camera(rng, origin, (0.0, 0.0, 15.0), 0.55, 0.4, shear=0.25 if case in (1, 3) else 0.0)
    light = api.srt_light()
    light.position[:] = [3.0, 8.0, -6.0]; light.intensity[:] = [4000.0, 3500.0, 3000.0]
    scene = ctx.create_scene(v, a)
    for spp in (32, 64, 3):
        common = dict(spp=spp, jitter=1, seed=case, shade=api.SRT_SHADE_SHADOW, background=(0.1, 0.2, 0.3))
        brute = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BRUTE, **common))
        grid = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, **common))
        for k, name in enumerate(("rgb", "hit_id", "hit_t")):
            assert np.array_equal(brute[k], grid[k]), f"case {case} spp {spp}: {name} differs from the exhaustive kernel"
    ids = np.unique(brute[1])
    assert (brute[1] == -1).any() or len(ids) > 50        # a real frame: many triangles visible
    # several layers are visible through the holes (the walk must not stop early) ...
    first = np.cumsum([0] + [len(layers[k]) for k in order])
    seen = {int(np.searchsorted(first, i, side="right") - 1) for i in ids if i >= 0}
    assert len(seen) >= 4, seen
    scene.release()


def test_grids_on_cornell_box(ctx):
    """A dozen wall-sized triangles around the camera: every triangle is 'global' or covers hundreds of
    tiles; the grid path must still reproduce the oracle."""
    W, H = 96, 96
    v, a, cam, light = api.gen_cornell_box(W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=33, jitter=1, seed=4, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH)
    got = scene.render(cam, light, p)
    ref = O.render(v, a, cam, light, p)
    assert_parity(got, ref, "cornell through the grids")
    scene.release()


@pytest.mark.parametrize("view", range(10))
def test_grids_equal_bvh_only_on_mesh_views(ctx, view, monkeypatch):
    """The 100K-triangle benchmark mesh from ten other viewpoints (high, low and grazing cameras, one inside the
    terrain's bounding box looking along it; lights overhead, low on the horizon and inside the box): the grid paths
    (SRT_GRID / SRT_LIGHT_GRID default on) against the same kernels walking the BVH only.  Bit-identical."""
    rng = np.random.default_rng(500 + view)
    W, H = 320, 184
    v, a, cam0, light0 = api.gen_mesh(1, 100_000, W, H)
    if not hasattr(test_grids_equal_bvh_only_on_mesh_views, "scene"):
        test_grids_equal_bvh_only_on_mesh_views.scene = ctx.create_scene(v, a)      # one BVH build for the ten views
    scene = test_grids_equal_bvh_only_on_mesh_views.scene
    height = [40.0, 12.0, 3.0, 80.0, 6.0, 25.0, 1.5, 60.0, 9.0, 18.0][view]
    origin = np.array([rng.uniform(-60, 60), height, rng.uniform(-60, 60)])
    target = np.array([rng.uniform(-30, 30), rng.uniform(-2, 6), rng.uniform(-30, 30)])
    cam = _camera(rng, origin, target, 0.2 + 0.7 * rng.random(), 0.15 + 0.4 * rng.random())
    light = api.srt_light()
    lp = [(30, 60, 25), (-80, 8, 10), (0, 5, 0), (5, 200, -5), (120, 30, -90)][view % 5]
    light.position[:] = [float(x) for x in lp]; light.intensity[:] = [4000.0, 3800.0, 3500.0]
    for spp in (32, 1):
        p = api.make_params(W, H, spp=spp, jitter=1, seed=view, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                            background=(0.5, 0.7, 0.9), collect_stats=1)
        monkeypatch.setenv("SRT_GRID", "1"); monkeypatch.setenv("SRT_LIGHT_GRID", "1")
        grid = scene.render(cam, light, p)
        monkeypatch.setenv("SRT_GRID", "0"); monkeypatch.setenv("SRT_LIGHT_GRID", "0")
        bvh = scene.render(cam, light, p)
        for k, name in enumerate(("rgb", "hit_id", "hit_t")):
            assert np.array_equal(grid[k], bvh[k]), f"view {view} spp {spp}: {name} differs from the BVH-only render"
        assert grid[3].shadow_rays == bvh[3].shadow_rays
        # the grids only ever remove node visits; printed with -s for the curious
        assert grid[3].box_tests <= bvh[3].box_tests
        print(f"view {view} spp {spp}: box tests {grid[3].box_tests} with grids, {bvh[3].box_tests} without; "
              f"hit fraction {(bvh[1] >= 0).mean():.2f}")
