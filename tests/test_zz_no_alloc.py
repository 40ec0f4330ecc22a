"""GPU test of srtRenderDeviceNoAlloc (include/srt_render.h): the render whose per-frame scratch lives in a static
device array of the library, for callers that cannot call cudaMalloc.  No counterpart in the reference.
Called through the C ABI and compared with the existing entry point
(srtRenderDevice) on the same scene, two calls in a row, within the tolerance the render tests use (hit ids
exact, distances and radiance 1e-5 relative -- in practice bit-identical).

This file sorts last on purpose (`pytest --collect-only`): a failure here
must not hide the tests before it.  It passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from simpleraytracer_b200 import api

pytestmark = pytest.mark.gpu
REL = 1e-5
COMMON = dict(jitter=1, seed=5, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH, background=(0.5, 0.7, 0.9))


def frames(torch, W, H):
    return (torch.full((H, W, 3), -3.0, device="cuda"), torch.full((H, W), -7, dtype=torch.int32, device="cuda"),
            torch.full((H, W), -3.0, device="cuda"))


def render_pair(torch, scene, cam, light, p):
    """(existing entry point, no-alloc entry point) -> two (rgb, hit_id, hit_t) triples as numpy arrays."""
    W, H = p.width, p.height
    s = torch.cuda.current_stream().cuda_stream
    want = frames(torch, W, H)
    scene.render_device(cam, light, p, want[0].data_ptr(), want[1].data_ptr(), want[2].data_ptr(), s)
    scene.stats()
    torch.cuda.synchronize()
    got = frames(torch, W, H)
    scene.render_device_no_alloc(cam, light, p, got[0].data_ptr(), got[1].data_ptr(), got[2].data_ptr(), s)
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in want], [t.cpu().numpy() for t in got]


def assert_same(got, want, what):
    assert np.array_equal(got[1], want[1]), f"{what}: {np.count_nonzero(got[1] != want[1])} hit ids differ"
    fin = np.isfinite(want[2])
    assert np.array_equal(np.isfinite(got[2]), fin), what
    np.testing.assert_allclose(got[2][fin], want[2][fin], rtol=REL, atol=0, err_msg=what)
    np.testing.assert_allclose(got[0], want[0], rtol=REL, atol=1e-7, err_msg=what)


def test_no_alloc_render_equals_render_device_on_two_calls_in_a_row(ctx):
    import torch
    v, a, cam, light = api.gen_mesh(11, 20000, 640, 360)
    scene = ctx.create_scene(v, a)
    # call 1: 640x360, thread-per-pixel kernel + screen grid in the static scratch
    p1 = api.make_params(640, 360, spp=2, **COMMON)
    want, got = render_pair(torch, scene, cam, light, p1)
    assert_same(got, want, "call 1 (640x360, 2 spp)")
    assert (got[1] >= 0).any() and (got[1] == -1).any()          # geometry and sky: a real frame
    # call 2, straight after: a SMALLER frame of another scene, sample-parallel kernel -- the array still holds
    # call 1's lists and counts, the call must clear what it needs
    v2, a2, cam2, light2 = api.gen_mesh(12, 3000, 200, 120)
    scene2 = ctx.create_scene(v2, a2)
    p2 = api.make_params(200, 120, spp=32, **COMMON)
    want2, got2 = render_pair(torch, scene2, cam2, light2, p2)
    assert_same(got2, want2, "call 2 (200x120, 32 spp)")
    # and the first scene again, other modes: exhaustive search (no grid), 8 spp kernel (light grid off, BVH shadows)
    for kw in (dict(spp=1, accel=api.SRT_ACCEL_BRUTE), dict(spp=8), dict(spp=1, shade=api.SRT_SHADE_ALBEDO)):
        p = api.make_params(640, 360, **{**COMMON, **kw})
        w, g = render_pair(torch, scene, cam, light, p)
        assert_same(g, w, str(kw))
    # no device memory is taken by further calls (the first ones loaded the kernels)
    free0 = torch.cuda.mem_get_info()[0]
    out = frames(torch, 640, 360)
    free1 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        scene.render_device_no_alloc(cam, light, p1, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free1 <= free0
    assert np.array_equal(out[0].cpu().numpy(), want[0])
    scene2.release()
    scene.release()


def test_no_alloc_render_rejects_sizes_beyond_the_static_scratch(ctx):
    import torch
    out = torch.zeros(8, 8, 3, device="cuda")
    n = api.SRT_NOALLOC_MAX_TRIANGLES + 1
    v, a, cam, light = api.gen_mesh(1, n, 64, 64)
    big = ctx.create_scene(v, a)
    with pytest.raises(api.MLError, match=f"{n} triangles.*at most {api.SRT_NOALLOC_MAX_TRIANGLES}"):
        big.render_device_no_alloc(cam, light, api.make_params(8, 8, **COMMON), out.data_ptr())
    big.release()
    v, a, cam, light = api.gen_mesh(1, 500, 64, 64)
    scene = ctx.create_scene(v, a)
    # 1032 x 1024 = 129 x 128 tiles = 16512 > 16384
    with pytest.raises(api.MLError, match=f"16512 8x8 tiles.*at most {api.SRT_NOALLOC_MAX_TILES}"):
        scene.render_device_no_alloc(cam, light, api.make_params(1032, 1024, **COMMON), out.data_ptr())
    # the largest frame that fits renders, and equals the existing entry point
    p = api.make_params(1024, 1024, spp=1, **COMMON)
    want, got = render_pair(torch, scene, cam, light, p)
    assert_same(got, want, "1024x1024")
    # the usual argument checks still apply
    with pytest.raises(api.MLError, match="sample count"):
        scene.render_device_no_alloc(cam, light, api.make_params(8, 8, spp=0, **COMMON), out.data_ptr())
    scene.release()
