"""Parity at BASELINE.json's FULL sizes (configs 3, 4 and 5), through the C ABI.

The CPU oracle's BVH path finishes these in seconds on the GPU box's host cores (bench.py's
cpu_baseline times the same calls), so the whole frame is compared, not a crop:
  * config 3: 1920x1080, 64 spp, Lambert + shadow ray, 100K triangles (the headline workload),
  * config 4: 3840x2160, 1 spp, 1M triangles,
  * config 5: the 6-layer 3x3 conv net on a 1920x1080 image, fp32 and bf16.
Beside the oracle, size-independent properties that need no oracle at all: exhaustive search ==
BVH traversal on every pixel of a full frame, bit-identical re-runs, and band composition.
Bar (north_star): hit ids bit-exact, hit distance and radiance <= 1e-5 relative, fp32 inference
<= 1e-3 of the output range, bf16 <= 3e-2."""
import numpy as np
import pytest

import oracle_lib as O
from simpleraytracer_b200 import api, graphdef_writer as gw
from test_render_gpu import assert_parity

pytestmark = pytest.mark.gpu

MESH_SEED, SAMPLE_SEED = 1, 1          # bench.py's WORKLOAD
SKY = (0.5, 0.7, 0.9)


def test_config3_headline_frame_vs_oracle(ctx):
    """Every pixel of the benchmarked frame (sample-parallel BVH kernel) against the oracle."""
    W, H, SPP = 1920, 1080, 64
    v, a, cam, light = api.gen_mesh(MESH_SEED, 100_000, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=SPP, jitter=1, seed=SAMPLE_SEED, shade=api.SRT_SHADE_SHADOW,
                        accel=api.SRT_ACCEL_BVH, background=SKY)
    got = scene.render(cam, light, p)
    ref = O.render(v, a, cam, light, p, threads=0, use_bvh=True)
    assert_parity(got, ref, "config 3, full 1080p x 64 spp")
    assert got[3].primary_rays == W * H * SPP == int(ref[3][0])
    assert got[3].shadow_rays == int(ref[3][1])
    # bit-identical re-run (ordered accumulation, no atomics on the frame)
    again = scene.render(cam, light, p)
    assert np.array_equal(again[0], got[0]) and np.array_equal(again[1], got[1]) and np.array_equal(again[2], got[2])
    scene.release()


def test_config3_exhaustive_search_equals_bvh_on_the_full_frame(ctx):
    """No oracle involved: the brute-force kernel (every ray x every triangle) and the BVH kernels
    must agree on the closest hit of all 2M pixels, ids AND distances bit for bit."""
    W, H = 1920, 1080
    v, a, cam, light = api.gen_mesh(MESH_SEED, 100_000, W, H)
    scene = ctx.create_scene(v, a)
    common = dict(spp=1, jitter=1, seed=SAMPLE_SEED, shade=api.SRT_SHADE_SHADOW, background=SKY)
    brute = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BRUTE, **common))
    bvh = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, **common))
    assert np.array_equal(brute[1], bvh[1]), f"{np.count_nonzero(brute[1] != bvh[1])} hit ids differ"
    assert np.array_equal(brute[2], bvh[2])
    assert np.array_equal(brute[0], bvh[0])
    assert brute[3].shadow_rays == bvh[3].shadow_rays
    # band composition at full size: two half frames == the frame
    top = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, y0=0, y1=536, **common))
    bot = scene.render(cam, light, api.make_params(W, H, accel=api.SRT_ACCEL_BVH, y0=536, y1=H, **common))
    assert np.array_equal(top[0][:536], bvh[0][:536]) and np.array_equal(bot[0][536:], bvh[0][536:])
    assert top[3].primary_rays + bot[3].primary_rays == W * H
    scene.release()


def test_config4_4k_1m_triangles_vs_oracle(ctx):
    W, H = 3840, 2160
    v, a, cam, light = api.gen_mesh(MESH_SEED, 1_000_000, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=1, jitter=1, seed=SAMPLE_SEED, shade=api.SRT_SHADE_SHADOW,
                        accel=api.SRT_ACCEL_BVH, background=SKY)
    got = scene.render(cam, light, p)
    ref = O.render(v, a, cam, light, p, threads=0, use_bvh=True)
    assert_parity(got, ref, "config 4, 4K x 1M triangles")
    assert got[3].primary_rays == W * H and got[3].shadow_rays == int(ref[3][1])
    scene.release()


def test_config5_conv_net_1080p_vs_oracle(ctx, tmp_path):
    """The benchmarked network (3x3 convs 3->64, 4 x 64->64, 64->3, ReLU, residual) on one full
    1080p image: fp32 direct-conv kernels and the bf16 tcgen05 kernels against the fp32 oracle."""
    W, H = 1920, 1080
    path = tmp_path / "net.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=64, depth=6, seed=0)
    model = ctx.create_model(path)
    i, _ = model.info()
    i.width, i.height = W, H
    model.set_input_info(i)
    x = np.random.default_rng(3).random((H, W, 3), dtype=np.float32)
    inp = ctx.create_image(W, H, 3); out = ctx.create_image(W, H, 3)
    inp.write(x)
    ref = O.run_denoiser(x, layers, alpha=0.1)
    scale = float(np.abs(ref).max())
    def elem_rel(got):
        """relative error over the elements that carry signal (|ref| > 5 % of the range): north_star's '1e-3 relative'"""
        m = np.abs(ref) > 5e-2 * scale
        return float((np.abs(got - ref)[m] / np.abs(ref)[m]).max())
    model.set_precision(api.MLX_PRECISION_FP32)             # the CUDA-core fp32 kernels (the default is bf16x3)
    assert model.tensor_core_layers() == 0
    model.infer(inp, out)
    got32 = out.read()
    err32 = float(np.abs(got32 - ref).max()) / scale
    assert err32 <= 1e-3, f"fp32 path: {err32:.3e} of the output range"
    assert elem_rel(got32) <= 1e-3, f"fp32 path: element-relative {elem_rel(got32):.3e}"
    model.set_precision(api.MLX_PRECISION_BF16)
    assert model.tensor_core_layers() == 6
    model.infer(inp, out)
    err16 = float(np.abs(out.read() - ref).max()) / scale
    assert err16 <= 3e-2, f"bf16 path: {err16:.3e} of the output range"
    model.set_precision(api.MLX_PRECISION_BF16X3)           # split operands: fp32-grade on the tensor cores
    assert model.tensor_core_layers() == 6
    model.infer(inp, out)
    got_x3 = out.read()
    err_x3 = float(np.abs(got_x3 - ref).max()) / scale
    assert err_x3 <= 1e-4, f"bf16x3 path: {err_x3:.3e} of the output range"
    assert elem_rel(got_x3) <= 1e-3, f"bf16x3 path (the default precision): element-relative {elem_rel(got_x3):.3e}"
    for obj in (inp, out, model):
        obj.release()


def test_unet_1080p_vs_oracle(ctx, tmp_path):
    """A two-level U-Net (MaxPool down, Keras-style nearest upsampling, skip concats, FusedBatchNorm, global
    residual) on one full 1080p image, default precision (bf16x3 tensor-core convolutions) and fp32, vs the oracle."""
    W, H = 1920, 1080
    path = tmp_path / "unet.pb"
    _, prog = gw.make_unet(path, channels=3, width=16, levels=2, seed=2, up="nearest", down="maxpool")
    model = ctx.create_model(path)
    i, _ = model.info()
    i.width, i.height = W, H
    model.set_input_info(i)
    x = np.random.default_rng(5).random((H, W, 3), dtype=np.float32)
    inp = ctx.create_image(W, H, 3); out = ctx.create_image(W, H, 3)
    inp.write(x)
    ref = O.run_program(x, prog)
    scale = float(np.abs(ref).max())
    assert model.tensor_core_layers() >= 3      # encoder + bottleneck convs; 48- / 96-channel decoder convs stay on the CUDA cores
    model.infer(inp, out)
    err = float(np.abs(out.read() - ref).max()) / scale
    assert err <= 1e-4, f"U-Net, default precision: {err:.3e} of the output range"
    model.set_precision(api.MLX_PRECISION_FP32)
    model.infer(inp, out)
    err = float(np.abs(out.read() - ref).max()) / scale
    assert err <= 1e-3, f"U-Net, fp32: {err:.3e} of the output range"
    for obj in (inp, out, model):
        obj.release()
