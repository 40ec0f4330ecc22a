"""GPU tests of the multi-GPU frame: interleaved row blocks stored straight into one
frame buffer, completion published by the render kernel's last block (device-side flags), consumed by a spin
kernel / srtGatherFrameHost on the frame owner.  The assembled frame must equal the single-GPU frame bit for
bit.  Every mechanism is first exercised with several replicas on ONE GPU (the flags, tickets, chunking and
host gather are identical), then on real peers when the box has them."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from simpleraytracer_b200 import api

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def gpu_count():
    import torch
    return torch.cuda.device_count()


def device_to_host(ptr, arr):
    """cudaMemcpy through the CUDA runtime torch has loaded (same primary context as the library's)."""
    import ctypes as C
    import torch
    torch.cuda.init()
    rt = C.CDLL("libcudart.so.12")
    rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    rt.cudaMemcpy.restype = C.c_int
    assert rt.cudaMemcpy(arr.ctypes.data, ptr, arr.nbytes, 2) == 0       # cudaMemcpyDeviceToHost


COMMON = dict(jitter=1, seed=3, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH, background=(0.5, 0.7, 0.9))


@pytest.mark.parametrize("shape", [
    dict(W=150, H=77, spp=2, n=3, tris=5000),       # ragged: rows not a multiple of 8 * N, thread-per-pixel kernel
    dict(W=96, H=200, spp=32, n=2, tris=20000),     # sample-parallel kernel + screen grid
    dict(W=64, H=40, spp=1, n=4, tris=300),         # fewer row blocks (5) than chunks x ranks: empty launches signal too
])
def test_scene_group_replicas_on_one_gpu_equal_single_render(ctx, shape):
    W, H, n = shape["W"], shape["H"], shape["n"]
    v, a, cam, light = api.gen_mesh(7, shape["tris"], W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=shape["spp"], **COMMON)
    want, _, _, st1 = scene.render(cam, light, p)
    group = api.SceneGroup([0] * n, v, a)
    assert group.size() == n
    for chunks in ("1", "3"):
        os.environ["SRT_GROUP_CHUNKS"] = chunks     # read at every call
        got, st = group.render(cam, light, p)
        assert np.array_equal(got, want)
        assert st.primary_rays == st1.primary_rays and st.shadow_rays == st1.shadow_rays
        assert st.kernel_ms > 0
    # frame left on GPU 0 (no host copy): read it back with a plain cudaMemcpy
    _, st = group.render(cam, light, p, want_host=False)
    ptr = group.frame_pointer()
    assert ptr
    dev = np.empty((H, W, 3), np.float32)
    device_to_host(ptr, dev)
    assert np.array_equal(dev, want)
    # page-locked destination: chunks copied straight from the device
    import torch
    pinned = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory()
    got, _ = group.render(cam, light, p, rgb_host=pinned.numpy())
    assert np.array_equal(pinned.numpy(), want)
    os.environ.pop("SRT_GROUP_CHUNKS", None)
    group.release()
    scene.release()


def test_signal_flags_and_gather_between_two_contexts(ctx):
    """The one-process-per-GPU protocol of bench.py, with both 'ranks' on one GPU: every rank renders its row
    blocks into rank 0's image with srtRenderDeviceSignal, rank 0 gathers with srtGatherFrameHost."""
    W, H, N = 128, 120, 2
    v, a, cam, light = api.gen_mesh(5, 8000, W, H)
    ranks = [api.Context(device=0) for _ in range(N)]
    scenes = [c.create_scene(v, a) for c in ranks]
    want = scenes[0].render(cam, light, api.make_params(W, H, spp=4, **COMMON))[0]
    frame = ranks[0].create_image(W, H, 3)
    d_frame = frame.device_pointer(for_write=True)
    d_flags = ranks[0].device_alloc(4 * 16 * N)
    for epoch, chunks in ((1, 1), (2, 2), (3, 4)):
        whole = api.make_params(W, H, spp=4, row_block_stride=N, **COMMON)
        rows, n_chunks = api.chunk_layout(whole, chunks)
        assert rows % (8 * N) == 0 and rows * n_chunks >= H
        host = np.full((H, W, 3), -1.0, np.float32)
        for r in reversed(range(N)):
            p = api.make_params(W, H, spp=4, row_block_stride=N, row_block_phase=r, **COMMON)
            scenes[r].render_device_signal(cam, light, p, d_frame, d_flags + 4 * r, epoch, chunk_stride=N,
                                           n_chunks=n_chunks)
        # Both 'ranks' share ONE GPU here: the producers are drained before the consumer's spin kernel is
        # launched (kernels that wait on one another must not be separate launches on one GPU); the genuinely
        # concurrent wait is exercised on real peers (test_scene_group_on_real_peers, the bench test).
        for sc in scenes:
            sc.stats()
        ranks[0].gather_frame_host(whole, d_frame, d_flags, N, n_chunks, epoch, host)
        assert np.array_equal(host, want), f"epoch {epoch}"
    # device-side wait only (the device-timed bench step): frame complete once the wait kernel has run
    for r in range(N):
        p = api.make_params(W, H, spp=4, row_block_stride=N, row_block_phase=r, **COMMON)
        scenes[r].render_device_signal(cam, light, p, d_frame, d_flags + 4 * r, 9, chunk_stride=N)
    scenes[1].stats()                            # drained first: one GPU (see above)
    ranks[0].stream_wait_flags(d_flags, N, 9)
    scenes[0].stats()                            # synchronises rank 0's stream
    assert ranks[0].flag_wait_timeouts() == 0
    assert np.array_equal(frame.read(), want)
    # st.grid_build_ms is reported for signalled device renders
    st = scenes[0].stats()
    assert 0 < st.grid_build_ms < st.kernel_ms
    assert st.light_grid_build_ms > 0
    ranks[0].device_free(d_flags)
    frame.release()
    for s in scenes:
        s.release()
    for c in ranks:
        c.release()


def test_a_missing_rank_times_out_instead_of_hanging(ctx):
    W, H = 64, 32
    v, a, cam, light = api.gen_mesh(1, 500, W, H)
    scene = ctx.create_scene(v, a)
    frame = ctx.create_image(W, H, 3)
    d_flags = ctx.device_alloc(64)
    whole = api.make_params(W, H, row_block_stride=2, **COMMON)
    p0 = api.make_params(W, H, row_block_stride=2, row_block_phase=0, **COMMON)
    scene.render_device_signal(cam, light, p0, frame.device_pointer(for_write=True), d_flags, 1, chunk_stride=2)
    host = np.zeros((H, W, 3), np.float32)
    before = ctx.flag_wait_timeouts()
    with pytest.raises(api.MLError, match="completion flag"):      # rank 1 never renders
        ctx.gather_frame_host(whole, frame.device_pointer(), d_flags, 2, 1, 1, host, timeout_ms=50)
    assert ctx.flag_wait_timeouts() == before + 1
    ctx.device_free(d_flags)
    frame.release()
    scene.release()


def test_srt_render_cli_multi_gpu_equals_single(tmp_path):
    """`srt_render -devices ...` (the C++ client of srtRenderGroup) writes the same bytes as one GPU."""
    cli = ROOT / "simpleraytracer_b200" / "bin" / "srt_render"
    if not cli.exists():
        pytest.skip("srt_render not built")
    n = gpu_count()
    devices = ",".join(str(d) for d in range(n)) if n >= 2 else "0,0,0"
    args = ["-s", "mesh", "-n", "20000", "-w", "320", "-h", "200", "-spp", "8"]
    one, many = tmp_path / "one.bin", tmp_path / "many.bin"
    subprocess.run([str(cli), *args, "-o", str(one)], check=True, capture_output=True)
    r = subprocess.run([str(cli), *args, "-devices", devices, "-o", str(many)], check=True, capture_output=True, text=True)
    assert f"{max(n, 3) if n < 2 else n} GPU(s)" in r.stderr
    assert one.read_bytes() == many.read_bytes() and one.stat().st_size == 320 * 200 * 12


def test_scene_group_on_real_peers(ctx):
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    W, H = 640, 360
    v, a, cam, light = api.gen_mesh(2, 50000, W, H)
    scene = ctx.create_scene(v, a)
    group = api.SceneGroup(list(range(n)), v, a)
    for spp in (1, 32):
        p = api.make_params(W, H, spp=spp, **COMMON)
        want = scene.render(cam, light, p)[0]
        for _ in range(3):                   # repeated frames: epochs advance, flags are never reset
            got, st = group.render(cam, light, p)
            assert np.array_equal(got, want)
    group.release()
    scene.release()


def test_bench_multi_process_frame_check():
    """bench.py under torchrun (one process per GPU, frame and flags IPC-mapped from rank 0): the frame the N
    ranks assemble equals the frame rank 0 renders alone -- for the peer-store path and for the reduce fallback."""
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    for extra in ([], ["--exchange", "reduce_sum"]):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
               "--master-addr", "127.0.0.1", "--master-port", "29611", str(ROOT / "bench.py"), "--gpus", "2",
               "--steps", "2", "--warmup", "3", "--spp", "8", "--no-cpu-baseline", "--no-brute", "--no-inference",
               "--no-extra-configs", *extra]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
        assert r.returncode == 0, r.stderr[-2000:]
        line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
        assert line["n_gpus"] == 2 and line["frame_check"]["equal"] is True
        assert line["config"]["exchange"] == ("reduce_sum" if extra else "peer_store_nvlink+device_flags")
