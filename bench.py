#!/usr/bin/env python
"""bench.py -- headline benchmark of the render hot path (BASELINE.json metric:
"Mrays/s (closest-hit+shade, 1080p)").

Workload (BASELINE.json configs[2]): 1920x1080, 64 spp, Lambert + one shadow ray per lit hit, seeded synthetic
100,000-triangle mesh, candidate grids in front of the BVH.  One "step" = one full frame (132,710,400 primary
rays + the shadow rays they spawn).  At N > 1 the frame is sharded in interleaved 8-row blocks over the ranks
(scene replicated); every rank's render kernel stores its pixels straight into rank 0's frame over NVLink and
its last block publishes a completion flag there; rank 0 waits for the flags on the device.  No collective is
on the data path (torch.distributed / NCCL is plumbing: handles, barriers, max-over-ranks of the timings).
Total work is fixed: scaling = "strong".

The same JSON line carries three more objects: `config2` (Cornell box 512x512, 1 spp: launch-bound, us per
frame), `config4` (3840x2160, 1 spp, 1M triangles at the run's N, with an NCCL row-band gather timed beside the
peer-store exchange) and `inference` (BASELINE configs[4]: 1080p image-to-image conv net, batch 8).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

`--impl reference` times the CPU implementation of the same path on the host cores.  The reference repository
contains no renderer, so that arm is the build-authored CPU oracle ("port"), on a bounded
sample of the same workload (the first spp_sample of the 64 samples of every pixel); it loads nothing of the
product (scene arrays come from oracle/libsrtscenes.so).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    "config3": dict(name="render_1080p_64spp_shadow_100k_tris", width=1920, height=1080, spp=64,
                    triangles=100_000, mesh_seed=1, sample_seed=1),
    "config4": dict(name="render_4k_1spp_shadow_1m_tris", width=3840, height=2160, spp=1,
                    triangles=1_000_000, mesh_seed=1, sample_seed=1),
}
WORKLOAD = dict(WORKLOADS["config3"])
CPU_SAMPLE_SPP = 64           # CPU sample: all 64 spp of every pixel (~10 s on 16 host cores)
FLOP_PER_TRI_TEST = 45        # SURVEY.md section 8d conventions
FLOP_PER_BOX_TEST = 21
BACKGROUND = (0.5, 0.7, 0.9)
# Flag chunks of a multi-GPU host frame.  1: one work-list render (fastest render, rows finish in no order), the copy
# pipelined in pieces behind its single flag; > 1: the render itself in that many chunks (natural order), each copied
# as soon as its flag arrives.  Measured at N = 8 (profiles/r02_e2e_chunks.txt).
# Which wins depends on what bounds the frame (N = 2: config 3, render-bound, 63.1 Grays/s with 4 chunks against 41.9 with
# 1; config 4, copy-bound, 2.67 against 1.25): 4 when the frame has >= 8 spp, else 1.  BENCH_E2E_CHUNKS overrides.
E2E_CHUNKS_ENV = int(os.environ.get("BENCH_E2E_CHUNKS", "0"))


def e2e_chunks(spp):
    return E2E_CHUNKS_ENV if E2E_CHUNKS_ENV > 0 else (4 if spp >= 8 else 1)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp", type=int, default=None, help=argparse.SUPPRESS)  # debugging only
    ap.add_argument("--no-cpu-baseline", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-brute", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-inference", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-extra-configs", action="store_true", help=argparse.SUPPRESS)   # skip config2 / config4
    # headline workload override (not a bench line of the contract): config4 = 4K, 1 spp, 1M triangles
    ap.add_argument("--workload", default="config3", choices=["config3", "config4"], help=argparse.SUPPRESS)
    # N > 1 exchange: auto = peer stores + device flags, reduce_sum only if the peer mapping cannot be opened
    ap.add_argument("--exchange", default="auto", choices=["auto", "reduce_sum"], help=argparse.SUPPRESS)
    return ap.parse_args()


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """One long-running `nvidia-smi -lms 50` whose lines are collected while the GPU legs run (warm-up, timed
    region, e2e, extra configs): clocks, power, utilisation and throttle reasons every 50 ms."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,utilization.gpu,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device_index):
        self.index = device_index
        self.samples = []
        self.marks = {}
        self.proc = None
        self._thread = None

    def _run(self):
        for line in self.proc.stdout:
            parts = [x.strip() for x in line.strip().split(",")]
            if len(parts) >= 9:
                try:
                    float(parts[1]); float(parts[2])
                except ValueError:
                    continue
                self.samples.append((time.perf_counter(), parts))

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        except OSError:
            self.proc = None
        return self

    def mark(self, name):
        """Names the interval that starts now (summary reports samples per interval)."""
        self.marks[name] = time.perf_counter()

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=3)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=3)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "samples_under_load": 0}

        def num(s, default=0.0):
            try:
                return float(s)
            except ValueError:
                return default
        load = [p for _, p in self.samples if num(p[4]) >= 50.0]
        use = load if load else [p for _, p in self.samples]
        sm = sorted(num(p[1]) for p in use)
        reasons = [n for k, n in enumerate(self.REASONS)
                   if any(p[5 + k].lower().startswith("active") for _, p in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz_under_load": sm[0], "sm_max_mhz": num(self.samples[0][1][2]),
                "power_w_max": max(num(p[3]) for _, p in self.samples), "reasons": reasons,
                "samples": len(self.samples), "samples_under_load": len(load), "period_ms": 50,
                "span": "warm-up + timed steps + e2e + extra configs + inference (one sampler for the whole GPU arm)"}


def physical_gpu_index(local_rank: int):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return vis.split(",")[local_rank].strip()      # index or UUID: nvidia-smi -i takes both
        except IndexError:
            pass
    return local_rank


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return json.loads(p.read_text()), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def ncu_fields(key: str) -> dict:
    """Per-launch fields that only an ncu capture can give (DRAM traffic, issue-slot utilisation ...), read
    from profiles/ncu_fields.json (written by tools/ncu_fields.py from a committed ncu CSV).  Every entry
    records the sha256 of the kernel's source file at capture time; if the source has changed since, the
    fields are reported as null with the reason instead of silently describing an older kernel."""
    path = ROOT / "profiles" / "ncu_fields.json"
    if not path.exists():
        return {"traffic": None, "traffic_source": "profiles/ncu_fields.json absent"}
    entry = json.loads(path.read_text()).get(key)
    if entry is None:
        return {"traffic": None, "traffic_source": f"no ncu capture recorded for {key}"}
    src = ROOT / entry["source_file"]
    sha = hashlib.sha256(src.read_bytes()).hexdigest() if src.exists() else None
    if sha != entry["source_sha256"]:
        return {"traffic": None, "traffic_source": f"{entry['source_file']} changed since the ncu capture "
                                                   f"{entry.get('csv')} (sha mismatch): re-profile"}
    out = {k: v for k, v in entry.items() if k not in ("source_file", "source_sha256")}
    out["traffic_source"] = f"ncu --set full, dram__bytes_read+write per launch ({entry.get('csv')}); source sha matches"
    return out


# --------------------------------------------------------------------------- CPU arm (oracle: test infrastructure)
REFERENCE_ARM_BUDGET_S = 150.0   # CPU seconds the timed steps of `--impl reference` may take in total


def oracle():
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib as O   # bench.py's CPU-baseline legs are allowed to execute oracle/
    return O


def oracle_params(O, wl, spp):
    p = O.orc_params()
    p.width, p.height, p.x0, p.x1, p.y0, p.y1 = wl["width"], wl["height"], 0, wl["width"], 0, wl["height"]
    p.spp, p.jitter, p.seed, p.shade = spp, 1, wl["sample_seed"], 2          # 2 = Lambert + shadow ray
    p.t_min, p.shadow_eps = 1e-4, 1e-3
    p.background[:] = BACKGROUND
    return p


def cpu_sample(v, a, cam, light, wl, spp):
    """One bounded sample of the workload on the host cores through the oracle's BVH path: the first `spp`
    samples of every pixel.  The oracle's own BVH build is timed apart (the GPU arm reports bvh_build_ms
    outside the frame time as well).  Returns (Mrays/s, trace seconds, rays, threads, bvh build seconds)."""
    O = oracle()
    p = oracle_params(O, wl, spp)
    t0 = time.perf_counter()
    _, _, _, counts, threads = O.render(v, a, cam, light, p, threads=0, use_bvh=True)
    dt = time.perf_counter() - t0
    setup = O.last_setup_ms() * 1e-3
    trace = max(dt - setup, 1e-9)
    rays = int(counts[0]) + int(counts[1])
    return rays / trace / 1e6, trace, rays, threads, setup


def run_reference_arm(args):
    """`--impl reference`: the CPU implementation of the path, all host threads.  Loads oracle/ only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O = oracle()
    wl = WORKLOAD
    spp = args.spp or wl["spp"]
    v, a, cam, light = O.gen_mesh(wl["mesh_seed"], wl["triangles"], wl["width"], wl["height"])
    # One untimed calibration pass (2 spp) sizes the per-step sample so that the whole run stays within
    # ~150 s of CPU time whatever --steps is: the first n of the frame's spp samples of every pixel.
    _, cal_s, _, _, _ = cpu_sample(v, a, cam, light, wl, min(2, spp))
    per_spp = max(cal_s / min(2, spp), 1e-3)
    sample_spp = int(max(1, min(CPU_SAMPLE_SPP, spp, REFERENCE_ARM_BUDGET_S / (max(args.steps, 1) * per_spp))))
    total_rays, total_s, threads, setup_s = 0, 0.0, 0, 0.0
    for _ in range(args.steps):
        _, dt, rays, threads, setup = cpu_sample(v, a, cam, light, wl, sample_spp)
        total_rays += rays
        total_s += dt
        setup_s += setup
    value = total_rays / total_s / 1e6
    sample = (f"first {sample_spp} of {spp} spp of every pixel of the {wl['width']}x{wl['height']} frame per step "
              f"({total_rays // args.steps} rays), oracle BVH path; its BVH build ({setup_s / args.steps * 1e3:.0f} ms "
              f"per step) is outside the timed trace, like bvh_build_ms of the GPU arm")
    line = {
        "impl": "reference", "metric": "Mrays/s (closest-hit+shade, 1080p)", "value": value, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "width": wl["width"], "height": wl["height"], "spp": spp,
                   "triangles": wl["triangles"], "note": "reference repo has no renderer; CPU oracle port",
                   "bvh_build_ms": setup_s / args.steps * 1e3},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- GPU arm: plumbing
class Env:
    """Process-group plumbing of the GPU arm: one process per GPU (torchrun), NCCL for handles / barriers /
    reductions of timings.  Nothing here is on the data path of the peer-store exchange."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            # NCCL prints its "NCCL version ..." banner (VERSION and WARN levels) and its log to stdout, and honours
            # NCCL_DEBUG_FILE only above the VERSION level: raise VERSION to WARN and send everything to stderr, so
            # that stdout carries the one JSON line
            if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            dist.init_process_group("nccl", device_id=self.dev)
        # A dedicated torch stream is made current for everything below: the render kernels are launched on it
        # through the C ABI (a NULL stream would mean "the context's own stream"), torch.distributed's
        # collectives order against it, and the timing events are recorded on it.
        self.stream = torch.cuda.Stream(device=self.dev)
        torch.cuda.set_stream(self.stream)
        assert self.stream.cuda_stream != 0
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)   # > 126 MB L2
        self._token = torch.zeros(1, dtype=torch.float32, device=self.dev)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def device_barrier(self):
        """Aligns the ranks ON THE DEVICE: a one-element all-reduce enqueued on the launching stream (the host does not
        wait).  Work enqueued behind it starts on every GPU when the collective completes there, within a microsecond
        or two of the others, whatever the jitter with which the host threads left the host barrier."""
        if self.world > 1:
            self.dist.all_reduce(self._token)

    def reduce(self, values, op="max"):
        """MAX / SUM / MIN over the ranks of a list of floats (float64)."""
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            ops = {"max": self.dist.ReduceOp.MAX, "sum": self.dist.ReduceOp.SUM, "min": self.dist.ReduceOp.MIN}
            self.dist.all_reduce(t, op=ops[op])
        return [float(x) for x in t.tolist()]

    def gather(self, value):
        """Every rank's float, as a list indexed by rank."""
        t = self.torch.zeros(self.world, dtype=self.torch.float64, device=self.dev)
        t[self.rank] = value
        if self.world > 1:
            self.dist.all_reduce(t)
        return [float(x) for x in t.tolist()]


class FrameRig:
    """One workload resident on this rank's GPU and the N-rank frame it renders into.

    N = 1: the frame is a local device buffer.  N > 1: interleaved 8-row blocks (block b belongs to rank
    b % N); rank 0 owns the frame (an ml_image) and a flag array, both IPC-mapped by every other rank; a step
    is srtRenderDeviceSignal on every rank (pixels stored straight into rank 0's frame, completion flag
    published by the render kernel's last block) plus, on rank 0, srtStreamWaitFlags.  If the mapping cannot
    be opened (or --exchange reduce_sum) the ranks sum-reduce zero-padded local frames over NCCL instead
    (bit-exact: x + 0 == x)."""

    def __init__(self, env, ctx, wl, spp, exchange_pref="auto", vertices=None):
        from simpleraytracer_b200 import api, sharding
        self.api, self.sharding, self.env, self.ctx, self.wl = api, sharding, env, ctx, wl
        torch, dist = env.torch, env.dist
        W, H = wl["width"], wl["height"]
        self.W, self.H, self.spp = W, H, spp
        if vertices is None:
            self.v, self.a, self.cam, self.light = api.gen_mesh(wl["mesh_seed"], wl["triangles"], W, H)
        else:
            self.v, self.a, self.cam, self.light = vertices
        self.scene = ctx.create_scene(self.v, self.a, build_bvh=True)
        self.info = self.scene.info()
        self.shade = wl.get("shade", api.SRT_SHADE_SHADOW)
        self.jitter = wl.get("jitter", 1)
        self.params = self.make_params(stride=env.world if env.world > 1 else 0, phase=env.rank if env.world > 1 else 0)
        self.my_rows = len(sharding.interleaved_rows(H, env.world, env.rank))
        self.frame = torch.zeros(H, W, 3, device=env.dev)       # local frame (N = 1, the fallback, solo renders)
        self.target_ptr, self.exchange = self.frame.data_ptr(), "none"
        self.frame_img, self.d_flags, self.epoch, self._opened = None, 0, 0, []
        self.flag_words = 64 * max(env.world, 1)                 # [16 chunks][N ranks] + the "consumed" word
        if env.world > 1:
            handles = [None, None]
            if env.rank == 0:
                self.frame_img = ctx.create_image(W, H, 3)
                self.d_flags = ctx.device_alloc(4 * self.flag_words)
                handles = [self.frame_img.ipc_export(), ctx.ipc_export_pointer(self.d_flags)]
            dist.broadcast_object_list(handles, src=0)
            ok = 1
            try:
                if env.rank == 0:
                    self.target_ptr = self.frame_img.device_pointer(for_write=True)
                else:
                    self.target_ptr = ctx.ipc_open(handles[0])
                    self._opened.append(self.target_ptr)
                    self.d_flags = ctx.ipc_open(handles[1])
                    self._opened.append(self.d_flags)
            except api.MLError as e:
                print(f"[rank {env.rank}] peer mapping unavailable ({e}); falling back to reduce", file=sys.stderr)
                ok = 0
            ok = int(env.reduce([ok], "min")[0])
            self.exchange = "peer_store_nvlink+device_flags" if ok and exchange_pref == "auto" else "reduce_sum"
            if self.exchange == "reduce_sum":
                self.target_ptr = self.frame.data_ptr()
        self.d_consumed = self.d_flags + 4 * (self.flag_words - 1) if self.d_flags else 0

    def make_params(self, stride=0, phase=0, collect_stats=0):
        api = self.api
        return api.make_params(self.W, self.H, spp=self.spp, jitter=self.jitter, seed=self.wl["sample_seed"],
                               shade=self.shade, accel=api.SRT_ACCEL_BVH, background=BACKGROUND,
                               row_block_stride=stride, row_block_phase=phase, collect_stats=collect_stats)

    # ---- one frame, device resident (enqueue only; the caller brackets it with events / synchronises)
    def step(self, params=None):
        env, p = self.env, params or self.params
        s = env.stream.cuda_stream
        if self.exchange == "peer_store_nvlink+device_flags":
            self.epoch += 1
            self.scene.render_device_signal(self.cam, self.light, p, self.target_ptr, self.d_flags + 4 * env.rank,
                                            self.epoch, chunk_stride=env.world, n_chunks=1, stream=s)
            if env.rank == 0:      # device-side wait: every rank's rows have landed in this GPU's memory
                self.ctx.stream_wait_flags(self.d_flags, env.world, self.epoch, stream=s)
        elif self.exchange == "reduce_sum":
            if env.rank == 0:
                self.frame.zero_()                 # the previous step's sum must not be summed again
            self.scene.render_device(self.cam, self.light, p, self.target_ptr, 0, 0, s)
            self.sharding.reduce_interleaved(self.frame, env.rank)
        else:
            self.scene.render_device(self.cam, self.light, p, self.target_ptr, 0, 0, s)

    def assembled_frame(self):
        """Rank 0: host copy of the frame the last step assembled."""
        if self.exchange == "peer_store_nvlink+device_flags":
            self.frame_img.device_pointer(for_write=True)      # peers wrote the device copy behind the image's back
            return self.frame_img.read()
        return self.frame.cpu().numpy()

    def solo_frame(self):
        """Rank 0 renders the same frame ALONE (single-GPU path) into its local buffer."""
        self.scene.render_device(self.cam, self.light, self.make_params(), self.frame.data_ptr(), 0, 0,
                                 self.env.stream.cuda_stream)
        self.scene.stats()
        self.env.torch.cuda.synchronize()
        return self.frame.cpu().numpy()

    def frame_check(self):
        """Untimed: one N-rank frame against rank 0's solo render of the same frame, bit for bit."""
        env = self.env
        env.barrier()
        self.step()
        env.barrier()
        res = None
        if env.rank == 0:
            got = self.assembled_frame()
            want = self.solo_frame()
            bad = int(np.count_nonzero((got.view(np.uint32) != want.view(np.uint32)).any(axis=2)))
            res = {"equal": bad == 0, "pixels": self.W * self.H, "mismatched_pixels": bad,
                   "compared": f"frame assembled by {env.world} ranks ({self.exchange}) vs rank 0 alone, uint32 view"}
            if self.exchange == "peer_store_nvlink+device_flags":
                res["flag_wait_timeouts"] = self.ctx.flag_wait_timeouts()
        env.barrier()
        return res

    # ---- one frame into a HOST buffer on rank 0 (the e2e deliverable at N > 1)
    def e2e_step(self, host_rgb):
        env = self.env
        s = env.stream.cuda_stream
        if self.exchange == "peer_store_nvlink+device_flags":
            self.epoch += 1
            whole = self.make_params(stride=env.world)
            # the previous frame must have left rank 0's buffer before anyone overwrites it: device-side
            # wait on rank 0's "consumed" word (peers poll it over NVLink), no host barrier in the loop
            self.ctx.stream_wait_flags(self.d_consumed, 1, self.epoch - 1, stream=s)
            self.scene.render_device_signal(self.cam, self.light, self.params, self.target_ptr,
                                            self.d_flags + 4 * env.rank, self.epoch, chunk_stride=env.world,
                                            n_chunks=e2e_chunks(self.spp), stream=s)
            if env.rank == 0:
                self.ctx.gather_frame_host(whole, self.target_ptr, self.d_flags, env.world, e2e_chunks(self.spp), self.epoch,
                                           host_rgb)
                self.ctx.stream_write_flag(self.d_consumed, self.epoch, stream=s)
            env.stream.synchronize()
        elif self.exchange == "reduce_sum":
            self.step()
            if env.rank == 0:
                host_rgb[...] = self.frame.cpu().numpy()
            else:
                env.stream.synchronize()
        else:
            import ctypes as C
            api = self.api
            st = api.srt_render_stats()
            rc = self.scene.lib.srtRender(self.scene.handle, C.byref(self.cam), C.byref(self.light),
                                          C.byref(self.params), host_rgb.ctypes.data, None, None, C.byref(st))
            if rc != api.ML_OK:
                raise RuntimeError(self.scene.error())

    def sync_consumed(self):
        """Before a run of e2e steps: bring the 'consumed' word up to the current epoch."""
        if self.exchange == "peer_store_nvlink+device_flags":
            self.env.barrier()
            if self.env.rank == 0:
                self.ctx.stream_write_flag(self.d_consumed, self.epoch, stream=self.env.stream.cuda_stream)
            self.env.barrier()

    def release(self):
        self.env.barrier()
        for ptr in self._opened:
            self.ctx.ipc_close(ptr)
        self.env.barrier()
        if self.frame_img is not None:
            self.frame_img.release()
            self.ctx.device_free(self.d_flags)
        self.scene.release()


# --------------------------------------------------------------------------- GPU arm: measurements
def median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2] if xs else None


def timed_device_steps(env, rig, steps, warmup):
    """K frames, inputs resident: every step is bracketed by CUDA events on the launching stream.  Before it,
    untimed: a host barrier, the L2 flush, and a device-side barrier on the launching stream, so that the ranks'
    GPUs start the step together (the hosts leave a host barrier tens of microseconds apart, which at 8 ranks
    was charged to a 0.59 ms step) with the step's launches already queued behind the flush, as they are at N = 1.
    On rank 0 the bracket ends behind the device-side flag wait, i.e. when the whole frame is in its memory.
    Returns per-rank lists (step_ms, own_span_ms, grid_build_ms) and the instrumented counts."""
    torch = env.torch
    for _ in range(max(warmup, 3)):
        rig.step()
    env.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    span, grid = [], []
    for k in range(steps):
        env.barrier()
        env.flush.zero_()
        env.device_barrier()
        ev[k][0].record(env.stream)
        rig.step()
        ev[k][1].record(env.stream)
        torch.cuda.synchronize()
        st = rig.scene.stats()
        span.append(st.kernel_ms)
        grid.append(st.grid_build_ms)
    env.barrier()
    return [s.elapsed_time(e) for s, e in ev], span, grid


def timed_e2e_steps(env, rig, steps, host_rgb):
    """K frames into rank 0's HOST buffer through the public calls; wall clock, max over ranks."""
    rig.sync_consumed()
    rig.e2e_step(host_rgb)                       # warm-up (staging buffers, page faults of the destination)
    rig.e2e_step(host_rgb)
    rig.sync_consumed()
    t0 = time.perf_counter()
    for _ in range(steps):
        rig.e2e_step(host_rgb)
    env.barrier()
    return env.reduce([time.perf_counter() - t0], "max")[0]


def instrumented_counts(env, rig):
    """One untimed pass of the counting kernel variants: rays and box / triangle tests of the frame."""
    p = rig.make_params(stride=rig.params.row_block_stride, phase=rig.params.row_block_phase, collect_stats=1)
    env.barrier()
    rig.step(p)
    env.torch.cuda.synchronize()
    st = rig.scene.stats()
    mine = [st.primary_rays, st.shadow_rays, st.box_tests, st.tri_tests]
    total = env.reduce(mine, "sum")
    env.barrier()
    return total, mine


def config2_leg(env, ctx):
    """BASELINE configs[1]: primary-ray closest hit of the 12-triangle Cornell box, 512x512, 1 spp, one GPU
    (rank 0; the other ranks idle).  262,144 rays: the frame is launch-bound, so the numbers are us per frame
    and Mrays/s, not a roofline fraction."""
    from simpleraytracer_b200 import api
    torch = env.torch
    res = None
    if env.rank == 0:
        W = H = 512
        v, a, cam, light = api.gen_cornell_box(W, H)
        scene = ctx.create_scene(v, a)
        frame = torch.zeros(H, W, 3, device=env.dev)
        host = np.zeros((H, W, 3), np.float32)
        res = {"workload": "cornell_box_512x512_1spp", "triangles": int(len(v)), "rays_per_frame": W * H,
               "bound": "launch latency (one 8-block-row launch sequence per frame; no roofline fraction)"}
        import ctypes as C
        for name, shade, accel in (("closest_hit", api.SRT_SHADE_ALBEDO, api.SRT_ACCEL_BVH),
                                   ("closest_hit_brute", api.SRT_SHADE_ALBEDO, api.SRT_ACCEL_BRUTE),
                                   ("shadow", api.SRT_SHADE_SHADOW, api.SRT_ACCEL_BVH)):
            p = api.make_params(W, H, spp=1, shade=shade, accel=accel, background=BACKGROUND)
            for _ in range(5):
                scene.render_device(cam, light, p, frame.data_ptr(), 0, 0, env.stream.cuda_stream)
            torch.cuda.synchronize()
            ms = []
            for _ in range(30):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(env.stream)
                scene.render_device(cam, light, p, frame.data_ptr(), 0, 0, env.stream.cuda_stream)
                e1.record(env.stream)
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
            st = scene.stats()
            rays = st.primary_rays + st.shadow_rays
            wall = []
            stt = api.srt_render_stats()
            for _ in range(12):
                t0 = time.perf_counter()
                rc = scene.lib.srtRender(scene.handle, C.byref(cam), C.byref(light), C.byref(p), host.ctypes.data,
                                         None, None, C.byref(stt))
                wall.append(time.perf_counter() - t0)
                if rc != api.ML_OK:
                    raise RuntimeError(scene.error())
            wall = wall[2:]
            res[name] = {"us_per_frame_median": median(ms) * 1e3, "us_per_frame_best": min(ms) * 1e3,
                         "mrays_per_s": rays / (median(ms) * 1e-3) / 1e6, "rays": int(rays),
                         "e2e_us_per_frame_median": median(wall) * 1e6,
                         "e2e_mrays_per_s": rays / median(wall) / 1e6, "d2h_bytes_per_frame": W * H * 12}
        scene.release()
    env.barrier()
    return res


def nccl_gather_leg(env, rig, steps):
    """The exchange BASELINE configs[3] names, timed beside the peer-store path: every rank renders a
    CONTIGUOUS row band into its own frame and the bands are gathered on rank 0 over NCCL
    (sharding.gather_bands: all_gather_into_tensor for equal bands, send / recv otherwise)."""
    torch = env.torch
    bands = rig.sharding.row_bands(rig.H, env.world)
    y0, y1 = bands[env.rank]
    api = rig.api
    p = api.make_params(rig.W, rig.H, spp=rig.spp, jitter=rig.jitter, seed=rig.wl["sample_seed"], shade=rig.shade,
                        accel=api.SRT_ACCEL_BVH, background=BACKGROUND, y0=y0, y1=y1)

    def step():
        if y1 > y0:
            rig.scene.render_device(rig.cam, rig.light, p, rig.frame.data_ptr(), 0, 0, env.stream.cuda_stream)
        rig.sharding.gather_bands(rig.frame, bands, env.rank)

    for _ in range(3):
        step()
    env.barrier()
    total, render = [], []
    for _ in range(steps):
        env.flush.zero_()
        env.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(env.stream)
        step()
        e1.record(env.stream)
        torch.cuda.synchronize()
        total.append(e0.elapsed_time(e1))
        render.append(rig.scene.stats().kernel_ms if y1 > y0 else 0.0)
    env.barrier()
    check = None
    if env.rank == 0:
        got = rig.frame.cpu().numpy()
        want = rig.solo_frame()
        check = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
    env.barrier()
    tot = env.reduce([sum(total)], "max")[0] / steps
    ren = env.reduce([sum(render)], "max")[0] / steps
    return {"render_plus_gather_ms": tot, "render_only_ms_max_rank": ren, "frame_equal": check,
            "sharding": "contiguous row bands (what an in-place all-gather needs), NCCL over NVLink",
            "bytes_gathered": rig.W * rig.H * 12}


# --------------------------------------------------------------------------- inference leg (config 5)
NET = dict(name="denoiser_conv3x3_w64_d6_residual", width=64, depth=6, batch=8)
INFER_REPS = 10


def cpu_inference_sample(layers, threads=0):
    """Oracle (fp32 direct conv, all host threads) on a 768x768 crop; scaled to 1080p images/s by pixels."""
    O = oracle()
    x = np.random.default_rng(3).random((768, 768, 3), dtype=np.float32)
    t0 = time.perf_counter()
    O.run_denoiser(x, layers, alpha=0.1, threads=threads)
    dt = time.perf_counter() - t0
    return (768 * 768) / (WORKLOAD["width"] * WORKLOAD["height"]) / dt, dt


def inference_leg(env, ctx, rig, no_cpu):
    """BASELINE config 5: the rendered 1080p frame feeds the image-to-image conv net, batch 8, bf16 on
    the tensor cores.  images/s device-resident and end to end from host buffers (mlxInferBatch), each from
    INFER_REPS batches timed on the device (CUDA events inside the library, mlxModelLastBatchMs: first upload
    enqueued -> last read-back done) -- median and best, max over ranks; the dominant kernel against the
    tensor roofline; render -> infer with no host bounce."""
    import tempfile
    from simpleraytracer_b200 import api, graphdef_writer as gw
    W, H = rig.W, rig.H
    rank, world = env.rank, env.world
    with tempfile.TemporaryDirectory() as d:
        path = Path(d) / "net.pb"
        _, layers = gw.make_denoiser(path, channels=3, width=NET["width"], depth=NET["depth"], seed=0)
        model = ctx.create_model(path)
    i, o = model.info()
    i.width, i.height = W, H
    model.set_input_info(i)
    n_img = max(NET["batch"] // world, 1)
    rng = np.random.default_rng(10 + rank)
    ins = [ctx.create_image(W, H, 3) for _ in range(n_img)]
    outs = [ctx.create_image(W, H, 3) for _ in range(n_img)]
    x = rng.random((H, W, 3), dtype=np.float32)
    for a in ins:
        a.write(x)                    # host copy authoritative: every infer uploads it (reference semantics)
    res = {}

    def timed(fn, reps):
        """reps batches; per batch the device span (library events) and the host wall time; max over ranks."""
        fn()
        dev_ms, wall_ms = [], []
        for _ in range(reps):
            env.barrier()
            t0 = time.perf_counter()
            fn()
            wall_ms.append((time.perf_counter() - t0) * 1e3)
            dev_ms.append(model.last_batch_ms())
        dev_ms = [env.reduce([t], "max")[0] for t in dev_ms]
        wall_ms = [env.reduce([t], "max")[0] for t in wall_ms]
        return {"device_ms_median": median(dev_ms), "device_ms_best": min(dev_ms),
                "wall_ms_median": median(wall_ms), "wall_ms_best": min(wall_ms), "reps": reps}

    def rate(t, key):
        return n_img * world / (t[key] * 1e-3)

    for prec, key in ((api.MLX_PRECISION_FP32, "fp32"), (api.MLX_PRECISION_BF16X3, "bf16x3"),
                      (api.MLX_PRECISION_BF16, "bf16")):
        model.set_precision(prec)
        reps = INFER_REPS if key != "fp32" else 5           # fp32 CUDA-core path: 0.11 s per batch of 8
        if key == "bf16x3":
            res["bf16x3_tc_layers"] = model.tensor_core_layers()
        # e2e: host buffers in, host buffers out (outputs' host pointers were handed out by read())
        for b in outs:
            b.read()
        t = timed(lambda: model.infer_batch(ins, outs), reps)
        res[key + "_e2e"] = {"images_per_s": rate(t, "wall_ms_median"), "images_per_s_best": rate(t, "wall_ms_best"),
                             "device_images_per_s": rate(t, "device_ms_median"), **t}
        # device resident: inputs and outputs stay in HBM
        for a in ins:
            a.device_pointer(False); a.device_pointer(True)
        fresh = [ctx.create_image(W, H, 3) for _ in range(n_img)]   # host pointer never handed out: no D2H
        t = timed(lambda: model.infer_batch(ins, fresh), reps)
        res[key] = {"images_per_s": rate(t, "device_ms_median"), "images_per_s_best": rate(t, "device_ms_best"),
                    "wall_images_per_s": rate(t, "wall_ms_median"), **t}
        if rank == 0:
            res[key + "_ops"] = model.profile(ins[0], fresh[0])
        for f in fresh:
            f.release()
        for a in ins:
            a.write(x)
    res["tc_layers"] = model.tensor_core_layers()
    res["launches_per_image"] = model.launch_count()

    # render -> infer on the device (full frame on every rank: replicas, no exchange)
    frame = ctx.create_image(W, H, 3)
    den = ctx.create_image(W, H, 3)
    p1 = rig.make_params()
    ms = []
    for k in range(6):
        env.barrier()
        t0 = time.perf_counter()
        rig.scene.render_to_image(rig.cam, rig.light, p1, frame)
        model.infer(frame, den)
        ms.append((time.perf_counter() - t0) * 1e3)
    res["render_plus_infer_ms"] = env.reduce([median(ms[1:])], "max")[0]
    cpu = None
    if rank == 0 and not no_cpu:
        v, dt = cpu_inference_sample(layers)
        cpu = {"value": v, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
               "sample": f"fp32 conv oracle on one 768x768 crop ({dt:.1f} s), scaled to 1920x1080 by pixel count"}
    for obj in ins + outs + [frame, den, model]:
        obj.release()
    if rank != 0:
        return None
    peaks, src = measured_peaks()
    ops = res.pop("bf16_ops")
    fp32_ops = res.pop("fp32_ops")
    x3_ops = res.pop("bf16x3_ops")
    # the dominant kernel: the launch with the most algorithmic work (the 64->64 layers), the slowest of those
    k = max(range(len(ops)), key=lambda j: (ops[j][1], ops[j][0]))
    ms, fl, by = ops[k]
    roof = {"bound": "tensor", "kernel": f"conv3x3_tc_rows_kernel<64,1> (op {k}: 3x3 conv 64->64 + bias + relu, bf16 in/out)",
            "achieved": fl / (ms * 1e-3) / 1e12, "peak": peaks.get("bf16_tflops"), "unit": "TFLOP/s",
            "frac": fl / (ms * 1e-3) / 1e12 / peaks["bf16_tflops"] if peaks.get("bf16_tflops") else None,
            "peak_source": f"MEASURED_PEAKS.json bf16_tflops (burst: kernel timed alone), {src}",
            "kernel_ms": ms, "algorithmic_flops_per_launch": fl, "algorithmic_bytes_per_launch": by,
            "hbm_gbs_achieved": by / (ms * 1e-3) / 1e9, "hbm_frac": by / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
            **ncu_fields("conv3x3_tc_rows_kernel"),
            "all_ops_ms": [round(o[0], 4) for o in ops], "sum_ops_ms": sum(o[0] for o in ops),
            "fp32_all_ops_ms": [round(o[0], 4) for o in fp32_ops]}
    return {"metric": "images/s (1080p image-to-image conv net, batch 8)", "unit": "images/s",
            "config": {"workload": "infer_1080p_" + NET["name"], "batch": NET["batch"], "images_per_rank": n_img,
                       "width": W, "height": H, "net": NET["name"], "weights": "random-init (He), seed 0",
                       "parallelism": f"images_dp{world}" if world > 1 else "single_gpu",
                       "timing": f"{INFER_REPS} batches per figure (fp32: 5), CUDA events inside mlxInferBatch; "
                                 "value = median device span, e2e.value = median host wall time of the call"},
            "dtype": "bf16", "value": res["bf16"]["images_per_s"], "value_best": res["bf16"]["images_per_s_best"],
            "e2e": {"value": res["bf16_e2e"]["images_per_s"], "value_best": res["bf16_e2e"]["images_per_s_best"],
                    "unit": "images/s", "h2d_bytes_per_step": NET["batch"] * W * H * 12,
                    "d2h_bytes_per_step": NET["batch"] * W * H * 12, "timing": res["bf16_e2e"]},
            "timing": res["bf16"],
            "fp32": {"value": res["fp32"]["images_per_s"], "e2e": res["fp32_e2e"]["images_per_s"],
                     "timing": res["fp32"], "e2e_timing": res["fp32_e2e"]},
            # fp32-grade results on the tensor cores: split bf16 operands, three MMAs per product
            # (MLX_PRECISION_BF16X3; bound 1e-4 of the range against the fp32 oracle, tests/test_model_gpu.py)
            "fp32_grade_tensor_core": {"value": res["bf16x3"]["images_per_s"], "e2e": res["bf16x3_e2e"]["images_per_s"],
                                       "value_best": res["bf16x3"]["images_per_s_best"],
                                       "timing": res["bf16x3"], "e2e_timing": res["bf16x3_e2e"],
                                       "tensor_core_layers": res["bf16x3_tc_layers"],
                                       "all_ops_ms": [round(o[0], 4) for o in x3_ops],
                                       "roofline": x3_roofline(x3_ops, peaks, src)},
            "tensor_core_layers": res["tc_layers"], "launches_per_image": res["launches_per_image"],
            "render_plus_infer_ms": res["render_plus_infer_ms"], "roofline": roof, "cpu_baseline": cpu}


def x3_roofline(ops, peaks, src):
    """Dominant kernel of the default-precision (bf16x3) path: the slowest of the launches with the most work
    (a 64->64 layer).  Three bf16 MMAs are issued per fp32-equivalent product, so the tensor-pipe figure counts 3x
    the layer's algorithmic flops; bytes: activations carried as hi + lo bf16 halves = 4 B per element either side."""
    k = max(range(len(ops)), key=lambda j: (ops[j][1], ops[j][0]))
    ms, fl, by = ops[k]
    peak = peaks.get("bf16_tflops")
    issued = 3.0 * fl / (ms * 1e-3) / 1e12
    return {"bound": "tensor", "kernel": f"conv3x3_tc_kernel<64> split mode (op {k}: 3x3 conv 64->64 + bias + relu, hi/lo bf16 in/out)",
            "achieved": issued, "peak": peak, "unit": "TFLOP/s", "frac": issued / peak if peak else None,
            "fp32_equivalent_tflops": fl / (ms * 1e-3) / 1e12,
            "peak_source": f"MEASURED_PEAKS.json bf16_tflops (burst: kernel timed alone), {src}",
            "kernel_ms": ms, "algorithmic_flops_per_launch": fl, "issued_bf16_flops_per_launch": 3.0 * fl,
            "algorithmic_bytes_per_launch": by, "hbm_gbs_achieved": by / (ms * 1e-3) / 1e9,
            **ncu_fields("conv3x3_tc_kernel_split")}


# --------------------------------------------------------------------------- GPU arm: one sharded-frame workload
def measure_frame_workload(env, ctx, wl, spp, args, steps, warmup, with_nccl_gather):
    """Device-resident frames, the N-rank frame check, end-to-end frames into rank 0's host buffer and (N > 1)
    the NCCL row-band gather alternative, for one workload.  Returns (report dict, rig); rank != 0 gets the
    collective parts only."""
    torch = env.torch
    rig = FrameRig(env, ctx, wl, spp, args.exchange)
    W, H, world = rig.W, rig.H, env.world
    (primary, shadow, boxes, tris), mine = instrumented_counts(env, rig)
    rays = primary + shadow

    check = rig.frame_check() if world > 1 else None
    if world > 1 and rig.exchange != "reduce_sum":
        # the fallback exchange, exercised once on the same frame
        keep = (rig.exchange, rig.target_ptr)
        rig.exchange, rig.target_ptr = "reduce_sum", rig.frame.data_ptr()
        fb = rig.frame_check()
        rig.exchange, rig.target_ptr = keep
        if env.rank == 0:
            check["reduce_sum_fallback_equal"] = fb["equal"]

    step_ms, span_ms, grid_ms = timed_device_steps(env, rig, steps, warmup)
    total_ms = env.reduce([sum(step_ms)], "max")[0]
    value = rays * steps / (total_ms * 1e-3) / 1e6
    spans = env.gather(float(np.mean(span_ms)))
    light_grid_ms = rig.scene.stats().light_grid_build_ms

    # ---- e2e: frames into a HOST buffer on rank 0 (N = 1: srtRender; N > 1: chunked gather behind the flags)
    host_rgb = np.zeros((H, W, 3), np.float32)
    e2e_s = timed_e2e_steps(env, rig, steps, host_rgb)
    e2e_equal = None
    if env.rank == 0 and world > 1:
        e2e_equal = bool(np.array_equal(host_rgb.view(np.uint32), rig.solo_frame().view(np.uint32)))
    pinned = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory()
    e2e_pin_s = timed_e2e_steps(env, rig, steps, pinned.numpy())
    del pinned

    nccl = nccl_gather_leg(env, rig, steps) if (with_nccl_gather and world > 1) else None

    rep = None
    if env.rank == 0:
        own = float(np.mean(span_ms))
        rep = {
            "value": value, "ms_per_step": total_ms / steps, "steps": steps,
            "rays_per_step": rays, "primary_rays": primary, "shadow_rays": shadow, "box_tests": boxes, "tri_tests": tris,
            "rank0_share_of_rays": (mine[0] + mine[1]) / rays if rays else 1.0,
            "rank0_device_span_ms": own, "screen_grid_build_ms": float(np.mean(grid_ms)),
            "light_grid_build_ms": light_grid_ms,
            "light_grid_note": "built once per light position (first frame), reused while the light stays put: "
                               "outside the timed steps",
            "bvh_build_ms": rig.info.bvh_build_ms, "scene_upload_ms": rig.info.upload_ms,
            "bvh_nodes": int(rig.info.n_bvh_nodes), "scene_device_bytes": int(rig.info.device_bytes),
            "exchange": rig.exchange, "frame_check": check,
            "e2e": {"value": rays * steps / e2e_s / 1e6, "unit": "Mrays/s", "ms_per_frame": e2e_s / steps * 1e3,
                    "h2d_bytes_per_step": 152, "d2h_bytes_per_step": W * H * 12,
                    "host_buffer": "pageable (a reference-style caller's std::vector / numpy array) on rank 0, "
                                   "the WHOLE frame" + ("" if world == 1 else f", gathered in {e2e_chunks(rig.spp)} flag chunk(s) behind the device flags"),
                    "value_page_locked_buffer": rays * steps / e2e_pin_s / 1e6,
                    "frame_equal_single_gpu": e2e_equal},
        }
        if world > 1:
            rep["render_only_ms"] = max(spans)                  # slowest rank's own grid build + render kernel
            rep["render_plus_exchange_ms"] = total_ms / steps   # until rank 0 has seen every rank's flag
            rep["breakdown_ms"] = {
                "grid_build": float(np.mean(grid_ms)), "render": own - float(np.mean(grid_ms)),
                "signal_wait": max(float(np.mean(step_ms)) - own, 0.0), "skew": max(spans) - min(spans),
                "per_rank_span": spans, "ideal_step": None,
                "note": "rank 0's timeline: grid_build + render = its own span (library events); signal_wait = time "
                        "its stream then spent in the flag wait (slower ranks + flag latency); skew = slowest minus "
                        "fastest rank span"}
            rep["nccl_gather_alternative"] = nccl
    return rep, rig


def brute_roofline(env, ctx, rig, fp32_peak):
    """The closest-hit intersection kernel against the FP32 roofline: every ray x every triangle at 1080p."""
    api = rig.api
    bp = api.make_params(rig.W, rig.H, spp=1, jitter=1, seed=rig.wl["sample_seed"], shade=api.SRT_SHADE_ALBEDO,
                         accel=api.SRT_ACCEL_BRUTE, background=BACKGROUND)
    s = env.stream.cuda_stream
    rig.scene.render_device(rig.cam, rig.light, bp, rig.frame.data_ptr(), 0, 0, s)   # warm-up
    rig.scene.stats()
    times = []
    for _ in range(2):
        env.flush.zero_()
        rig.scene.render_device(rig.cam, rig.light, bp, rig.frame.data_ptr(), 0, 0, s)
        times.append(rig.scene.stats().kernel_ms)
    ms = min(times)
    tests = float(rig.W) * rig.H * rig.wl["triangles"]
    tf = tests * FLOP_PER_TRI_TEST / (ms * 1e-3) / 1e12
    return {"kernel": "render_brute_kernel<4>", "workload": f"{rig.W}x{rig.H} x 1 spp x {rig.wl['triangles']} triangles, closest hit",
            "ms": ms, "tests": tests, "achieved": tf, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": tf / fp32_peak if fp32_peak else None, "mrays_per_s": rig.W * rig.H / (ms * 1e-3) / 1e6,
            "bound": "fp32"}


def run_b200_arm(args):
    from simpleraytracer_b200 import api
    env = Env()
    rank, world = env.rank, env.world
    ctx = api.Context(device=env.local_rank)
    wl = WORKLOAD
    spp = args.spp or wl["spp"]
    sampler = ClockSampler(physical_gpu_index(env.local_rank)).start()
    t_wall = time.perf_counter()

    main, rig = measure_frame_workload(env, ctx, wl, spp, args, args.steps, args.warmup,
                                       with_nccl_gather=args.workload == "config4")
    wall_s = time.perf_counter() - t_wall

    fp32_peak, brute = None, None
    if rank == 0:
        fp32_peak = ctx.measure_fp32_peak(5)
        if not args.no_brute:
            brute = brute_roofline(env, ctx, rig, fp32_peak)
    env.barrier()

    inference = None
    if not args.no_inference:
        try:
            inference = inference_leg(env, ctx, rig, args.no_cpu_baseline)
        except Exception as e:  # noqa: BLE001  -- the headline render numbers must still be reported
            inference = {"error": f"{type(e).__name__}: {e}"}

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cv, dt, crays, threads, setup = cpu_sample(rig.v, rig.a, rig.cam, rig.light, wl, min(CPU_SAMPLE_SPP, spp))
        cpu = {"value": cv, "unit": "Mrays/s", "cores": threads, "kind": "port",
               "sample": f"first {min(CPU_SAMPLE_SPP, spp)} of {spp} spp of every pixel of the {wl['width']}x{wl['height']} frame "
                         f"({crays} rays, {dt:.1f} s), oracle BVH path; its BVH build ({setup * 1e3:.0f} ms) timed apart"}
    rig.release()

    config2 = config4 = None
    if not args.no_extra_configs:
        try:
            config2 = config2_leg(env, ctx)
        except Exception as e:  # noqa: BLE001
            config2 = {"error": f"{type(e).__name__}: {e}"}
        if args.workload != "config4":
            try:
                w4 = WORKLOADS["config4"]
                rep4, rig4 = measure_frame_workload(env, ctx, w4, w4["spp"], args, min(args.steps, 5), 3,
                                                    with_nccl_gather=True)
                rig4.release()
                if rank == 0:
                    config4 = {"workload": w4["name"], "width": w4["width"], "height": w4["height"], "spp": w4["spp"],
                               "triangles": w4["triangles"], "n_gpus": world, "unit": "Mrays/s",
                               "accel": "screen grid + light grid candidate lists, BVH fallback", **rep4}
            except Exception as e:  # noqa: BLE001
                config4 = {"error": f"{type(e).__name__}: {e}"}
    sampler.stop()

    if rank != 0:
        if world > 1:
            env.dist.destroy_process_group()
        return 0

    # ---- rank 0: report
    flops = main["box_tests"] * FLOP_PER_BOX_TEST + main["tri_tests"] * FLOP_PER_TRI_TEST
    share = main["rank0_share_of_rays"]
    kern_ms = main["rank0_device_span_ms"]
    achieved = flops * share / (kern_ms * 1e-3) / 1e12
    rows0 = rig.my_rows
    kernel_name = "render_bvh_spp_kernel<false,1>" if spp >= 32 else "render_bvh_kernel"
    roofline = {"bound": "fp32", "kernel": kernel_name + " (screen grid + light grid candidate lists, BVH fallback)",
                "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                "frac": achieved / fp32_peak if fp32_peak else None,
                **(ncu_fields(kernel_name) if world == 1 and spp == wl["spp"] else
                   {"traffic": None, "traffic_source": "ncu capture exists for the single-GPU full-spp launch only"}),
                "peak_source": "srtMeasureFp32Peak live FFMA microbenchmark (MEASURED_PEAKS.json has no FP32 figure)",
                "algorithmic_flops_per_launch": flops * share,
                "box_tests": main["box_tests"], "tri_tests": main["tri_tests"], "kernel_ms": kern_ms,
                "algorithmic_bytes_per_launch": rows0 * wl["width"] * 12 + main["scene_device_bytes"],
                "note": "instruction-issue bound, not FP32-pipe bound; flops = box_tests*21 + tri_tests*45 counted by the "
                        "instrumented kernel variant on the same frame -- the grids REMOVE algorithmic work (84 box + 4.6 triangle "
                        "tests per ray with the BVH alone, 0.04 + 3.7 with the grids), so TFLOP/s falls while Mrays/s rises; "
                        "kernel_ms includes the per-frame screen-grid build (4 small launches before the render kernel)"}
    launches_per_step = 6 * world + (1 if world > 1 and main["exchange"] != "reduce_sum" else 0)
    config = {"workload": wl["name"], "width": wl["width"], "height": wl["height"], "spp": spp,
              "triangles": wl["triangles"], "shade": "lambert+shadow", "accel": "bvh+grids",
              "parallelism": f"interleaved_row_blocks{world}" if world > 1 else "single_gpu",
              "exchange": main["exchange"], "l2_flush_between_steps": True,
              "step_start": "host barrier, L2 flush, device-side barrier on the launching stream (untimed), then the start event",
              "rays_per_step": main["rays_per_step"], "primary_rays": main["primary_rays"],
              "shadow_rays": main["shadow_rays"], "bvh_nodes": main["bvh_nodes"],
              "bvh_build_ms": main["bvh_build_ms"], "scene_upload_ms": main["scene_upload_ms"],
              "screen_grid_build_ms": main["screen_grid_build_ms"], "light_grid_build_ms": main["light_grid_build_ms"],
              "light_grid_note": main["light_grid_note"], "wall_s_render_legs": wall_s}
    if world > 1:
        bd = dict(main["breakdown_ms"])
        bd["ideal_step"] = None
        config[f"n{world}_breakdown_ms"] = bd
        config["render_only_ms"] = main["render_only_ms"]
        config["render_plus_exchange_ms"] = main["render_plus_exchange_ms"]
    line = {
        "metric": "Mrays/s (closest-hit+shade, 1080p)" if args.workload == "config3" else "Mrays/s (closest-hit+shade, 4K)",
        "value": main["value"], "unit": "Mrays/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": main["ms_per_step"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config, "frame_check": main["frame_check"],
        "roofline": roofline, "intersect_roofline": brute, "cpu_baseline": cpu,
        "e2e": main["e2e"], "config2": config2, "config4": config4, "inference": inference,
        "gpu_launches": args.steps * launches_per_step, "clocks": sampler.summary(),
        "peaks": {"fp32_tflops_measured_live": fp32_peak, "hbm_gbs": measured_peaks()[0].get("hbm_gbs"),
                  "source": measured_peaks()[1]},
    }
    if args.workload == "config4":
        line["nccl_gather_alternative"] = main.get("nccl_gather_alternative")
    print(json.dumps(line), flush=True)
    ctx.release()
    if world > 1:
        env.dist.destroy_process_group()
    ok = True
    for fc in (main["frame_check"], (config4 or {}).get("frame_check")):
        if fc and (not fc["equal"] or fc.get("reduce_sum_fallback_equal") is False):
            ok = False
    if not ok:
        print("bench.py: the N-rank frame differs from the single-GPU frame", file=sys.stderr)
    return 0 if ok else 1


def main():
    args = parse_args()
    if args.workload == "config4":
        WORKLOAD.clear()
        WORKLOAD.update(WORKLOADS["config4"])
        args.no_brute = True          # 4K x 1M brute force is 3.7e14 FLOP
        args.no_inference = True
    if args.impl == "reference":
        run_reference_arm(args)
        return 0
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
