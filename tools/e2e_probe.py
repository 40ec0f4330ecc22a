#!/usr/bin/env python
"""Wall time of srtRender (host buffers) against the device-resident render of the benchmark frame."""
import sys, time, ctypes as C
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from simpleraytracer_b200 import api
W,H=1920,1080
v,a,cam,light=api.gen_mesh(1,100000,W,H)
ctx=api.Context(); scene=ctx.create_scene(v,a)
p=api.make_params(W,H,spp=64,jitter=1,seed=1,shade=api.SRT_SHADE_SHADOW,accel=api.SRT_ACCEL_BVH,background=(0.5,0.7,0.9))
host=np.zeros((H,W,3),np.float32)
for it in range(4):
    st=api.srt_render_stats()
    t0=time.perf_counter()
    rc=scene.lib.srtRender(scene.handle,C.byref(cam),C.byref(light),C.byref(p),host.ctypes.data,None,None,C.byref(st))
    t1=time.perf_counter()
    print(f"wall {1e3*(t1-t0):.2f} ms kernel_ms {st.kernel_ms:.2f}")
import torch
fr=torch.zeros(H,W,3,device='cuda'); torch.cuda.synchronize()
for it in range(3):
    t0=time.perf_counter(); scene.render_device(cam,light,p,fr.data_ptr(),0,0,0); s=scene.stats(); t1=time.perf_counter()
    print(f"device wall {1e3*(t1-t0):.2f} kernel {s.kernel_ms:.2f}")
