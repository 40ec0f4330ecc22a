"""Triangle / box tests per ray of the config-3 frame with the grids on and off (instrumented kernel)."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from simpleraytracer_b200 import api
W, H = 1920, 1080
v, a, cam, light = api.gen_mesh(1, 100000, W, H)
ctx = api.Context()
scene = ctx.create_scene(v, a)
frame = ctx.create_image(W, H, 3)
p = api.make_params(W, H, spp=64, jitter=1, seed=7, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                    background=(0.5, 0.7, 0.9))
p.collect_stats = 1
st = scene.render_to_image(cam, light, p, frame)
rays = st.primary_rays + st.shadow_rays
print(f"primary {st.primary_rays} shadow {st.shadow_rays} box/ray {st.box_tests / rays:.2f} tri/ray {st.tri_tests / rays:.2f} "
      f"kernel_ms {st.kernel_ms:.2f}")
