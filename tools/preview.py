"""Render previews through the C ABI and save them as PNG (gpurun_out/)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracer2_b200 as rt
from PIL import Image
def render(kind, scale, w, h, spp, bounces, name, env=0, tex=512):
    sc = rt.Scene(kind, scale, tex, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights); g.refrence_mode = 1; g.environment = env
    ctx = rt.Renderer(0); rt.Model.from_scene(ctx, sc); lp = rt.LightPasses(ctx, w, h); lp.update_uniform(g)
    lp.execute(1, rt.LaunchParams(spp=spp, max_bounces=bounces))
    d = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)[:, :, :3]
    st = ctx.stats()
    img = np.clip(d / (1 + d), 0, 1) ** (1 / 2.2)
    Image.fromarray((img * 255).astype(np.uint8)).save(f"gpurun_out/{name}.png")
    print(name, "rays", st.rays_traced, "mean", d.mean())
    ctx.close()
render(rt.SCENE_HALL, 1.0, 960, 540, 64, 4, "hall")
render(rt.SCENE_CORNELL, 1.0, 512, 512, 256, 5, "cornell")
render(rt.SCENE_MANYLIGHT, 0.3, 960, 540, 32, 3, "manylight")
render(rt.SCENE_INSTANCED, 0.3, 960, 540, 16, 3, "instanced", env=1)
