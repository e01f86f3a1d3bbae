"""one K1 frame shape under ncu's launch list: rank 0 of N bands of the bench workload (spp = 4 N), speculative or sequential.
usage: spec_profile.py N frames"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import sys, torch
import raytracer2_b200 as rt
n = int(sys.argv[1]); frames = int(sys.argv[2]) if len(sys.argv) > 2 else 2
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
ctx = rt.Renderer(0)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
p = rt.LaunchParams(spp=4 * n, max_bounces=4, band_rows=8 if n > 1 else 0, band_rank=0, band_world=n)
for f in range(frames):
    lp.execute(1 + f, p); ctx.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); lp.execute(9, p); ctx.sync(); e1.record(); torch.cuda.synchronize()
print("rays", ctx.stats().rays_traced, "launches", ctx.stats().kernel_launches)
ctx.close()
