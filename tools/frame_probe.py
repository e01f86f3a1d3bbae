"""per-frame device time of K1 on C2 for a range of frame indices (looking for slow frames)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import sys, torch, numpy as np
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
ctx = rt.Renderer(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
p = rt.LaunchParams(spp=4, max_bounces=4)
for i in range(3):
    lp.execute(1 + i, p); ctx.sync()
out = []
for f in list(range(1, 33)) + [100, 101, 102, 103, 104, 200, 201, 300, 301, 302, 303, 304, 500]:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); lp.execute(f, p); e1.record(); torch.cuda.synchronize()
    out.append((f, round(e0.elapsed_time(e1), 2), ctx.stats().rays_traced))
print(out)
pp = rt.LaunchParams(spp=4, max_bounces=4, flags=rt.LAUNCH_PROFILE)
slow = max(out, key=lambda t: t[1])[0]
for f in (slow, 1):
    lp.execute(f, pp); ctx.sync(); st = ctx.stats()
    print("frame", f, "trace_ms", st.trace_ns / 1e6, "shade_ms", st.shade_ns / 1e6)
