import torch, json
import raytracer2_b200 as rt
W,H=1920,1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
ctx = rt.Renderer(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
import os
for n in (1,8):
    p = rt.LaunchParams(spp=4*n, max_bounces=4, band_rows=8 if n>1 else 0, band_rank=0, band_world=n)
    for i in range(3): lp.execute(1+i, p)
    ctx.sync()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(5): lp.execute(10+i, p)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/5; rays=ctx.stats().rays_traced
    print(n, round(ms,3), "ms", rays, round(rays/ms/1e3,1), "Mrays/s per GPU")
    pp = rt.LaunchParams(spp=4*n, max_bounces=4, band_rows=8 if n>1 else 0, band_rank=0, band_world=n, flags=rt.LAUNCH_PROFILE)
    lp.execute(50, pp); ctx.sync(); st=ctx.stats()
    print("   trace_ms", round(st.trace_ns/1e6,3), "shade_ms", round(st.shade_ns/1e6,3), "launches", st.trace_launches, "per-launch trace us", round(st.trace_ns/1e3/st.trace_launches,1), "shade us", round(st.shade_ns/1e3/st.trace_launches,1))
