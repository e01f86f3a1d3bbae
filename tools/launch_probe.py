"""host enqueue time of rt2_render vs its device time for one rank's share of an N-way sharded 1080p frame (is the wavefront loop launch-bound?)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import time, torch
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
ctx = rt.Renderer(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
for n in (1, 2, 4, 8):
    p = rt.LaunchParams(spp=4 * n, max_bounces=4, band_rows=8 if n > 1 else 0, band_rank=0, band_world=n)
    for i in range(3): lp.execute(1 + i, p)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host = 0.0
    e0.record()
    for i in range(6):
        t0 = time.perf_counter(); lp.execute(10 + i, p); host += time.perf_counter() - t0
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 6; st = ctx.stats()
    print(f"rank 0 of {n}: device {ms:.2f} ms/step, host enqueue {host / 6 * 1e3:.2f} ms/step, {st.kernel_launches} launches, {st.rays_traced / ms / 1e3:.1f} Mrays/s")
