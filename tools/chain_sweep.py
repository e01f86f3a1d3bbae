import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import os, sys, torch
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
for rpl in (0, 2, 3, 4):
  for chains in (2, 3, 4, 6, 8):
    os.environ["RT2_CHAINS"] = str(chains); os.environ["RT2_RAYS_PER_LANE"] = str(rpl)
    ctx = rt.Renderer(0)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
    row = []
    for n in (2, 3, 6):
        p = rt.LaunchParams(spp=4 * n, max_bounces=4, band_rows=8 if n > 1 else 0, band_rank=0, band_world=n)
        for i in range(3): lp.execute(1 + i, p)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(4): lp.execute(10 + i, p)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 4; rays = ctx.stats().rays_traced
        row.append(f"N={n}: {rays / ms / 1e3:7.1f}")
    print(f"rays/lane {rpl} chains {chains}:  " + "   ".join(row))
    ctx.close()
