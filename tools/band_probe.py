"""per-GPU throughput of one rank's share of the bench workload (rank 0 of band_world N, spp = 4 N) for several chain counts:
emulates the weak-scaling bench on one GPU (no gather)."""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import os, sys, torch
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
for chains in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0").split(",")]:
    if chains: os.environ["RT2_CHAINS"] = str(chains)
    else: os.environ.pop("RT2_CHAINS", None)
    ctx = rt.Renderer(0)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
    row = []
    for n in (1, 2, 4, 8):
        p = rt.LaunchParams(spp=4 * n, max_bounces=4, band_rows=8 if n > 1 else 0, band_rank=0, band_world=n)
        for i in range(4): lp.execute(1 + i, p); ctx.sync()
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(5): lp.execute(10 + i, p)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5; rays = ctx.stats().rays_traced
        row.append(f"N={n}: {rays / ms / 1e3:7.1f}")
    print(f"chains {chains or 'auto'}:  " + "   ".join(row) + "   Mrays/s per GPU")
    ctx.close()
