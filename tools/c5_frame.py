"""C5 (flattened instancing, sky on, 1080p, 4 spp x 4 bounces, K1 + K13) on one GPU, a few frames — the command ncu wraps for the
C5 trace-kernel capture (tools/gpu_prof_c5.sh) and a quick timing line.  usage: python tools/c5_frame.py [--frames N] [--overhead]"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, json, time
import numpy as np
import raytracer2_b200 as rt

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=6); ap.add_argument("--overhead", action="store_true"); ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--band-world", type=int, default=0, help="render only rank 0's 8-row bands of this many ranks (one rank of N, emulated)")
a = ap.parse_args()
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_INSTANCED, a.scale, 512, 0)
d = sc.desc.contents
cam = rt.overhead_camera(sc, W / H) if a.overhead else sc.camera(W / H)
g = rt.default_gconst(cam, W, H, sc.num_lights, env_size=(d.sky_width, d.sky_height))
g.refrence_mode = 1; g.environment = 1
ctx = rt.Renderer(0)
t0 = time.perf_counter()
rt.Model.from_scene(ctx, sc); ctx.sync()
print(f"upload + BVH build: {time.perf_counter() - t0:.2f} s", flush=True)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
post = rt.PostProcessPass(ctx)
p = rt.LaunchParams(spp=4, max_bounces=4, band_rows=8 if a.band_world > 1 else 0, band_rank=0, band_world=a.band_world if a.band_world > 1 else 0)
out = []
for f in range(1, a.frames + 1):
    ctx.sync(); t0 = time.perf_counter()
    lp.execute(f, p)
    ctx.sync(); st = ctx.stats()                 # K13 resets the per-pass statistics
    post.execute(f)
    ctx.sync(); dt = time.perf_counter() - t0
    out.append({"frame": f, "ms": round(dt * 1e3, 3), "rays": st.rays_traced, "hits": st.rays_hit, "launches": st.kernel_launches, "schedule": st.sample_schedule,
                "Mrays_s": round(st.rays_traced / dt / 1e6, 1)})
print(json.dumps(out))
ctx.close()
