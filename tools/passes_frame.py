"""A few frames of (a) C3: 1 M triangles / 10 000 lights, 1080p, K3 + K9 with 8 local + 1 BRDF candidates and both visibility rays, or
(b) the default ReSTIR-GI frame on the C2 hall (K3 -> K4 -> K5 -> K8), or (c) the same with K6 + K7 — the command ncu wraps for the
per-kernel captures of the ReSTIR passes (tools/gpu_prof_passes.sh).  usage: python tools/passes_frame.py c3|gi|gi67 [--frames N]"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, json, time
import raytracer2_b200 as rt

ap = argparse.ArgumentParser()
ap.add_argument("what", choices=["c3", "gi", "gi67"]); ap.add_argument("--frames", type=int, default=4)
a = ap.parse_args()
W, H = 1920, 1080
ctx = rt.Renderer(0)
if a.what == "c3":
    sc = rt.Scene(rt.SCENE_MANYLIGHT, 1.0, 64, 0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H)
    cam = sc.camera(W / H)
    g = rt.default_gconst(cam, W, H, sc.num_lights)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 8; isp.num_primary_brdf_samples = 1; isp.enable_initial_visibility = 1
    g.restir_di.shading_params.enable_final_visibility = 1
else:
    sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H)
    cam = sc.camera(W / H)
    g = rt.default_gconst(cam, W, H, sc.num_lights)
    g.refrence_mode = 0
    if a.what == "gi67":
        g.enable_temporal_resampling = 1; g.enable_spatial_resampling = 1
rt.PrepareLightsTasks(ctx).execute()
rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
out = []
for f in range(1, a.frames + 1):
    rt.advance_frame(g, cam, W, H, f); lp.update_uniform(g)
    ctx.sync(); t0 = time.perf_counter()
    lp.execute(f)
    ctx.sync(); dt = time.perf_counter() - t0
    st = ctx.stats()
    out.append({"frame": f, "ms": round(dt * 1e3, 3), "rays": st.rays_traced, "launches": st.kernel_launches})
print(json.dumps(out))
ctx.close()
