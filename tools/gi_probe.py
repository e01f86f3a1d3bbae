"""device time of the ReSTIR frame variants on C2 at 1080p: default GI chain, + temporal/spatial, + DI (not the bench metric)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import json, torch
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
cam = sc.camera(W / H)
ctx = rt.Renderer(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H)
rt.PrepareLightsTasks(ctx).execute()
rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
out = {}
for name, kw in [("default_gi", {}), ("gi_temporal_spatial", dict(t=1, s=1)), ("gi_ts_raytraced_bias", dict(t=1, s=1, bias=3)), ("di_only", dict(gi=0, di=1)), ("di_plus_gi", dict(di=1))]:
    g = rt.default_gconst(cam, W, H, sc.num_lights)
    g.refrence_mode = 0; g.enable_restir_gi = kw.get("gi", 1); g.enable_restir_di = kw.get("di", 0)
    g.enable_temporal_resampling = kw.get("t", 0); g.enable_spatial_resampling = kw.get("s", 0)
    g.restir_gi.temporal_resampling_params.temporal_bias_correction_mode = kw.get("bias", 2)
    g.restir_gi.spatial_resampling_params.spatial_bias_correction_mode = kw.get("bias", 2)
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = 4
    g.restir_di.initial_sampling_params.num_primary_brdf_samples = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    for f in range(1, 4):
        rt.advance_frame(g, cam, W, H, f); lp.update_uniform(g); lp.execute(f)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for f in range(4, 4 + n):
        rt.advance_frame(g, cam, W, H, f); lp.update_uniform(g); lp.execute(f)
    e1.record(); torch.cuda.synchronize()
    st = ctx.stats()
    ms = e0.elapsed_time(e1) / n
    out[name] = {"ms_per_frame": round(ms, 3), "rays_per_frame": st.rays_traced, "Mrays_per_s": round(st.rays_traced / ms / 1e3, 1), "launches": st.kernel_launches}
print(json.dumps(out, indent=1))
