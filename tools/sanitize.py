"""Workload for compute-sanitizer (memcheck / racecheck / initcheck): small frames through every pass family.
Stands in for the reference's always-on Vulkan validation layer (src/context.rs:139-171).  Run as
    compute-sanitizer --tool memcheck python tools/sanitize.py
RT2_CHAINS=8 is set by the caller to force the multi-stream wavefront chains on the small hall."""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import numpy as np
import raytracer2_b200 as rt


def k1(kind, scale, w, h, spp, bounces, env=0, tex=32):
    sc = rt.Scene(kind, scale, tex, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1; g.environment = env
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    lp.trace_primary()
    lp.execute(1, rt.LaunchParams(spp=spp, max_bounces=bounces))
    lp.execute(2, rt.LaunchParams(spp=spp, max_bounces=bounces, band_rows=8, band_rank=1, band_world=3))
    rt.PostProcessPass(ctx).execute(1)
    ctx.sync()
    s = ctx.stats()
    ctx.close()
    return s.rays_traced


def restir(kind, scale, w, h, tex=32):
    sc = rt.Scene(kind, scale, tex, 0)
    cam = sc.camera(w / h)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    total = 0
    for kw in ({}, dict(t=1, s=1), dict(t=1, s=1, bias=3), dict(gi=0, di=1), dict(di=1)):
        g = rt.default_gconst(cam, w, h, sc.num_lights)
        g.refrence_mode = 0; g.enable_restir_gi = kw.get("gi", 1); g.enable_restir_di = kw.get("di", 0)
        g.enable_temporal_resampling = kw.get("t", 0); g.enable_spatial_resampling = kw.get("s", 0)
        g.restir_gi.temporal_resampling_params.temporal_bias_correction_mode = kw.get("bias", 2)
        g.restir_gi.spatial_resampling_params.spatial_bias_correction_mode = kw.get("bias", 2)
        g.restir_di.initial_sampling_params.num_primary_local_light_samples = 4
        g.restir_di.initial_sampling_params.num_primary_brdf_samples = 1
        g.restir_di.initial_sampling_params.enable_initial_visibility = 1
        g.restir_di.shading_params.enable_final_visibility = 1
        for f in range(1, 3):
            rt.advance_frame(g, cam, w, h, f); lp.update_uniform(g); lp.execute(f)
        ctx.sync()
        total += ctx.stats().rays_traced
    lp.execute_presampeling(3, False)
    ctx.sync()
    ctx.close()
    return total


if __name__ == "__main__":
    print("C1 512x512:", k1(rt.SCENE_CORNELL, 1.0, 512, 512, 1, 2))
    print("hall 0.125, 641x360 (odd width, chains):", k1(rt.SCENE_HALL, 0.125, 641, 360, 2, 3))
    print("instanced 0.05 + sky:", k1(rt.SCENE_INSTANCED, 0.05, 320, 180, 2, 3, env=1))
    print("ReSTIR frames (GI, GI+K6/K7, ray-traced bias, DI, DI+GI, K11) hall 0.125 320x180:", restir(rt.SCENE_HALL, 0.125, 320, 180))
    print("sanitize workload done")
