"""BASELINE.md result table: device throughput of configs C1, C3, C4 (1 GPU share), C5 with the CPU oracle timed on a
bounded sample of the same workload.  Prints one JSON object.  (C2 is bench.py's own line.)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import json, os, sys, time
import numpy as np, torch
import raytracer2_b200 as rt
from oracle import oracle_ffi as O

out = {}
cores = os.cpu_count() or 1


def gpu_time(fn, n=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def setup(kind, scale, w, h, tex=512):
    sc = rt.Scene(kind, scale, tex, 0)
    ctx = rt.Renderer(0)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
    t0 = time.perf_counter(); rt.Model.from_scene(ctx, sc); ctx.sync(); build_s = time.perf_counter() - t0
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    lp = rt.LightPasses(ctx, w, h)
    return sc, ctx, g, lp, build_s


def k1(name, kind, scale, w, h, spp, bounces, cpu_rows=None, cpu_spp=1, env=0):
    sc, ctx, g, lp, build_s = setup(kind, scale, w, h)
    g.refrence_mode = 1; g.environment = env
    lp.update_uniform(g)
    p = rt.LaunchParams(spp=spp, max_bounces=bounces)
    frame = [1]
    def step():
        lp.execute(frame[0], p); frame[0] += 1
    ms = gpu_time(step)
    rays = ctx.stats().rays_traced
    osc = O.OracleScene(sc.desc)
    rows = cpu_rows or h
    t0 = time.perf_counter()
    crays = osc.render_reference(g, 1, w, h, spp=cpu_spp, max_bounces=bounces, row0=0, row1=rows, threads=cores)[3]
    cdt = time.perf_counter() - t0
    out[name] = {"tris": sc.num_triangles, "size": [w, h], "spp": spp, "bounces": bounces, "gpu_ms": round(ms, 3), "rays": rays,
                 "gpu_Mrays_s": round(rays / ms / 1e3, 1), "cpu_Mrays_s": round(crays / cdt / 1e6, 2), "cpu_cores": cores,
                 "cpu_sample": f"{rows} rows x {cpu_spp} spp, {crays} rays in {cdt:.1f} s", "scene_upload_and_bvh_build_s": round(build_s, 3)}
    ctx.close()


k1("C1_cornell_512_1spp_2b", rt.SCENE_CORNELL, 1.0, 512, 512, 1, 2, cpu_spp=16)
k1("C1_cornell_512_64spp_5b", rt.SCENE_CORNELL, 1.0, 512, 512, 64, 5, cpu_spp=16)
k1("C4_hall_4k_4spp_4b_one_gpu", rt.SCENE_HALL, 1.0, 3840, 2160, 4, 4, cpu_rows=540, cpu_spp=1)
k1("C5_instanced_1080p_4spp_4b_env", rt.SCENE_INSTANCED, 1.0, 1920, 1080, 4, 4, cpu_rows=270, cpu_spp=1, env=1)

# C3: many lights — K10 + K12 + K3 + K9 per frame
w, h = 1920, 1080
sc, ctx, g, lp, build_s = setup(rt.SCENE_MANYLIGHT, 1.0, w, h)
g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
g.restir_di.initial_sampling_params.num_primary_local_light_samples = 8
g.restir_di.initial_sampling_params.num_primary_brdf_samples = 1
g.restir_di.initial_sampling_params.enable_initial_visibility = 1
g.restir_di.shading_params.enable_final_visibility = 1
pl = rt.PrepareLightsTasks(ctx); mp = rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS)
lp.update_uniform(g)
ms_prep = gpu_time(lambda: (pl.execute(), mp.execute_mip_generation()))
frame = [1]
def step3():
    lp.execute(frame[0]); frame[0] += 1
ms = gpu_time(step3)
rays = ctx.stats().rays_traced
out["C3_manylight_1080p_K3_K9"] = {"tris": sc.num_triangles, "lights": sc.num_lights, "gpu_ms_K3_K9": round(ms, 3), "gpu_ms_K10_K12": round(ms_prep, 3), "rays": rays,
                                   "gpu_Mrays_s": round(rays / ms / 1e3, 1), "light_candidates_per_s_G": round(8 * w * h / ms / 1e6, 2),
                                   "scene_upload_and_bvh_build_s": round(build_s, 3)}
ctx.close()
print(json.dumps(out, indent=1))
