"""One rank's share of C5 at N = 8 (rank 0's 8-row bands of a 1080p frame), alone on a GPU, under different K1 schedules: which one
suits a share of 259 k pixels whose paths are sequential chains of up to 20 rays?  (measurement only)
usage: python tools/c5_share_sweep.py [--band-world 8] [--frames 8] [--overhead]"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, json, time
import raytracer2_b200 as rt

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=8); ap.add_argument("--overhead", action="store_true"); ap.add_argument("--band-world", type=int, default=8)
ap.add_argument("--scene", default="instanced"); ap.add_argument("--band-rows", default="8"); ap.add_argument("--band-rank", default="0")
ap.add_argument("--variants", type=int, default=0, help="only the first N knob sets (0 = all)")
a = ap.parse_args()
W, H = 1920, 1080
kind = {"instanced": rt.SCENE_INSTANCED, "hall": rt.SCENE_HALL}[a.scene]
sc = rt.Scene(kind, 1.0, 512 if kind == rt.SCENE_INSTANCED else 64, 0)
d = sc.desc.contents
cam = rt.overhead_camera(sc, W / H) if a.overhead else sc.camera(W / H)
g = rt.default_gconst(cam, W, H, sc.num_lights, env_size=(d.sky_width, d.sky_height))
g.refrence_mode = 1; g.environment = 1 if d.sky_width else 0

VARIANTS = [
    {}, {"RT2_CHAINS": "1"}, {"RT2_CHAINS": "2"}, {"RT2_CHAINS": "4"}, {"RT2_CHAINS": "16"},
    {"RT2_RAYS_PER_LANE": "0"}, {"RT2_RAYS_PER_LANE": "2"}, {"RT2_RAYS_PER_LANE": "4"}, {"RT2_RAYS_PER_LANE": "8"},
    {"RT2_CHAINS": "4", "RT2_RAYS_PER_LANE": "4"}, {"RT2_CHAINS": "2", "RT2_RAYS_PER_LANE": "4"},
    {"RT2_PATH_MODE": "1"}, {"RT2_PATH_MODE": "1", "RT2_WARP_P": "64"}, {"RT2_PATH_MODE": "1", "RT2_WARP_P": "128"}, {"RT2_PATH_MODE": "1", "RT2_WARP_P": "256"},
    {"RT2_PATH_MODE": "1", "RT2_WARP_P": "128", "RT2_WARP_BLOCKS": "4"}, {"RT2_PATH_MODE": "1", "RT2_WARP_P": "256", "RT2_WARP_BLOCKS": "4"},
    {"RT2_PATH_MODE": "1", "RT2_WARP_P": "512", "RT2_WARP_BLOCKS": "2"},
]
KNOBS = sorted({k for v in VARIANTS for k in v})
if a.variants:
    VARIANTS = VARIANTS[:a.variants]
import itertools
for v, rows, rank in itertools.product(VARIANTS, [int(x) for x in a.band_rows.split(",")], [int(x) for x in a.band_rank.split(",")]):
    for k in KNOBS:
        _os.environ.pop(k, None)
    _os.environ.update(v)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
    if d.sky_width:
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    bw = a.band_world
    p = rt.LaunchParams(spp=4, max_bounces=4, band_rows=rows if bw > 1 else 0, band_rank=rank, band_world=bw if bw > 1 else 0)
    ms = []
    for f in range(1, a.frames + 1):
        ctx.sync(); t0 = time.perf_counter()
        lp.execute(f, p)
        ctx.sync(); ms.append((time.perf_counter() - t0) * 1e3)
    st = ctx.stats()
    best = sorted(ms[2:])[len(ms[2:]) // 2]
    print(json.dumps({"knobs": v, "band_rows": rows, "band_rank": rank, "ms_median": round(best, 3), "rays": st.rays_traced, "Mrays_s": round(st.rays_traced / best / 1e3, 1), "launches": st.kernel_launches,
                      "schedule": st.sample_schedule}), flush=True)
    ctx.close()
