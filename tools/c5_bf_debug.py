import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))); _sys.path.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "tests"))
import numpy as np, time
import raytracer2_b200 as rt
from oracle import oracle_ffi as O
from test_gpu_fullsize import _primary_rays
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_INSTANCED, 1.0, 64, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1; g.environment = 1
osc = O.OracleScene(sc.desc)
ctx = rt.Renderer(0); rt.Model.from_scene(ctx, sc); lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
lp.trace_primary()
hits = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(H, W, 8)
ref = osc.trace_primary(g, W, H)
print("gpu vs oracle bvh, whole frame:", int((hits != ref).any(axis=2).sum()))
xs = np.arange(0, W, 8); ys = np.arange(800, 801)
rays = _primary_rays(g, xs, ys).reshape(-1, 8)
bf = osc.trace_rays(rays, flags=O.ORC_BRUTE_FORCE)
bv = osc.trace_rays(rays, flags=0)
got = hits[800, ::8]
print("trace_rays bvh vs gpu:", int((bv != got).any(axis=1).sum()), " bf vs gpu:", int((bf != got).any(axis=1).sum()))
for i in np.flatnonzero((bf != got).any(axis=1))[:10]:
    print(i, "gpu", got[i, :3], got[i, 3:6].view(np.float32), "| bf", bf[i, :3], bf[i, 3:6].view(np.float32), "| bvh", bv[i, :3], bv[i, 3:6].view(np.float32))
    print("   ray", rays[i])
