"""where does the e2e step go? (host timers around each C-ABI call)"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import time, torch, numpy as np
import raytracer2_b200 as rt
W, H = 1920, 1080
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights); g.refrence_mode = 1
ctx = rt.Renderer(0)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
p = rt.LaunchParams(spp=4, max_bounces=4)
host = torch.empty(W * H * 4, dtype=torch.float16).pin_memory()
host_np = np.empty(W * H * 4, np.float16)
for i in range(3):
    lp.execute(1 + i, p); ctx.sync()
for name, dst in (("pinned", host.data_ptr()), ("pageable", host_np.ctypes.data)):
    for i in range(4):
        t0 = time.perf_counter(); lp.update_uniform(g)
        t1 = time.perf_counter(); lp.execute(10 + i, p)
        t2 = time.perf_counter(); ctx.sync()
        t3 = time.perf_counter(); ctx.read_into(rt.BUF_DIFFUSE_F16, dst, W * H * 8)
        t4 = time.perf_counter()
        print(f"{name} set_constants {1e3*(t1-t0):.2f} ms, launch {1e3*(t2-t1):.2f} ms, sync {1e3*(t3-t2):.2f} ms, read {1e3*(t4-t3):.2f} ms")
