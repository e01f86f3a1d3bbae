"""Strong scaling through rt2_group_* in ONE process (rt2_group_create: the form a Rust host without a launcher uses): a fixed frame is
split into interleaved 8-row bands over N = 1, 2, 4, 8 GPUs of the box; the assembled frame is gathered every frame.  Workloads:
  c2s  C2 hall, 1920x1080, 4 spp TOTAL x 4 bounces, K1, rgba16f gather (the strong-scaling twin of bench.py's weak step)
  c4   C2 hall, 3840x2160, 16 accumulated frames x {4 spp, 4 bounces} (enable_accumulation, blend 1/frame), fp32 gather per frame
  c5   20.5 M instanced triangles, sky, 1920x1080, 4 spp x 4 bounces, K1 -> K13 tonemap -> RGBA8 gather (BASELINE config C5)
  gi   C2 hall, 1920x1080, the ReSTIR-GI chain with K6 + K7 on (phased frame + in-place exchanges), rgba16f gather
Timing: CUDA events on every device's render stream around the timed frames (after warm-up frames and a full sync); the figure is the
MAX over devices.  The last assembled frame of every N is compared with N = 1's bit for bit.
usage: python tools/strong_scaling.py [--workloads c2s,c4,c5,gi] [--ns 1,2,4,8] [--frames K]"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, hashlib, json, time
import numpy as np
import torch
import raytracer2_b200 as rt

ap = argparse.ArgumentParser()
ap.add_argument("--workloads", default="c2s,c4,c5,gi"); ap.add_argument("--ns", default="1,2,4,8")
ap.add_argument("--frames", type=int, default=0, help="timed frames (0 = the workload's default)"); ap.add_argument("--scale", type=float, default=1.0)
a = ap.parse_args()
ndev = torch.cuda.device_count()
NS = [n for n in (int(x) for x in a.ns.split(",")) if n <= ndev]


def workload(name):
    if name in ("c2s", "c4", "gi"):
        sc = rt.Scene(rt.SCENE_HALL, a.scale, 512, 0)
        W, H = (3840, 2160) if name == "c4" else (1920, 1080)
        cam = sc.camera(W / H)
        g = rt.default_gconst(cam, W, H, sc.num_lights)
        if name == "gi":
            g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 0
            g.enable_temporal_resampling = 1; g.enable_spatial_resampling = 1
            return dict(sc=sc, W=W, H=H, g=g, cam=cam, params=None, gather=rt.BUF_DIFFUSE_F16, frames=10, post=False, accumulate=False, move=True,
                        desc="C2 hall 1080p, ReSTIR-GI chain K3 K4 K5 K6 K7 K8 (temporal + spatial resampling), rgba16f gather")
        g.refrence_mode = 1
        if name == "c4":
            g.enable_accumulation = 1
            return dict(sc=sc, W=W, H=H, g=g, cam=cam, params=rt.LaunchParams(spp=4, max_bounces=4), gather=rt.BUF_DIFFUSE_F32, frames=16, post=False, accumulate=True, move=False,
                        desc="C4: C2 hall 3840x2160, 16 accumulated frames x {4 spp, 4 bounces}, fp32 gather per frame")
        return dict(sc=sc, W=W, H=H, g=g, cam=cam, params=rt.LaunchParams(spp=4, max_bounces=4), gather=rt.BUF_DIFFUSE_F16, frames=10, post=False, accumulate=False, move=False,
                    desc="C2 hall 1080p, 4 spp total x 4 bounces (K1), rgba16f gather")
    if name == "c5":
        sc = rt.Scene(rt.SCENE_INSTANCED, a.scale, 512, 0)
        d = sc.desc.contents
        W, H = 1920, 1080
        cam = sc.camera(W / H)
        g = rt.default_gconst(cam, W, H, sc.num_lights, env_size=(d.sky_width, d.sky_height))
        g.refrence_mode = 1; g.environment = 1
        return dict(sc=sc, W=W, H=H, g=g, cam=cam, params=rt.LaunchParams(spp=4, max_bounces=4), gather=rt.BUF_OUTPUT_RGBA8, frames=10, post=True, accumulate=False, move=False,
                    desc="C5: 10 000 x 2048 instanced triangles (flattened), sky, 1080p, 4 spp x 4 bounces, K1 -> K13 -> RGBA8 gather")
    raise SystemExit(f"unknown workload {name}")


def run(wl, n):
    sc, W, H, g = wl["sc"], wl["W"], wl["H"], wl["g"]
    frames = a.frames or wl["frames"]
    t0 = time.time()
    grp = rt.Group(list(range(n)))
    streams = []
    for i in range(n):                       # real streams, so that torch's events see the library's launches
        with torch.cuda.device(i):
            s = torch.cuda.Stream(device=i); streams.append(s)
            grp.ctx(i).set_stream(s.cuda_stream)
    grp.upload_scene(sc.desc); grp.resize(W, H)
    if wl["post"]:
        grp.generate_mips(rt.GenerateMipsPass.ENVIRONMENT)
    setup_s = time.time() - t0
    cam = sc.camera(W / H)

    def sequence(first_frame, count):
        for k in range(count):
            f = first_frame + k
            if wl["accumulate"]:
                g.blend_factor = 1.0 / (k + 1)
            if wl["move"]:
                cam.position[0] += 0.02
                rt.advance_frame(g, cam, W, H, f)
            g.frame = f
            grp.update_uniform(g)
            if wl["post"]:
                grp.execute(f, wl["params"], gather=-1); grp.postprocess(f); grp.gather(wl["gather"])
            else:
                grp.execute(f, wl["params"], gather=wl["gather"])

    warm = 4 if not wl["accumulate"] else frames          # accumulation: one whole warm-up sequence
    for k in range(warm if not wl["accumulate"] else 1):
        if wl["accumulate"]:
            sequence(1, frames)
        else:
            sequence(1 + k, 1)
        grp.sync()                                        # frame by frame: the sample-schedule policy reads the previous frame's statistics
    cam = sc.camera(W / H)
    ev = []
    for i in range(n):
        with torch.cuda.device(i):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(streams[i]); ev.append((e0, e1))
    sequence(1 if wl["accumulate"] else 100, frames)
    grp.wait_gather()
    for i in range(n):
        with torch.cuda.device(i):
            ev[i][1].record(streams[i])
    grp.sync()
    for i in range(n):
        torch.cuda.synchronize(i)
    ms = max(e0.elapsed_time(e1) for e0, e1 in ev)
    rays = sum(grp.ctx(i).stats().rays_traced for i in range(n)) if not wl["post"] else None
    if wl["post"]:                                        # K13 resets the pass statistics: count the rays of one more K1 frame
        grp.execute(100 + frames - 1, wl["params"], gather=-1); grp.sync()
        rays = sum(grp.ctx(i).stats().rays_traced for i in range(n))
        grp.postprocess(100 + frames - 1); grp.gather(wl["gather"])
    frame = grp.read(np.uint8)
    digest = hashlib.sha256(frame.tobytes()).hexdigest()[:16]
    grp.close()
    return dict(n_gpus=n, ms_per_frame=round(ms / frames, 4), frames=frames, rays_per_frame=rays, Mrays_s=round(rays / (ms / frames) / 1e3, 1) if rays else None,
                frame_sha=digest, setup_s=round(setup_s, 1))


out = {}
for name in a.workloads.split(","):
    wl = workload(name)
    rows = []
    for n in NS:
        r = run(wl, n)
        r["speedup"] = round(rows[0]["ms_per_frame"] / r["ms_per_frame"], 3) if rows else 1.0
        r["efficiency"] = round(r["speedup"] / (n / NS[0]), 3)
        r["equals_first_bit_exact"] = (r["frame_sha"] == rows[0]["frame_sha"]) if rows else True
        rows.append(r)
        print(name, json.dumps(r), flush=True)
    out[name] = {"workload": wl["desc"], "rows": rows}
print(json.dumps(out))
