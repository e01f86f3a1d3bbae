"""BASELINE config C5 as specified: the 20.5 M-triangle instanced scene (10 000 instances of one 2048-triangle BLAS in a TLAS),
sky on, 1080p, 4 spp x 4 bounces, K12 (environment PDF mips) once, then per frame K1 on this rank's 8-row bands -> K13 tonemap ->
NCCL gather of the RGBA8 bands (sharded K13 + gather of RGBA8).  STRONG scaling: the frame is fixed.  Run alone (N=1)
or under torch.distributed.run; rank 0 prints one JSON object.  --check compares the gathered RGBA8 image bit for bit with a
whole-frame render on rank 0."""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, json, os
import numpy as np, torch
import raytracer2_b200 as rt
from raytracer2_b200.sharding import BandPlan, gather_frame, unpack_gathered

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=10)
ap.add_argument("--check", action="store_true")
args = ap.parse_args()
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

W, H, BAND = 1920, 1080, 8
sc = rt.Scene(rt.SCENE_INSTANCED, 1.0, 512, 0)
d = sc.desc.contents
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights, env_size=(d.sky_width, d.sky_height))
g.refrence_mode = 1; g.environment = 1
ctx = rt.Renderer(local)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H); lp.update_uniform(g)
rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()      # K12, once
post = rt.PostProcessPass(ctx)
plan = BandPlan(H, BAND, world)
send = torch.zeros(plan.max_rows * W, dtype=torch.int32, device="cuda")
recv = torch.zeros(world * plan.max_rows * W, dtype=torch.int32, device="cuda")
params = rt.LaunchParams(spp=4, max_bounces=4, band_rows=BAND if world > 1 else 0, band_rank=rank, band_world=world)


rays_last = [0]


def frame(f, p):
    lp.execute(f, p)                 # K1 on this rank's bands
    rays_last[0] = f
    post.execute(f)                  # K13 (rows of other ranks hold stale data and are not gathered)
    if world > 1:
        ctx.pack_bands(rt.BUF_OUTPUT_RGBA8, send.data_ptr(), plan.rows(rank) * W * 4, BAND, rank, world)
        gather_frame(plan, rank, send, recv, dist)


for f in range(1, 4):
    frame(f, params)
lp.execute(3, params); ctx.sync()
rays_rank = float(ctx.stats().rays_traced)        # rays of one frame of this rank (frames differ by a fraction of a percent)
torch.cuda.synchronize()
if dist: dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for f in range(10, 10 + args.frames):
    frame(f, params)
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
rays = torch.tensor([rays_rank], dtype=torch.float64, device="cuda")
if dist:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(rays, op=dist.ReduceOp.SUM)
ms, rays = float(ms.item()), float(rays.item())
out = {"config": f"C5: {sc.num_triangles} instanced tris ({d.num_instances} instances of one BLAS), sky {d.sky_width}x{d.sky_height}, {W}x{H}, 4 spp x 4 bounces, K1 + K13 per frame, "
                 f"{BAND}-row bands over {world} GPU(s), RGBA8 bands gathered per frame",
       "n_gpus": world, "rays_per_frame": rays, "ms_per_frame": round(ms / args.frames, 3), "Mrays_s": round(rays * args.frames / ms / 1e3, 1)}
if args.check and world > 1:
    torch.cuda.synchronize()
    last = 10 + args.frames - 1
    got = unpack_gathered(plan, recv.view(world * plan.max_rows, W).cpu().numpy(), np.zeros((H, W), np.int32))
    if rank == 0:
        lp.execute(last, rt.LaunchParams(spp=4, max_bounces=4)); post.execute(last)
        full = ctx.read(rt.BUF_OUTPUT_RGBA8, np.int32).reshape(H, W)
        out["sharded_equals_whole_frame_bit_exact"] = bool(np.array_equal(full, got))
if rank == 0:
    print(json.dumps(out))
if dist:
    dist.barrier(); dist.destroy_process_group()
