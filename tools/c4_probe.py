"""BASELINE config C4: C2 hall at 3840x2160, 16 progressive frames x {4 spp, 4 bounces} with the reference's fp32
accumulation (enable_accumulation=1, blend_factor = 1/frame: src/main.rs:551-554), rows sharded over the ranks
(STRONG scaling: the 4K frame is fixed, each rank renders its interleaved 8-row bands), fp32 bands gathered over NCCL
(a) once per frame and (b) once after the last frame — both timed.  Run alone (N=1) or under torch.distributed.run.
Rank 0 prints one JSON object.  With --check the gathered image is compared bit for bit against a whole-frame render
of the same 16 frames on rank 0."""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import argparse, json, os
import numpy as np, torch
import raytracer2_b200 as rt
from raytracer2_b200.sharding import BandPlan, gather_frame, unpack_gathered

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=16)
ap.add_argument("--check", action="store_true")
args = ap.parse_args()
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

W, H, BAND = 3840, 2160, 8
sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights)
g.refrence_mode = 1; g.enable_accumulation = 1
ctx = rt.Renderer(local)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); ctx.set_stream(stream.cuda_stream)
rt.Model.from_scene(ctx, sc)
lp = rt.LightPasses(ctx, W, H)
plan = BandPlan(H, BAND, world)
send = torch.zeros(plan.max_rows * W * 4, dtype=torch.float32, device="cuda")
recv = torch.zeros(world * plan.max_rows * W * 4, dtype=torch.float32, device="cuda")


def sequence(params, gather_each_frame, gather_at_end=True):
    rays = 0
    for frame in range(1, args.frames + 1):
        g.blend_factor = 1.0 / frame; g.frame = frame
        lp.update_uniform(g)
        lp.execute(frame, params)
        last = frame == args.frames
        if world > 1 and (gather_each_frame or (last and gather_at_end)):
            ctx.pack_bands(rt.BUF_DIFFUSE_F32, send.data_ptr(), plan.rows(rank) * W * 16, BAND, rank, world)
            gather_frame(plan, rank, send, recv, dist)
    return rays


def timed(params, gather_each_frame):
    sequence(params, gather_each_frame)                      # warm-up pass over all frames
    torch.cuda.synchronize()
    if dist: dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); sequence(params, gather_each_frame); e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    rays = torch.tensor([float(ctx.stats().rays_traced)], dtype=torch.float64, device="cuda")   # last frame of this rank
    if dist:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    return float(ms.item()), float(rays.item())


params = rt.LaunchParams(spp=4, max_bounces=4, band_rows=BAND if world > 1 else 0, band_rank=rank, band_world=world)
ms_each, rays_frame = timed(params, True)
ms_end, _ = timed(params, False)
out = {"config": f"C4: hall 262144 tris, {W}x{H}, {args.frames} frames x 4 spp x 4 bounces accumulated in fp32, {BAND}-row bands over {world} GPU(s)",
       "n_gpus": world, "rays_per_frame": rays_frame,
       "gather_every_frame": {"ms_total": round(ms_each, 3), "ms_per_frame": round(ms_each / args.frames, 3), "Mrays_s": round(rays_frame * args.frames / ms_each / 1e3, 1)},
       "gather_at_end": {"ms_total": round(ms_end, 3), "ms_per_frame": round(ms_end / args.frames, 3), "Mrays_s": round(rays_frame * args.frames / ms_end / 1e3, 1)},
       "gather_bytes_per_frame": world * plan.max_rows * W * 16 if world > 1 else 0}

if args.check and world > 1:
    # bands of all ranks after the last frame == the whole-frame render of the same frames, bit for bit
    torch.cuda.synchronize()
    got = unpack_gathered(plan, recv.view(world * plan.max_rows, W, 4).cpu().numpy(), np.zeros((H, W, 4), np.float32))
    if rank == 0:
        ctx.write(rt.BUF_DIFFUSE_F32, np.zeros((H, W, 4), np.float32))
        sequence(rt.LaunchParams(spp=4, max_bounces=4), False, False)
        full = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(H, W, 4)
        out["sharded_equals_whole_frame_bit_exact"] = bool(np.array_equal(full.view(np.uint32), got.view(np.uint32)))
if rank == 0:
    print(json.dumps(out))
if dist:
    dist.barrier(); dist.destroy_process_group()
