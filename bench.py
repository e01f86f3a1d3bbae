#!/usr/bin/env python
"""bench.py — Mrays/s of the reference path tracer (K1, refrence.rgen) on BASELINE config C2:
procedural Sponza-scale hall, 262 144 triangles, 64 textured materials, 1920x1080, 4 bounces,
4 spp per GPU.

  python bench.py --gpus N --steps K --warmup W          (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference ...                   (CPU oracle on the host cores; see below)

A step is one rt2_render of the whole 1080p frame.  With N ranks the frame is split into interleaved
16-row bands, every rank traces 4 spp x N (so per-GPU work is fixed: weak scaling), and
the fp32 bands are all-gathered with NCCL over NVLink inside the timed region.
`value` is device-timed with inputs resident; `e2e` goes through the C ABI with host buffers:
rt2_set_constants (pinned host -> device) + rt2_render + rt2_read_async of the rgba16f accumulation image into
pinned host memory, every step, two frames in flight (the transfer of frame n overlaps the render of frame n+1).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, SPP, BOUNCES = 1920, 1080, 4, 4
BAND_ROWS = 8
TRIS = 262144
# algorithmic bytes per ray, SURVEY.md §8(d): B_ray = 80*D + 96 + h*(12+192+128+16*tex), D = ceil(log8(T/2)) = 6
B_TRAVERSE = 80 * 6 + 96          # k_trace's share: 576 B/ray
B_SHADE = 12 + 192 + 128 + 16     # k_shade's share: 348 B/ray
B_RAY = B_TRAVERSE + B_SHADE      # 924 B/ray


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum per k_trace launch from the committed ncu --set full capture
    (profiles/*_summary.json, same workload, N=1); None when no capture is committed."""
    import glob
    for p in sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r*_summary.json")), reverse=True):
        try:
            return json.load(open(p))["traffic_bytes_per_launch"]
        except Exception:
            continue
    return None


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML in-process every
    5 ms (nvidia-smi -lms needs ~200 ms to start, longer than a short timed region); nvidia-smi is the fallback."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu, self.nvml, self.h, self.stop_flag = [], None, gpu_index, None, None, False

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
        return pynvml, h

    def start(self):
        try:
            self.nvml, self.h = self._nvml_handle()
            self.max_mhz = self.nvml.nvmlDeviceGetMaxClockInfo(self.h, self.nvml.NVML_CLOCK_SM)
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                mhz = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
                try:
                    bits = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    bits = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((mhz, bits))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.t.join(timeout=1)
            sm = sorted(r[0] for r in self.rows)
            reasons = sorted({name for _, bits in self.rows for bit, name in self.REASONS if bits & bit})
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(sm), "source": "nvml, 5 ms"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm), "source": "nvidia-smi -lms 100"}


def cpu_oracle_run(spp, bounces, threads=0, rows=None):
    """Times the CPU oracle (the restatement of the reference's shader math; the reference has no CPU
    implementation of this path) on the C2 scene.  Returns (Mrays/s, rays, seconds, cores)."""
    import raytracer2_b200 as rt
    from oracle import oracle_ffi as O
    sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
    g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights)
    g.refrence_mode = 1
    osc = O.OracleScene(sc.desc)
    cores = threads or os.cpu_count() or 1
    r1 = rows or H
    osc.render_reference(g, 1, W, H, spp=1, max_bounces=1, row0=0, row1=min(8, r1), threads=cores)   # warm
    t0 = time.perf_counter()
    _, _, _, rays = osc.render_reference(g, 1, W, H, spp=spp, max_bounces=bounces, row0=0, row1=r1, threads=cores)
    dt = time.perf_counter() - t0
    return rays / dt / 1e6, rays, dt, cores, (osc, g)


def run_reference_arm(args, rank, world):
    if rank != 0:
        return 0
    # bounded sample per step: the full 1080p frame at 1 spp x 4 bounces (8.3 M rays), all host threads
    import raytracer2_b200 as rt
    from oracle import oracle_ffi as O
    sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
    g = rt.default_gconst(sc.camera(W / H), W, H, sc.num_lights)
    g.refrence_mode = 1
    osc = O.OracleScene(sc.desc)
    cores = os.cpu_count() or 1
    rows = H
    for _ in range(args.warmup):
        osc.render_reference(g, 1, W, H, spp=1, max_bounces=BOUNCES, row0=0, row1=64, threads=cores)
    rays_total, t0 = 0, time.perf_counter()
    for s in range(args.steps):
        _, _, _, rays = osc.render_reference(g, 1 + s, W, H, spp=1, max_bounces=BOUNCES, row0=0, row1=rows, threads=cores)
        rays_total += rays
    dt = time.perf_counter() - t0
    v = rays_total / dt / 1e6
    sample = f"{W}x{rows} rows, 1 spp x {BOUNCES} bounces per step ({rays_total // max(args.steps,1)} rays/step), {cores} threads"
    line = {"impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / max(args.steps, 1) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2 procedural hall {sc.num_triangles} tris, 64 sRGB textures 512^2, {W}x{H}, {BOUNCES} bounces, {SPP} spp per GPU (spp={SPP * max(args.gpus, 1)})",
                       "arm": "CPU oracle (SAH BVH2, std::thread over rows): the reference itself is Rust + GLSL on Vulkan ray tracing, has no CPU path and cannot run on this box; each step is a bounded sample of the workload, 1 spp of the same frame",
                       "rays_per_step": rays_total // max(args.steps, 1)},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--scale", type=float, default=1.0, help="scene scale (1.0 = BASELINE size); for quick checks only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference_arm(args, rank, world)

    import numpy as np
    import torch
    import raytracer2_b200 as rt

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the path has no CPU fallback (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world
    spp = SPP * n_gpus     # fixed 4 spp of work per GPU

    sc = rt.Scene(rt.SCENE_HALL, args.scale, 512, 0)
    cam = sc.camera(W / H)
    g = rt.default_gconst(cam, W, H, sc.num_lights)
    g.refrence_mode = 1
    ctx = rt.Renderer(local_rank)
    # a real (non-default) torch stream: the library, torch's events and NCCL all run on it, so
    # torch.cuda.Event timing sees exactly the kernels librt2cuda launches
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx.set_stream(stream.cuda_stream)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H)
    lp.update_uniform(g)
    params = rt.LaunchParams(spp=spp, max_bounces=BOUNCES, band_rows=BAND_ROWS if n_gpus > 1 else 0, band_rank=rank, band_world=n_gpus)

    # band gather plumbing (raytracer2_b200/sharding.py): each rank's rows packed contiguously -> all_gather
    from raytracer2_b200.sharding import BandPlan, gather_frame
    plan = BandPlan(H, BAND_ROWS, n_gpus)
    # two send/recv pairs: the gather of frame n (NCCL's own stream) overlaps the render of frame n+1; a pair is reused only
    # after the collective that last used it has completed (handle.wait() = a stream-level wait, the host never blocks)
    send = [torch.zeros(plan.max_rows * W * 4, dtype=torch.float32, device="cuda") for _ in range(2)]
    recv = [torch.zeros(n_gpus * plan.max_rows * W * 4, dtype=torch.float32, device="cuda") if n_gpus > 1 else None for _ in range(2)]
    pending = [None, None]
    step_no = [0]

    def step(frame):
        lp.execute(frame, params)
        if n_gpus > 1:
            b = step_no[0] & 1
            step_no[0] += 1
            if pending[b] is not None:
                pending[b].wait()
            ctx.pack_bands(rt.BUF_DIFFUSE_F32, send[b].data_ptr(), plan.rows(rank) * W * 16, BAND_ROWS, rank, n_gpus)
            pending[b] = gather_frame(plan, rank, send[b], recv[b], dist, async_op=True)

    def drain():
        for b in range(2):
            if pending[b] is not None:
                pending[b].wait()
                pending[b] = None

    for i in range(args.warmup):
        step(1 + i)
        ctx.sync()          # warm-up frames complete one by one: the library picks its sample schedule from the previous frame's
                            # miss statistics and allocates the speculative path buffers on the first frame that uses them
    drain()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = 0
    launches = 0
    ev0.record()
    for i in range(args.steps):
        step(100 + i)
    drain()                 # every gather has landed before the end event
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    st = ctx.stats()
    rays_per_step = st.rays_traced
    launches = st.kernel_launches * args.steps
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None

    # per-kernel time of the dominant kernel (k_trace): one profiled step with CUDA events around each launch
    prof = rt.LaunchParams(spp=spp, max_bounces=BOUNCES, band_rows=params.band_rows, band_rank=rank, band_world=n_gpus, flags=rt.LAUNCH_PROFILE)
    lp.execute(500, prof)
    ctx.sync()
    pst = ctx.stats()
    trace_ms, shade_ms, trace_launches = pst.trace_ns / 1e6, pst.shade_ns / 1e6, max(1, pst.trace_launches)

    # ---- e2e through the C ABI with host buffers --------------------------------------------------------------
    # Two frames in flight, as the reference's swapchain loop has (src/context.rs:30,639-694): frame n's image travels to
    # pinned host memory on the context's copy stream (rt2_read_async) while frame n+1 renders; every step still uploads its
    # GConst from pinned staging and reads its result back, and the timed region ends only after the last transfer landed.
    host_out = [torch.empty(W * H * 4, dtype=torch.float16).pin_memory() for _ in range(2)]
    for i in range(2):
        lp.update_uniform(g); step(200 + i); ctx.read_into_async(rt.BUF_DIFFUSE_F16, host_out[i & 1].data_ptr(), W * H * 8)
    drain()
    ctx.sync()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        g.frame = 300 + i
        lp.update_uniform(g)                                                                  # pinned host -> device, 1424 B
        step(300 + i)
        ctx.read_into_async(rt.BUF_DIFFUSE_F16, host_out[i & 1].data_ptr(), W * H * 8)        # device -> pinned host
    drain()
    ctx.read_fence()                                                                          # e1 waits for every transfer
    e1.record()
    ctx.sync()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1)

    t = torch.tensor([ms, e2e_ms, float(rays_per_step), float(launches), trace_ms, shade_ms], dtype=torch.float64, device="cuda")
    if dist:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, e2e_ms = tmax[0].item(), tmax[1].item()
        total_rays_step, launches = tsum[2].item(), int(tsum[3].item())
    else:
        total_rays_step = float(rays_per_step)
    value = total_rays_step * args.steps / (ms * 1e-3) / 1e6
    e2e_value = total_rays_step * args.steps / (e2e_ms * 1e-3) / 1e6
    if rank == 0:
        peak, peak_kind = measured_peak_hbm()
        rays_per_launch = rays_per_step / trace_launches
        achieved = rays_per_launch * B_TRAVERSE / (trace_ms / trace_launches * 1e-3) / 1e9
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2 procedural hall {sc.num_triangles} tris, 64 sRGB textures 512^2, {W}x{H}, {BOUNCES} bounces, {SPP} spp per GPU (spp={spp})",
                       "sharding": "whole frame" if n_gpus == 1 else f"{BAND_ROWS}-row bands interleaved over {n_gpus} ranks, scene replicated, NCCL all_gather of fp32 bands per step",
                       "l2": "no flush: working set (BVH+triangles+shading records+textures+path state ~270 MB) exceeds the 126 MB L2",
                       "rays_per_step": total_rays_step},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": 1424, "d2h_bytes_per_step": W * H * 8, "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "k_trace", "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic_bytes(),
                         "algorithmic_bytes_per_ray": B_TRAVERSE, "rays_per_launch": rays_per_launch, "avg_launch_ms": trace_ms / trace_launches,
                         "trace_share_of_step": trace_ms / (trace_ms + shade_ms) if trace_ms + shade_ms > 0 else None,
                         "whole_step": {"bytes_per_ray": B_RAY, "achieved": value * 1e6 / n_gpus * B_RAY / 1e9, "frac": value * 1e6 / n_gpus * B_RAY / 1e9 / peak}},
        }
        if n_gpus == 1 and not args.no_cpu_baseline:
            v, crays, dt, cores, _ = cpu_oracle_run(1, BOUNCES)                 # probe: one sample per pixel
            cspp = max(1, min(64, int(round(15.0 / max(dt, 1e-3)))))            # then ~15 s of CPU work on the same frame
            if cspp > 1:
                v, crays, dt, cores, _ = cpu_oracle_run(cspp, BOUNCES)
            line["cpu_baseline"] = {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "port",
                                    "sample": f"same C2 scene and camera, {W}x{H}, {cspp} spp x {BOUNCES} bounces = {crays} rays in {dt:.1f} s (CPU oracle, {cores} threads)"}
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
