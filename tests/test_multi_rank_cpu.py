"""N>1 host logic on CPU: the row-band plan (raytracer2_b200/sharding.py) and the band gather, exercised by
two real processes over `gloo` (127.0.0.1).  Each rank renders only its own bands with the CPU oracle (the
checker; the product's kernels apply the same (y / band_rows) % world rule, tests/test_gpu_parity.py covers
that on the device), packs them, all-gathers, and rank 0 reassembles — the result must be bit-identical to the
whole-frame render, which is the property SURVEY §8e asks of the sharded path."""
import os
import socket
import sys

import numpy as np
import pytest

from raytracer2_b200.sharding import BandPlan, pack_bands, unpack_gathered

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_band_plan_partitions_rows():
    for (h, br, world) in [(1080, 16, 1), (1080, 16, 2), (1080, 16, 8), (1080, 16, 3), (37, 5, 4), (7, 16, 2), (16, 16, 8)]:
        p = BandPlan(h, br, world)
        owner = np.full(h, -1)
        for r in range(world):
            for (y0, y1) in p.bands(r):
                assert 0 <= y0 < y1 <= h and y0 % br == 0 and y1 - y0 <= br
                assert (owner[y0:y1] == -1).all()
                owner[y0:y1] = r
            assert p.rows(r) == (owner == r).sum() <= p.max_rows
        assert (owner >= 0).all()
        assert all(owner[y] == p.owner(y) for y in range(h))
        # interleaving keeps the ranks within one band of each other
        assert max(p.rows(r) for r in range(world)) - min(p.rows(r) for r in range(world)) <= br
    p = BandPlan(1080, 16, 8)                       # the 8-GPU bench layout: 68 bands (last one 8 rows)
    assert p.num_bands == 68 and p.max_rows == 9 * 16 and p.rows(3) == 8 * 16 + 8 and p.rows(7) == 8 * 16
    with pytest.raises(ValueError):
        BandPlan(0, 16, 2)
    with pytest.raises(ValueError):
        p.bands(8)


def test_pack_unpack_round_trip_single_process():
    rng = np.random.default_rng(3)
    for (h, br, world) in [(37, 5, 4), (1080, 16, 8), (9, 4, 3)]:
        p = BandPlan(h, br, world)
        frame = rng.integers(0, 2**32, size=(h, 6, 4), dtype=np.uint32)
        gathered = np.full((world * p.max_rows, 6, 4), 0xDEADBEEF, np.uint32)
        for r in range(world):
            pack_bands(p, r, frame, gathered[r * p.max_rows:(r + 1) * p.max_rows])
        out = np.zeros_like(frame)
        unpack_gathered(p, gathered, out)
        assert np.array_equal(out, frame)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _rank_main(rank, world, port, w, h, band_rows, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import raytracer2_b200 as rt
    from raytracer2_b200.sharding import BandPlan, gather_frame, pack_bands, unpack_gathered
    from oracle import oracle_ffi as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 16, 0)                 # scene replicated: every rank builds its own copy
        g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
        g.refrence_mode = 1
        osc = O.OracleScene(sc.desc)
        plan = BandPlan(h, band_rows, world)
        d32 = np.zeros((h, w, 4), np.float32)
        rays = 0
        for (y0, y1) in plan.bands(rank):                           # this rank's rows only
            rays += osc.render_reference(g, 7, w, h, spp=2, max_bounces=3, row0=y0, row1=y1, diffuse_f32=d32, threads=2)[3]
        mine = np.zeros(h, bool)
        for (y0, y1) in plan.bands(rank):
            mine[y0:y1] = True
        assert (d32[~mine] == 0).all()                              # nothing rendered outside the rank's bands
        send = torch.zeros(plan.max_rows, w, 4)
        pack_bands(plan, rank, torch.from_numpy(d32), send)
        recv = torch.full((world * plan.max_rows, w, 4), float("nan"))
        gather_frame(plan, rank, send, recv, dist)
        recv2 = torch.full((world * plan.max_rows, w, 4), float("nan"))      # the double-buffered form bench.py uses: handle, wait later
        handle = gather_frame(plan, rank, send, recv2, dist, async_op=True)
        handle.wait()
        assert torch.equal(recv2.view(torch.int32), recv.view(torch.int32))
        t = torch.tensor([float(rays)], dtype=torch.float64)
        dist.all_reduce(t)                                          # whole-job ray count, as bench.py sums it
        tm = torch.tensor([float(rank + 1)]); dist.all_reduce(tm, op=dist.ReduceOp.MAX)   # max-over-ranks timing path
        assert tm.item() == world
        if rank == 0:
            frame = torch.zeros(h, w, 4)
            unpack_gathered(plan, recv, frame)
            whole, wrays = osc.render_reference(g, 7, w, h, spp=2, max_bounces=3, threads=2)[2:4]
            q.put((np.array_equal(frame.numpy().view(np.uint32), whole.view(np.uint32)), int(t.item()), int(wrays)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("h,band_rows", [(40, 16), (33, 4)])
def test_two_ranks_over_gloo_reassemble_the_whole_frame_bit_exactly(h, band_rows):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, 48, h, band_rows, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        if p.is_alive():
            p.kill()
        assert p.exitcode == 0
    same, rays_sum, rays_whole = q.get(timeout=5)
    assert same
    assert rays_sum == rays_whole > 0           # the bands partition the work: no ray traced twice, none dropped


def _exchange_main(rank, world, port, w, h, band_rows, q):
    """Two ranks each hold a block-linear reservoir page and a row-major G-buffer plane that are valid only at their own pixels;
    after the exchange (pack own -> all_gather -> scatter the others') both hold the complete buffers."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from raytracer2_b200.sharding import BandPlan, exchange_entry_index, exchange_pack, exchange_unpack
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = BandPlan(h, band_rows, world)
        pitch = ((w + 15) // 16) * 256                               # calculate_reservoir_buffer_parameters: blockRowPitch
        entries = pitch * ((h + 15) // 16)
        rng = np.random.default_rng(5)                               # same seed on every rank: the "whole frame" truth
        truth = {"page": rng.integers(0, 2**32, (entries, 8), dtype=np.uint32), "plane": rng.integers(0, 2**32, (w * h, 1), dtype=np.uint32)}
        owner = (np.arange(h) // band_rows) % world
        ok = True
        for name, layout in (("page", 1), ("plane", 0)):
            idx = exchange_entry_index(w, h, layout, pitch)
            mine = idx[owner == rank].ravel()
            buf = np.full_like(truth[name], 0xDEADBEEF)
            buf[mine] = truth[name][mine]                            # only this rank's pixels are valid
            send = torch.from_numpy(exchange_pack(plan, rank, buf, w, layout, pitch).view(np.int32).copy())
            chunks = [torch.zeros_like(send) for _ in range(world)]
            dist.all_gather(chunks, send)
            gathered = np.stack([c.numpy().view(np.uint32) for c in chunks])
            exchange_unpack(plan, rank, gathered, buf, w, layout, pitch)
            pix = idx.ravel()                                        # entries that belong to a pixel (block padding stays poisoned)
            ok = ok and np.array_equal(buf[pix], truth[name][pix])
            pad = np.setdiff1d(np.arange(len(buf)), pix)
            ok = ok and (buf[pad] == 0xDEADBEEF).all()
        q.put((rank, bool(ok)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("w,h,band_rows", [(48, 40, 8), (37, 50, 8), (64, 33, 4)])
def test_two_ranks_over_gloo_exchange_reservoir_pages_and_gbuffer_planes(w, h, band_rows):
    """The K6 / K7 input exchange of rt2_group_render (rt2_group.cu), host-side mirror over gloo: block-linear and row-major addressing,
    8-row bands cutting the 16 x 16 reservoir blocks in half, ragged last bands."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_exchange_main, args=(r, 2, port, w, h, band_rows, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        if p.is_alive():
            p.kill()
        assert p.exitcode == 0
    got = sorted(q.get(timeout=5) for _ in range(2))
    assert got == [(0, True), (1, True)]
