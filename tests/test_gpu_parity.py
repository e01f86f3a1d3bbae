"""GPU parity tests: the CUDA path through the C ABI (librt2cuda.so) against the CPU oracle on the
same seeded inputs.  Bit-exact for hit ids / t / barycentrics and — because sin/cos, the triangle
test, texture filtering and fp16 stores are pinned — bit-exact for K1 radiance too."""
import ctypes as C

import numpy as np
import pytest
import torch

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

pytestmark = pytest.mark.gpu


def _gpu_setup(kind, scale, w, h, seed=0, tex=64):
    sc = rt.Scene(kind, scale, tex, seed)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    return sc, g, ctx, lp


@pytest.mark.parametrize("kind,scale,seed,w,h,brute", [
    (rt.SCENE_CORNELL, 1, 0, 512, 512, True),          # BASELINE config 1 size
    (rt.SCENE_FURNACE, 1, 0, 64, 64, True),
    (rt.SCENE_RANDOM, 1, 3, 256, 256, True),
    (rt.SCENE_HALL, 0.125, 0, 320, 180, True),
    (rt.SCENE_HALL, 0.5, 0, 480, 270, False),
    (rt.SCENE_MANYLIGHT, 0.1, 0, 320, 180, True),
    (rt.SCENE_INSTANCED, 0.05, 0, 320, 180, True),
])
def test_primary_hit_ids_bit_exact(kind, scale, seed, w, h, brute):
    sc, g, ctx, lp = _gpu_setup(kind, scale, w, h, seed)
    lp.trace_primary()
    got = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8)
    ref = O.OracleScene(sc.desc).trace_primary(g, w, h, flags=O.ORC_BRUTE_FORCE if brute else 0)
    st = ctx.stats()
    assert st.rays_traced == w * h
    assert st.rays_hit == int((ref[:, :, 1] != 0xFFFFFFFF).sum())
    assert np.array_equal(got[:, :, :3], ref[:, :, :3]), "instance/geometry/primitive ids differ"
    assert np.array_equal(got, ref), "t / barycentrics differ"
    ctx.close()


@pytest.mark.parametrize("kind,scale,w,h,spp,bounces", [
    (rt.SCENE_CORNELL, 1, 512, 512, 1, 2),       # BASELINE config 1
    (rt.SCENE_CORNELL, 1, 128, 128, 12, 5),      # the reference's hard-coded 12 spp x 5 bounces (launch params zeroed)
    (rt.SCENE_FURNACE, 1, 64, 64, 3, 4),
    (rt.SCENE_RANDOM, 1, 160, 160, 4, 4),
    (rt.SCENE_HALL, 0.125, 320, 180, 4, 4),      # textured (linear + nearest), transforms, emissive strips
    (rt.SCENE_MANYLIGHT, 0.1, 256, 144, 2, 3),
])
def test_reference_path_tracer_bit_exact(kind, scale, w, h, spp, bounces):
    sc, g, ctx, lp = _gpu_setup(kind, scale, w, h)
    params = rt.LaunchParams(spp=0 if spp == 12 else spp, max_bounces=0 if bounces == 5 else bounces)
    lp.execute(1, params)
    d16 = ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4)
    s16 = ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4)
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)
    r16, rs16, r32, rays = O.OracleScene(sc.desc).render_reference(g, 1, w, h, spp=spp, max_bounces=bounces)
    assert ctx.stats().rays_traced == rays
    assert np.array_equal(d32.view(np.uint32), r32.view(np.uint32)), f"fp32 radiance differs in {(d32 != r32).any(axis=2).sum()} pixels"
    assert np.array_equal(d16, r16) and np.array_equal(s16, rs16)
    assert np.isfinite(d32).all() and d32.max() > 0
    ctx.close()


def test_instanced_path_tracer_with_environment_bit_exact():
    """Two-level traversal + environment lookups.  atan2/asin of the equirect mapping follow the transcendental pin
    (binary64, rounded once) on both sides; a last-bit disagreement of the two binary64 libraries can only show where
    the result sits within ~1e-16 of a binary32 rounding boundary, so at most a couple of pixels may differ."""
    w, h = 192, 108
    sc, g, ctx, lp = _gpu_setup(rt.SCENE_INSTANCED, 0.05, w, h)
    g.environment = 1
    lp.update_uniform(g)
    lp.execute(1, rt.LaunchParams(spp=8, max_bounces=3))
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)
    r32, rays = O.OracleScene(sc.desc).render_reference(g, 1, w, h, spp=8, max_bounces=3)[2:4]
    assert ctx.stats().rays_traced == rays
    nbad = int((d32.view(np.uint32) != r32.view(np.uint32)).any(axis=2).sum())
    assert nbad <= 2, f"{nbad} pixels differ"
    assert np.allclose(d32, r32, rtol=1e-5, atol=1e-6)
    ctx.close()


def test_frames_seed_the_rng_and_accumulate_like_the_reference():
    w = h = 96
    sc, g, ctx, lp = _gpu_setup(rt.SCENE_CORNELL, 1, w, h)
    osc = O.OracleScene(sc.desc)
    g.enable_accumulation = 1
    d16 = np.zeros((h, w, 4), np.uint16); s16 = np.zeros((h, w, 4), np.uint16); d32 = np.zeros((h, w, 4), np.float32)
    for frame in range(1, 5):
        g.blend_factor = 1.0 / frame          # the override path, src/main.rs:551-554 (Q13)
        g.frame = frame
        lp.update_uniform(g)
        lp.execute(frame, rt.LaunchParams(spp=2, max_bounces=3))
        osc.render_reference(g, frame, w, h, spp=2, max_bounces=3, diffuse_f16=d16, specular_f16=s16, diffuse_f32=d32)
    assert np.array_equal(ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4), d16)
    assert np.array_equal(ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4), s16)
    assert np.array_equal(ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4).view(np.uint32), d32.view(np.uint32))
    ctx.close()


def test_row_band_sharding_is_bit_identical_to_single_context():
    """RNG is keyed on absolute pixel + frame, so N contexts rendering interleaved 16-row
    bands reproduce the single-context image bit for bit (here both 'ranks' run on cuda:0)."""
    w, h = 200, 150   # neither a multiple of 8/16: ragged tiles and a ragged last band
    sc, g, ctx, lp = _gpu_setup(rt.SCENE_HALL, 0.125, w, h)
    lp.execute(3, rt.LaunchParams(spp=2, max_bounces=3))
    full = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4).copy()
    rays_full = ctx.stats().rays_traced
    merged = np.zeros_like(full)
    rays = 0
    for world in (2, 3):
        merged[:] = 0
        rays = 0
        for rank in range(world):
            ctx.write(rt.BUF_DIFFUSE_F32, np.zeros_like(full))
            lp.execute(3, rt.LaunchParams(spp=2, max_bounces=3, band_rows=16, band_rank=rank, band_world=world))
            part = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)
            rows = [y for y in range(h) if (y // 16) % world == rank]
            other = [y for y in range(h) if (y // 16) % world != rank]
            assert not part[other].any(), "a rank wrote rows it does not own"
            merged[rows] = part[rows]
            rays += ctx.stats().rays_traced
            # rt2_pack_bands: this rank's bands back to back, in frame order, in one call
            send = torch.zeros(len(rows) * w * 4, dtype=torch.float32, device="cuda")
            ctx.pack_bands(rt.BUF_DIFFUSE_F32, send.data_ptr(), send.numel() * 4, 16, rank, world)
            ctx.sync()
            assert np.array_equal(send.cpu().numpy().reshape(len(rows), w, 4).view(np.uint32), part[rows].view(np.uint32))
            with pytest.raises(rt.Rt2Error):
                ctx.pack_bands(rt.BUF_DIFFUSE_F32, send.data_ptr(), send.numel() * 4 - 16, 16, rank, world)
        assert np.array_equal(merged.view(np.uint32), full.view(np.uint32))
        assert rays == rays_full
    ctx.close()


@pytest.mark.parametrize("band_rows,world", [(8, 8), (4, 5), (1, 2)])
def test_pack_bands_matches_band_plan(band_rows, world):
    """The packed send buffer of every rank, scattered with sharding.BandPlan, rebuilds the frame (the NCCL gather's layout)."""
    from raytracer2_b200.sharding import BandPlan
    w, h = 96, 77
    sc, g, ctx, lp = _gpu_setup(rt.SCENE_CORNELL, 1, w, h)
    lp.execute(5, rt.LaunchParams(spp=1, max_bounces=2))
    full = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4).copy()
    plan = BandPlan(h, band_rows, world)
    out = np.zeros_like(full)
    for rank in range(world):
        n = plan.rows(rank)
        send = torch.zeros(max(n, 1) * w * 4, dtype=torch.float32, device="cuda")
        ctx.pack_bands(rt.BUF_DIFFUSE_F32, send.data_ptr(), n * w * 16, band_rows, rank, world)
        ctx.sync()
        packed = send.cpu().numpy().reshape(-1, w, 4)
        for (y0, y1, off) in plan.pack_offsets(rank):
            out[y0:y1] = packed[off:off + (y1 - y0)]
    assert np.array_equal(out.view(np.uint32), full.view(np.uint32))
    ctx.close()


def test_converged_radiance_rel_mse_vs_libm_oracle():
    """north_star: converged radiance within relMSE < 1e-3 of the oracle.  The oracle runs with glibc
    sin/cos here (ORC_TRIG_LIBM), i.e. independent of the pinned polynomial the kernels use."""
    w = h = 48
    sc, g, ctx, lp = _gpu_setup(rt.SCENE_CORNELL, 1, w, h)
    lp.execute(1, rt.LaunchParams(spp=4096, max_bounces=3))
    got = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)[:, :, :3].astype(np.float64)
    ref = O.OracleScene(sc.desc).render_reference(g, 2, w, h, spp=4096, max_bounces=3, flags=O.ORC_TRIG_LIBM)[2][:, :, :3].astype(np.float64)
    # different frame index => independent sample sets: compare the two Monte-Carlo estimates
    rel_mse = np.mean((got - ref) ** 2) / np.mean(ref ** 2)
    assert rel_mse < 5e-3, rel_mse      # two independent 4096-spp estimates: 2x the variance of one
    # same seeds, libm vs pinned trig: paths decorrelate only through last-ulp direction differences
    ref_same = O.OracleScene(sc.desc).render_reference(g, 1, w, h, spp=4096, max_bounces=3, flags=O.ORC_TRIG_LIBM)[2][:, :, :3].astype(np.float64)
    rel_mse_same = np.mean((got - ref_same) ** 2) / np.mean(ref_same ** 2)
    assert rel_mse_same < 1e-3, rel_mse_same
    ctx.close()


def test_error_paths():
    ctx = rt.Renderer(0)
    with pytest.raises(rt.Rt2Error):
        rt.LightPasses(ctx, 64, 64).execute(1)          # no scene / constants yet
    sc = rt.Scene(rt.SCENE_CORNELL)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, 64, 64)
    with pytest.raises(rt.Rt2Error):
        lp.execute(1)                                   # constants missing
    g = rt.default_gconst(sc.camera(1.0), 64, 64, 2)
    rc = ctx._lib.rt2_set_constants(ctx._ctx, C.byref(g), 100)
    assert rc == -1 and b"1424" in ctx._lib.rt2_last_error(ctx._ctx)
    with pytest.raises(rt.Rt2Error):
        ctx.resize(0, 10)
    ctx.close()


def test_empty_scene_misses_everywhere():
    ctx = rt.Renderer(0)
    d = rt._ffi.SceneDesc()
    rt.Model(ctx, C.byref(d))
    cam = rt.Scene(rt.SCENE_CORNELL).camera(1.0)
    g = rt.default_gconst(cam, 32, 32, 0)
    g.refrence_mode = 1
    lp = rt.LightPasses(ctx, 32, 32)
    lp.update_uniform(g)
    lp.trace_primary()
    hits = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(32, 32, 8)
    assert (hits[:, :, :3] == 0xFFFFFFFF).all()
    lp.execute(1, rt.LaunchParams(spp=2, max_bounces=2))
    assert not ctx.read(rt.BUF_DIFFUSE_F32, np.float32).any()
    assert ctx.stats().rays_traced == 32 * 32 * 2
    ctx.close()


def test_full_size_image_is_independent_of_the_schedule(monkeypatch):
    """BASELINE config C2 at full size (262 144 triangles, 1920x1080, 4 spp, 4 bounces): the oracle needs minutes there, so the
    check is a size-independent property — the closest hit is a total order and the RNG is keyed on pixel + frame, hence the
    image must not depend on how rays are scheduled: 1 chain vs the default rule vs 8 chains with capped grids vs another
    trace-kernel variant, and 8 interleaved row-band ranks reassembled with the band plan, all bit-identical."""
    from raytracer2_b200.sharding import BandPlan
    w, h = 1920, 1080
    sc = rt.Scene(rt.SCENE_HALL, 1.0, 512, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    p = rt.LaunchParams(spp=4, max_bounces=4)

    def render(env, bands=None):
        for k in ("RT2_CHAINS", "RT2_RAYS_PER_LANE", "RT2_TRACE_VARIANT"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        lp = rt.LightPasses(ctx, w, h); lp.update_uniform(g)
        if bands is None:
            lp.execute(7, p)
            out = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(h, w, 4)
            rays = ctx.stats().rays_traced
        else:
            plan = BandPlan(h, 8, bands)
            out = np.zeros((h, w, 4), np.uint32); rays = 0
            for rank in range(bands):
                lp.execute(7, rt.LaunchParams(spp=4, max_bounces=4, band_rows=8, band_rank=rank, band_world=bands))
                n = plan.rows(rank)
                send = torch.zeros(n * w * 4, dtype=torch.float32, device="cuda")
                ctx.pack_bands(rt.BUF_DIFFUSE_F32, send.data_ptr(), n * w * 16, 8, rank, bands)
                ctx.sync()
                packed = send.cpu().numpy().view(np.uint32).reshape(n, w, 4)
                for (y0, y1, off) in plan.pack_offsets(rank):
                    out[y0:y1] = packed[off:off + (y1 - y0)]
                rays += ctx.stats().rays_traced
        ctx.close()
        return out, rays

    ref, rays = render({"RT2_CHAINS": "1"})
    assert w * h * 4 * 3 < rays <= w * h * 4 * 4          # closed hall: nearly every path runs all 4 bounces
    f = ref.view(np.float32)[:, :, :3]
    assert np.isfinite(f).all() and (f >= 0).all() and f.mean() > 0.01
    for env in ({}, {"RT2_CHAINS": "8", "RT2_RAYS_PER_LANE": "3"}, {"RT2_TRACE_VARIANT": "0"}):
        got, r = render(env)
        assert r == rays and np.array_equal(got, ref), env
    got, r = render({}, bands=8)
    assert r == rays and np.array_equal(got, ref)


def test_read_async_two_frames_in_flight_matches_blocking_read():
    """rt2_read_async (device snapshot + copy stream) returns what rt2_read returns, with two reads in flight while the
    next frames render, and rt2_set_constants no longer waits for the stream (pinned staging ring)."""
    import torch
    w, h = 96, 64
    sc = rt.Scene(rt.SCENE_CORNELL)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    n = w * h * 8
    bufs = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(4)]
    want = []
    for f in range(4):
        g.frame = 1 + f
        lp.update_uniform(g)
        lp.execute(1 + f, rt.LaunchParams(spp=1 + f, max_bounces=2))
        want.append(ctx.read(rt.BUF_DIFFUSE_F16, np.uint8).copy())
    for f in range(4):                       # same frames again, never syncing between them
        g.frame = 1 + f
        lp.update_uniform(g)
        lp.execute(1 + f, rt.LaunchParams(spp=1 + f, max_bounces=2))
        ctx.read_into_async(rt.BUF_DIFFUSE_F16, bufs[f].data_ptr(), n)
    ctx.read_fence()
    ctx.sync()
    for f in range(4):
        assert np.array_equal(bufs[f].numpy(), want[f]), f"frame {f}"
    assert not np.array_equal(want[0], want[3])
    with pytest.raises(rt.Rt2Error):
        ctx.read_into_async(rt.BUF_DIFFUSE_F16, bufs[0].data_ptr(), n - 8)
    ctx.close()


def test_constants_ring_keeps_every_frames_gconst_without_syncing():
    """rt2_set_constants stages GConst in a ring of pinned slots and never waits for the stream: six back-to-back frames with
    six different cameras (more than the ring holds) must each render with their own constants."""
    import torch
    w, h = 96, 64
    sc = rt.Scene(rt.SCENE_CORNELL)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    n = w * h * 8

    def gconst(k):
        cam = sc.camera(w / h)
        cam.position[0] = 0.15 * k - 0.3
        cam.position[1] = 0.05 * k
        g = rt.default_gconst(cam, w, h, sc.num_lights)
        g.refrence_mode = 1
        return g

    want = []
    for k in range(6):
        lp.update_uniform(gconst(k)); lp.execute(5, rt.LaunchParams(spp=1, max_bounces=2))
        want.append(ctx.read(rt.BUF_DIFFUSE_F16, np.uint8).copy())
    assert len({w_.tobytes() for w_ in want}) == 6
    bufs = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(6)]
    for k in range(6):
        lp.update_uniform(gconst(k)); lp.execute(5, rt.LaunchParams(spp=1, max_bounces=2))
        ctx.read_into_async(rt.BUF_DIFFUSE_F16, bufs[k].data_ptr(), n)
    ctx.sync()
    for k in range(6):
        assert np.array_equal(bufs[k].numpy(), want[k]), f"frame {k}"
    ctx.close()
