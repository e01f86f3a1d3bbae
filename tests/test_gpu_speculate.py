"""Speculative sample parallelism (rt2_render.cu, render_reference_spec): every sample of a pixel traced as a path of its own from
a speculated RNG position, validated and summed in the shader's order afterwards.  The image, the rgba16f stores and the ray
count must be bit-identical to the oracle's sequential loop (refrence.rgen:18-47) — on closed scenes (speculation always right),
on open ones (most pixels need fix-up rounds and the sequential fallback), with 0 / 1 / 5 prediction rounds before the candidate round, under row-band
sharding and with accumulation."""
import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

pytestmark = pytest.mark.gpu


def _setup(monkeypatch, kind, scale, w, h, speculate="1", rounds=None, env=0, tex=32):
    monkeypatch.setenv("RT2_SPECULATE", speculate)
    if rounds is None:
        monkeypatch.delenv("RT2_SPEC_ROUNDS", raising=False)
    else:
        monkeypatch.setenv("RT2_SPEC_ROUNDS", str(rounds))
    sc = rt.Scene(kind, scale, tex, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1; g.environment = env
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    return sc, g, ctx, lp


@pytest.mark.parametrize("kind,scale,w,h,spp,bounces,env,rounds", [
    (rt.SCENE_FURNACE, 1, 64, 64, 3, 4, 0, None),          # closed box: speculation (almost) never fails
    (rt.SCENE_HALL, 0.125, 320, 180, 4, 4, 0, None),       # the bench scene's shape: textured, closed hall
    (rt.SCENE_HALL, 0.125, 201, 77, 8, 3, 0, 1),           # ragged tiles
    (rt.SCENE_CORNELL, 1, 128, 128, 12, 5, 0, None),       # open front: many samples end on a miss -> fix-up rounds + fallback
    (rt.SCENE_CORNELL, 1, 128, 128, 12, 5, 0, 0),          # no speculative fix-up: everything pending goes through the fallback
    (rt.SCENE_CORNELL, 1, 96, 96, 6, 2, 0, 5),             # more fix-up rounds than a pixel can need (spp - 1)
    (rt.SCENE_RANDOM, 1, 160, 160, 4, 4, 0, None),
    (rt.SCENE_INSTANCED, 0.05, 192, 108, 8, 3, 1, None),   # flattened instancing + sky: misses carry radiance
    (rt.SCENE_MANYLIGHT, 0.1, 256, 144, 2, 3, 0, None),
])
def test_speculative_samples_bit_exact(monkeypatch, kind, scale, w, h, spp, bounces, env, rounds):
    sc, g, ctx, lp = _setup(monkeypatch, kind, scale, w, h, rounds=rounds, env=env)
    lp.execute(1, rt.LaunchParams(spp=spp, max_bounces=bounces))
    st = ctx.stats()
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)
    d16 = ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4)
    s16 = ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4)
    r16, rs16, r32, rays = O.OracleScene(sc.desc).render_reference(g, 1, w, h, spp=spp, max_bounces=bounces)
    n_rounds = 1 + (1 if rounds is None else rounds) + 1           # round 0, prediction fix-ups, candidate round
    assert st.kernel_launches == n_rounds * (2 + 2 * bounces) + 1, "the speculative schedule did not run"
    assert st.rays_traced == rays, "only valid samples may be counted"
    nbad = int((d32.view(np.uint32) != r32.view(np.uint32)).any(axis=2).sum())
    assert nbad <= (2 if env else 0), f"{nbad} pixels differ"          # sky lookups: the transcendental pin (see test_gpu_parity)
    if not env:
        assert np.array_equal(d16, r16) and np.array_equal(s16, rs16)
    assert np.isfinite(d32).all() and d32.max() > 0
    ctx.close()


@pytest.mark.parametrize("chains", [2, 3, 8])
def test_speculative_wavefront_chains_bit_exact(monkeypatch, chains):
    """The frame cut into wavefront chains at tile-row boundaries (own streams, queues, pending lists, candidate regions)."""
    monkeypatch.setenv("RT2_CHAINS", str(chains))
    w, h, spp, bounces = 201, 77, 5, 3
    sc, g, ctx, lp = _setup(monkeypatch, rt.SCENE_CORNELL if chains == 3 else rt.SCENE_HALL, 1.0 if chains == 3 else 0.125, w, h)
    lp.execute(2, rt.LaunchParams(spp=spp, max_bounces=bounces))
    st = ctx.stats()
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(h, w, 4)
    _, _, r32, rays = O.OracleScene(sc.desc).render_reference(g, 2, w, h, spp=spp, max_bounces=bounces)
    assert st.kernel_launches == chains * (3 * (2 + 2 * bounces) + 1)
    assert st.rays_traced == rays
    assert np.array_equal(d32, r32.view(np.uint32))
    ctx.close()


def test_speculative_bands_and_accumulation_bit_exact(monkeypatch):
    """Three accumulated frames, each rendered as 3 interleaved band ranks on one context: equals the oracle's whole frames."""
    w, h = 200, 150
    sc, g, ctx, lp = _setup(monkeypatch, rt.SCENE_HALL, 0.125, w, h)
    osc = O.OracleScene(sc.desc)
    g.enable_accumulation = 1
    d16 = np.zeros((h, w, 4), np.uint16); s16 = np.zeros((h, w, 4), np.uint16); d32 = np.zeros((h, w, 4), np.float32)
    rays_gpu = rays_ref = 0
    for frame in range(1, 4):
        g.blend_factor = 1.0 / frame
        g.frame = frame
        lp.update_uniform(g)
        for rank in range(3):
            lp.execute(frame, rt.LaunchParams(spp=4, max_bounces=3, band_rows=16, band_rank=rank, band_world=3))
            rays_gpu += ctx.stats().rays_traced
        rays_ref += osc.render_reference(g, frame, w, h, spp=4, max_bounces=3, diffuse_f16=d16, specular_f16=s16, diffuse_f32=d32)[3]
    assert rays_gpu == rays_ref
    assert np.array_equal(ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4).view(np.uint32), d32.view(np.uint32))
    assert np.array_equal(ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4), d16)
    ctx.close()


def test_auto_policy_speculates_in_closed_scenes_only(monkeypatch):
    """RT2_SPECULATE unset: the first frame runs sequentially, its (rays, hits) are read back without a sync, and later frames
    speculate if fewer than 2 % of the samples ended on a miss — in the closed hall, not in the open Cornell box."""
    monkeypatch.delenv("RT2_SPECULATE", raising=False)
    for kind, scale, closed in ((rt.SCENE_HALL, 0.125, True), (rt.SCENE_CORNELL, 1.0, False)):
        w, h = 160, 90
        sc = rt.Scene(kind, scale, 32, 0)
        g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
        g.refrence_mode = 1
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        lp = rt.LightPasses(ctx, w, h)
        lp.update_uniform(g)
        osc = O.OracleScene(sc.desc)
        launches = []
        for frame in range(1, 5):
            lp.execute(frame, rt.LaunchParams(spp=4, max_bounces=4))
            ctx.sync()
            launches.append((ctx.stats().kernel_launches, ctx.stats().sample_schedule))
            got = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(h, w, 4)
            ref = osc.render_reference(g, frame, w, h, spp=4, max_bounces=4)[2].view(np.uint32)
            assert np.array_equal(got, ref), (kind, frame)
        # sequential: per wavefront chain a launch pair per walked (sample, bounce) — the primary ray once — and the init
        assert launches[0][1] == 0 and launches[0][0] % (2 * (1 + 4 * 3) + 1) == 0, launches
        if closed:
            assert launches[-1][1] == 1, launches         # speculation took over
        else:
            assert launches[-1] == launches[0], launches
        ctx.close()
