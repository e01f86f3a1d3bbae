"""Primary-hit reuse in K1 (rt2_render.cu, k1_shade_step / k_spec_gen): setupPrimaryRay has no jitter (GBufferHelpers.glsl:12-27) and
refrence.rgen:26 calls it anew for every sample, so the spp samples of a pixel issue the same first trace().  The product walks it once
and lets the later samples consume the kept hit record.  That must change nothing the shader can see: the fp32 radiance, the rgba16f
stores, rays_traced (the shader's trace() calls, equal to the oracle's count) and rays_hit are the same with RT2_PRIMARY_REUSE=0 and 1
under every K1 schedule; only rt2_stats.rays_walked drops — by exactly (spp - 1) walks per pixel where no speculation is involved."""
import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

pytestmark = pytest.mark.gpu

SCHEDULES = {"wavefront": {"RT2_SPECULATE": "0"}, "speculative": {"RT2_SPECULATE": "1"}, "persistent": {"RT2_SPECULATE": "0", "RT2_PATH_MODE": "1"},
             "wavefront-1-chain": {"RT2_SPECULATE": "0", "RT2_CHAINS": "1"}}


def _frame(monkeypatch, knobs, reuse, sc, g, w, h, spp, bounces, bands=None):
    for k in ("RT2_SPECULATE", "RT2_PATH_MODE", "RT2_CHAINS"):
        monkeypatch.delenv(k, raising=False)
    for k, v in knobs.items():
        monkeypatch.setenv(k, v)
    monkeypatch.setenv("RT2_PRIMARY_REUSE", "1" if reuse else "0")
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    p = rt.LaunchParams(spp=spp, max_bounces=bounces) if bands is None else rt.LaunchParams(spp=spp, max_bounces=bounces, band_rows=bands[0], band_rank=bands[1], band_world=bands[2])
    lp.execute(3, p)
    st = ctx.stats()
    out = (ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4).copy(), ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).copy(), st.rays_traced, st.rays_hit, st.rays_walked,
           st.wavefront_iterations)
    ctx.close()
    return out


@pytest.mark.parametrize("schedule", list(SCHEDULES))
@pytest.mark.parametrize("kind,scale,w,h,spp,bounces,env", [
    (rt.SCENE_HALL, 0.125, 200, 90, 4, 4, 0),           # closed: every sample runs its full length
    (rt.SCENE_CORNELL, 1, 96, 96, 5, 3, 0),             # open front: samples end early, sky pixels finish in one shade step
    (rt.SCENE_INSTANCED, 0.05, 160, 90, 3, 4, 1),       # flattened instancing + sky: the kept record carries the instance id
    (rt.SCENE_CORNELL, 1, 64, 64, 4, 1, 0),             # one bounce: a sample is its primary ray alone -> one walk per pixel
])
def test_primary_hit_reuse_changes_nothing_but_the_walks(monkeypatch, schedule, kind, scale, w, h, spp, bounces, env):
    sc = rt.Scene(kind, scale, 32, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1; g.environment = env
    off = _frame(monkeypatch, SCHEDULES[schedule], False, sc, g, w, h, spp, bounces)
    on = _frame(monkeypatch, SCHEDULES[schedule], True, sc, g, w, h, spp, bounces)
    assert np.array_equal(on[0].view(np.uint32), off[0].view(np.uint32)) and np.array_equal(on[1], off[1])
    assert on[2] == off[2] and on[3] == off[3]
    rays = O.OracleScene(sc.desc).render_reference(g, 3, w, h, spp=spp, max_bounces=bounces)[3]
    assert on[2] == rays, "rays_traced is the shader's trace() count whatever is walked"
    if schedule != "speculative":
        assert off[4] == off[2], "without reuse every trace() is a walk"
        assert on[4] == on[2] - w * h * (spp - 1), "one primary walk per pixel instead of spp"
        if schedule != "persistent":
            assert off[5] == spp * bounces and on[5] == 1 + spp * (bounces - 1)      # the dependent chain gets shorter too
    else:
        assert on[4] < off[4]


def test_primary_hit_reuse_under_row_bands(monkeypatch):
    kind, w, h, spp, bounces = rt.SCENE_HALL, 96, 70, 3, 3
    sc = rt.Scene(kind, 0.125, 32, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    whole = _frame(monkeypatch, SCHEDULES["wavefront"], True, sc, g, w, h, spp, bounces)
    rays = 0
    for rank in range(3):
        for name in ("wavefront", "speculative"):
            part = _frame(monkeypatch, SCHEDULES[name], True, sc, g, w, h, spp, bounces, bands=(8, rank, 3))
            rows = [y for y in range(h) if (y // 8) % 3 == rank]
            assert np.array_equal(part[0][rows].view(np.uint32), whole[0][rows].view(np.uint32)), (rank, name)
        rays += part[2]
    assert rays == whole[2]


@pytest.mark.parametrize("kind,scale,sky", [(rt.SCENE_MANYLIGHT, 0.1, 0), (rt.SCENE_CORNELL, 1.0, 0), (rt.SCENE_HALL, 0.125, 0)])
def test_k9_final_visibility_is_answered_by_a_clear_initial_ray(monkeypatch, kind, scale, sky):
    """K9 with both visibility tests on (di_fused_resampling.rgen:40-48, ShadingHelpers.glsl:33-38): the final ray has the initial ray's origin
    and direction and a t-interval inside the initial one, so a clear initial ray answers it (k_di_stage_c).  Reservoirs, both lighting images
    and the trace() count equal the every-ray-walked run's and the oracle's; rays_walked drops by the final rays that were not walked."""
    w, h, frame = 192, 108, 2
    out = {}
    for reuse in (0, 1):
        monkeypatch.setenv("RT2_VISIBILITY_REUSE", str(reuse))
        sc = rt.Scene(kind, scale, 32, 0)
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
        g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
        isp = g.restir_di.initial_sampling_params
        isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1; isp.enable_initial_visibility = 1
        g.restir_di.shading_params.enable_final_visibility = 1
        lp = rt.LightPasses(ctx, w, h)
        rt.PrepareLightsTasks(ctx).execute()
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
        lp.update_uniform(g)
        lp.execute(frame)
        st = ctx.stats()
        out[reuse] = (ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).copy(), ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).copy(), ctx.read(rt.BUF_SPECULAR_F16, np.uint16).copy(),
                      st.rays_traced, st.rays_hit, st.rays_walked)
        ctx.close()
    off, on = out[0], out[1]
    for k in range(3):
        assert np.array_equal(on[k], off[k])
    assert on[3] == off[3] and on[4] == off[4]
    assert off[5] == off[3], "without reuse every trace() is a walk"
    valid = int((on[0].reshape(-1, 6)[:, 0] != 0).sum())           # reservoirs that kept a light: their initial ray was clear
    assert 0.5 * valid <= on[3] - on[5] <= valid and valid > 0, (on[3], on[5], valid)


@pytest.mark.parametrize("kind,scale", [(rt.SCENE_MANYLIGHT, 0.1), (rt.SCENE_INSTANCED, 0.05), (rt.SCENE_HALL, 0.125)])
def test_visibility_ray_batches_end_at_their_first_hit(monkeypatch, kind, scale):
    """The visibility-ray batches of K9 / K5 (consumers read `p.geometryIndex == ~0u` only: bridge :218-231, ShadingHelpers.glsl:35-37) end a walk at
    its first hit (trace_dyn_body<ANYHIT>).  Nothing the passes write may depend on it: RT2_ANY_HIT=0 (closest hit always) gives the same reservoirs,
    images, ray and hit counts — with the visibility reuse off, so that every visibility ray is really walked."""
    w, h = 192, 108
    monkeypatch.setenv("RT2_VISIBILITY_REUSE", "0")
    out = {}
    for any_hit in (0, 1):
        monkeypatch.setenv("RT2_ANY_HIT", str(any_hit))
        sc = rt.Scene(kind, scale, 32, 0)
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        d = sc.desc.contents
        g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
        g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 1; g.environment = 1 if d.sky_width else 0
        isp = g.restir_di.initial_sampling_params
        isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1; isp.enable_initial_visibility = 1
        g.restir_di.shading_params.enable_final_visibility = 1
        lp = rt.LightPasses(ctx, w, h)
        rt.PrepareLightsTasks(ctx).execute()
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
        frames = []
        for frame in (1, 2):
            lp.update_uniform(g)
            lp.execute(frame)
            st = ctx.stats()
            frames.append((ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).copy(), ctx.read(rt.BUF_GI_RESERVOIRS, np.uint32).copy(), ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).copy(),
                           ctx.read(rt.BUF_SPECULAR_F16, np.uint16).copy(), st.rays_traced, st.rays_hit))
        out[any_hit] = frames
        ctx.close()
    for a, b in zip(out[0], out[1]):
        for k in range(4):
            assert np.array_equal(a[k], b[k]), k
        assert a[4] == b[4] and a[5] == b[5]
        assert a[5] > 0


@pytest.mark.parametrize("kind,scale,env", [(rt.SCENE_CORNELL, 1, 0), (rt.SCENE_HALL, 0.125, 0), (rt.SCENE_INSTANCED, 0.05, 1)])
def test_sparse_contributions_change_nothing(monkeypatch, kind, scale, env):
    """Speculative schedule: a per-bounce contribution that is +0 in every component is neither stored nor added (k_spec_shade / k_spec_reduce).
    x + (+0) == x bit for bit, so the image must equal the one with every contribution stored (RT2_SPARSE_CONTRIB=0) and the oracle's."""
    w, h, spp, bounces = 160, 96, 5, 4
    sc = rt.Scene(kind, scale, 32, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1; g.environment = env
    out = []
    for sparse in ("0", "1"):
        monkeypatch.setenv("RT2_SPARSE_CONTRIB", sparse)
        out.append(_frame(monkeypatch, SCHEDULES["speculative"], True, sc, g, w, h, spp, bounces))
    assert np.array_equal(out[0][0].view(np.uint32), out[1][0].view(np.uint32)) and np.array_equal(out[0][1], out[1][1])
    assert out[0][2] == out[1][2] and out[0][3] == out[1][3]
    ref = O.OracleScene(sc.desc).render_reference(g, 3, w, h, spp=spp, max_bounces=bounces)
    nbad = int((out[1][0].view(np.uint32) != ref[2].view(np.uint32)).any(axis=2).sum())
    assert nbad <= (2 if env else 0) and out[1][2] == ref[3]
