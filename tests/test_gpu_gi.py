"""GPU parity of the default ReSTIR-GI chain: K4 brdf_rays -> K5 shade_secondary_surfaces -> K8 gi_final_shading
through the C ABI (rt2_render with refrence_mode 0, enable_restir_gi 1) against the CPU oracle on the same G-buffer.
Compared bit for bit: the 48-byte secondary G-buffer, the packed 32-byte GI reservoirs (LogLuv radiance, snorm16
octahedral normal) and both rgba16f lighting images.  Tolerance: log2/pow are evaluated in binary64 and rounded once
on both sides, which can differ in the last bit for ~1e-5 of the inputs; such a pixel may differ, nothing else may."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

pytestmark = pytest.mark.gpu

GB_KEYS = [("depth", rt.BUF_GBUFFER_DEPTH, np.float32, 1), ("normals", rt.BUF_GBUFFER_NORMALS, np.uint32, 1),
           ("geo_normals", rt.BUF_GBUFFER_GEO_NORMALS, np.uint32, 1), ("diffuse_albedo", rt.BUF_GBUFFER_DIFFUSE_ALBEDO, np.uint32, 1),
           ("specular_rough", rt.BUF_GBUFFER_SPECULAR_ROUGH, np.uint32, 1)]


def _ctx(kind, scale, w, h, tex=64, sky=0):
    sc = rt.Scene(kind, scale, tex, 0)
    if sky:
        sc.add_sky(sky)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    d = sc.desc.contents
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
    g.refrence_mode = 0
    return sc, ctx, g, rt.LightPasses(ctx, w, h)


def _read_gb(ctx, w, h):
    return {k: ctx.read(b, dt).reshape(h, w) for k, b, dt, n in GB_KEYS}


def _light_inputs(sc, ctx):
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    mips = [ctx.read(rt.BUF_LOCAL_PDF_MIP0 + k, np.float32) for k in range(pm.value)]
    return {"lights": ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12), "geom_to_light": ctx.read(rt.BUF_GEOM_TO_LIGHT, np.uint32),
            "pdf_packed": np.concatenate([m.ravel() for m in mips]), "pdf_dims": (pw.value, ph.value, pm.value)}


def _run(kind, scale, w, h, frames=(1,), di=0, final_vis=0, mis=1, env=0, additive=1, oflags=O.ORC_BRUTE_FORCE, local=2, brdf=1, sky=0, presampling=False,
         env_samples=0):
    """sky: attach the procedural sky box of that width.  presampling: K11 every frame + RT2_LAUNCH_PRESAMPLING / ORC_PRESAMPLING — the sampler's
    RTXDI_ENABLE_PRESAMPLING 1 build, in which the environment light K10 writes at lights + 1 exists (K9 draws env_samples candidates from its tiles,
    K5 finds it through BRDF rays that miss)."""
    sc, ctx, g, lp = _ctx(kind, scale, w, h, sky=sky)
    g.restir_di.initial_sampling_params.num_primary_environment_samples = env_samples
    g.restir_di.initial_sampling_params.environment_map_importance_sampling = 1 if presampling else 0
    g.enable_restir_gi = 1; g.enable_restir_di = di; g.environment = env; g.enable_brdf_additive_blend = additive
    g.restir_gi.final_shading_params.enable_final_visibility = final_vis
    g.restir_gi.final_shading_params.enable_final_mis = mis
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = local
    g.restir_di.initial_sampling_params.num_primary_brdf_samples = brdf
    g.restir_di.shading_params.enable_final_visibility = 1
    if di:
        rt.PrepareLightsTasks(ctx).execute()
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    env_pdf, env_dims = None, (0, 0, 0)
    if presampling:
        oflags |= O.ORC_PRESAMPLING
        if sc.desc.contents.sky_width:
            import math
            rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
            sw, sh = sc.desc.contents.sky_width, sc.desc.contents.sky_height
            env_dims = (sw, sh, int(math.ceil(math.log2(max(sw, sh)))) + 1)
            env_pdf = np.concatenate([ctx.read(rt.BUF_ENV_PDF_MIP0 + k, np.uint16).ravel() for k in range(env_dims[2])])
    osc = O.OracleScene(sc.desc)
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    state = None
    out = []
    for frame in frames:
        g.frame = frame
        lp.update_uniform(g)
        if presampling:
            lp.execute_presampeling(frame, env_pdf is not None)
        lp.execute(frame, rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING) if presampling else None)
        gb = _read_gb(ctx, w, h)
        got = {"secondary": ctx.read(rt.BUF_SECONDARY_GBUFFER, np.uint32).reshape(-1, 12), "gi_reservoirs": ctx.read(rt.BUF_GI_RESERVOIRS, np.uint32).reshape(-1, 8),
               "diffuse": ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4), "specular": ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4)}
        st = ctx.stats()
        li = _light_inputs(sc, ctx) if di else None
        if di and presampling:
            li.update(ris=ctx.read(rt.BUF_RIS, np.uint32).reshape(-1, 2), ris_light_data=ctx.read(rt.BUF_RIS_LIGHT_DATA, np.uint32).reshape(-1, 8),
                      env_pdf=env_pdf, env_dims=env_dims)
        if state is None:
            state = {"secondary": np.zeros((pitch, 12), np.uint32), "gi_reservoirs": np.zeros((pitch * 2, 8), np.uint32),
                     "diffuse": np.zeros((h, w, 4), np.uint16), "specular": np.zeros((h, w, 4), np.uint16)}
        rays_di = 0
        if di:
            _, _, _, rays_di = osc.di_fused_resampling(g, frame, w, h, gb, li["lights"], li["geom_to_light"], li["pdf_packed"], *li["pdf_dims"],
                                                       g.restir_di.reservoir_buffer_params.reservoir_array_pitch,
                                                       diffuse_f16=state["diffuse"], specular_f16=state["specular"], flags=oflags,
                                                       ris=li.get("ris"), ris_light_data=li.get("ris_light_data"), env_pdf=li.get("env_pdf"), env_dims=li.get("env_dims", (0, 0, 0)))
        state = osc.gi_chain(g, frame, w, h, gb, pitch, light_inputs=li, state=state, flags=oflags)
        out.append((got, {k: v.copy() for k, v in state.items() if k != "rays"}, st.rays_traced - w * h, rays_di + sum(state["rays"]), gb))
    ctx.close()
    return out


def _check(got, ref, w, h, tol):
    for key in ("secondary", "gi_reservoirs"):
        assert got[key].shape == ref[key].shape
        bad = np.flatnonzero((got[key] != ref[key]).any(axis=1))
        assert len(bad) <= tol, f"{key}: {len(bad)} records differ, first {bad[:4]}: {got[key][bad[:1]]} vs {ref[key][bad[:1]]}"
    for key in ("diffuse", "specular"):
        nbad = int((got[key] != ref[key]).any(axis=2).sum())
        assert nbad <= tol, f"{key}: {nbad} pixels differ"


@pytest.mark.parametrize("kind,scale,w,h,kw", [
    (rt.SCENE_CORNELL, 1.0, 512, 512, dict()),                                         # the reference's default frame (src/main.rs:388-399) at config 1 size
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(mis=0, final_vis=1, additive=0)),
    (rt.SCENE_HALL, 0.125, 320, 180, dict(final_vis=1, oflags=0)),
    (rt.SCENE_INSTANCED, 0.05, 320, 180, dict(env=1, oflags=0)),                       # open scene under a sky: environment secondary surfaces
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict(di=1)),                                     # DI on: K5 samples the light list at the secondary hit
    (rt.SCENE_MANYLIGHT, 0.2, 320, 180, dict(di=1, final_vis=1, oflags=0)),
    # DI + GI under a sky with the environment light (presampling build): K9 draws environment candidates, K5's BRDF rays that miss select it
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(di=1, env=1, sky=128, presampling=True, env_samples=2)),
    (rt.SCENE_INSTANCED, 0.05, 320, 180, dict(di=1, env=1, oflags=0, presampling=True, env_samples=4, local=0)),
])
def test_gi_chain_bit_exact(kind, scale, w, h, kw):
    (got, ref, rays, rrays, gb), = _run(kind, scale, w, h, **kw)
    tol = max(2, w * h // 50000)
    _check(got, ref, w, h, tol)
    assert rays == rrays
    hit = gb["depth"] != 100000.0
    valid = (ref["gi_reservoirs"][:, 3] & 0xff) != 0
    assert hit.any() and valid.any()
    d = got["diffuse"].view(np.float16)[:, :, :3].astype(np.float32)
    assert np.isfinite(d).all() and d.max() > 0


def test_gi_chain_state_carries_over_frames():
    """The secondary G-buffer and the reservoirs persist: background pixels keep what an earlier frame left there."""
    outs = _run(rt.SCENE_CORNELL, 1.0, 160, 96, frames=(1, 2, 3))
    for (got, ref, rays, rrays, gb) in outs:
        _check(got, ref, 160, 96, 2)
        assert rays == rrays
    assert not np.array_equal(outs[0][0]["gi_reservoirs"], outs[1][0]["gi_reservoirs"])


def test_gi_default_frame_traces_three_rays_per_pixel_at_most():
    """the default frame is K3 + K4 + K5 = 3 rays / pixel (K4 skips background pixels and below-horizon samples)."""
    (got, ref, rays, rrays, gb), = _run(rt.SCENE_CORNELL, 1.0, 128, 128)
    n = 128 * 128
    hit = int((gb["depth"] != 100000.0).sum())
    assert rays == rrays and n < rays <= n + hit           # K5 traces every launch index, K4 only valid surfaces


# ---- N2: K6 temporal + K7 spatial resampling --------------------------------------------------------------------
def _run_resampling(kind, scale, w, h, nframes, temporal=1, spatial=1, move=True, bias=2, nspatial=1, permutation=0, oflags=O.ORC_BRUTE_FORCE, final_vis=0):
    sc = rt.Scene(kind, scale, 64, 0)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    cam = sc.camera(w / h)
    g = rt.default_gconst(cam, w, h, sc.num_lights)
    g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 0
    g.enable_temporal_resampling = temporal; g.enable_spatial_resampling = spatial
    g.restir_gi.temporal_resampling_params.temporal_bias_correction_mode = bias
    g.restir_gi.temporal_resampling_params.enable_permutation_sampling = permutation
    g.restir_gi.spatial_resampling_params.spatial_bias_correction_mode = bias
    g.restir_gi.spatial_resampling_params.num_spatial_samples = nspatial
    g.restir_gi.final_shading_params.enable_final_visibility = final_vis
    lp = rt.LightPasses(ctx, w, h)
    osc = O.OracleScene(sc.desc)
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    offsets = O.fill_neighbor_offset_buffer(8192)
    state = None
    gb_prev = {k: np.zeros((h, w), dt) for k, b, dt, n in GB_KEYS}          # the product zero-fills both G-buffer sets at allocation
    results = []
    for frame in range(1, nframes + 1):
        if move:
            cam.position[0] += 0.02; cam.position[1] += 0.01                 # a slow dolly: sub-pixel to few-pixel motion vectors
        rt.advance_frame(g, cam, w, h, frame)
        lp.update_uniform(g)
        lp.execute(frame)
        gb = _read_gb(ctx, w, h)
        motion = ctx.read(rt.BUF_MOTION_VECTORS, np.float32).reshape(h, w, 4)
        got = {"secondary": ctx.read(rt.BUF_SECONDARY_GBUFFER, np.uint32).reshape(-1, 12), "gi_reservoirs": ctx.read(rt.BUF_GI_RESERVOIRS, np.uint32).reshape(-1, 8),
               "diffuse": ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4), "specular": ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4)}
        rays = ctx.stats().rays_traced - w * h
        state = osc.gi_chain(g, frame, w, h, gb, pitch, state=state, flags=oflags, gb_prev=gb_prev, motion=motion, neighbor_offsets=offsets)
        results.append((got, {k: v.copy() for k, v in state.items() if k != "rays"}, rays, sum(state["rays"]), motion))
        # keep the oracle on the product's trajectory: a last-bit difference must not snowball through the history
        for k in ("secondary", "gi_reservoirs", "diffuse", "specular"):
            state[k] = got[k].copy()
        gb_prev = gb
    tab = ctx.read(rt.BUF_NEIGHBOR_OFFSETS, np.int8).reshape(-1, 2)
    ctx.close()
    return results, tab, offsets


@pytest.mark.parametrize("kind,scale,w,h,kw", [
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict()),                                          # reference defaults with both passes on
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(bias=3, nspatial=3, permutation=1, final_vis=1)),   # ray-traced bias correction: visibility rays in K6/K7
    (rt.SCENE_CORNELL, 1.0, 160, 96, dict(bias=0, nspatial=4)),                         # 1/M normalisation
    (rt.SCENE_HALL, 0.125, 320, 180, dict(nspatial=2, oflags=0)),
    (rt.SCENE_CORNELL, 1.0, 160, 96, dict(spatial=0)),                                  # temporal alone: page 1 is never written, no history (src/main.rs:241-247)
    (rt.SCENE_CORNELL, 1.0, 160, 96, dict(temporal=0, nspatial=3, move=False)),
])
def test_gi_resampling_bit_exact_over_frames(kind, scale, w, h, kw):
    results, tab, offsets = _run_resampling(kind, scale, w, h, 4, **kw)
    assert np.array_equal(tab, offsets)                                                 # neighbour-offset table (Q11 pin)
    tol = max(2, w * h // 50000)
    for (got, ref, rays, rrays, motion) in results:
        _check(got, ref, w, h, tol)
        assert abs(rays - rrays) <= tol                                                 # a pixel that differs may also trace differently
    if kw.get("move", True):
        assert np.abs(results[-1][4][:, :, :2]).max() > 0.05                           # the camera really moved
    if kw.get("temporal", 1) and kw.get("spatial", 1):
        M = results[-1][0]["gi_reservoirs"][:, 3] & 0xff
        assert M.max() > 2                                                              # history was found and merged


def test_gi_resampling_rejects_aliased_pages():
    sc, ctx, g, lp = _ctx(rt.SCENE_CORNELL, 1.0, 64, 64)
    g.enable_restir_gi = 1; g.enable_spatial_resampling = 1
    g.restir_gi.buffer_indices.spatial_resampling_output_buffer_index = g.restir_gi.buffer_indices.spatial_resampling_input_buffer_index
    lp.update_uniform(g)
    with pytest.raises(rt.Rt2Error):
        lp.execute(1)
    ctx.close()


# ---- row-band sharding of the per-pixel ReSTIR passes (K3 / K9 / K4 / K5 / K8 shard with no exchange) ---------------
@pytest.mark.parametrize("kind,scale,mode", [
    (rt.SCENE_HALL, 0.125, dict(gi=1, di=0)),                # the default frame K3 -> K4 -> K5 -> K8
    (rt.SCENE_HALL, 0.125, dict(gi=0, di=1)),                # K3 -> K9
    (rt.SCENE_CORNELL, 1.0, dict(gi=1, di=1)),               # K9 + the DI branch inside K5
    (rt.SCENE_INSTANCED, 0.05, dict(gi=1, di=0, env=1)),     # flattened instancing + sky
])
def test_restir_frames_shard_by_row_bands_bit_identically(kind, scale, mode):
    """Every rank renders its interleaved bands into the full-frame buffers of one context: after all ranks ran, the lighting
    images, G-buffer and packed reservoirs equal the whole-frame render bit for bit, and the ray counts add up."""
    w, h = 200, 150
    sc = rt.Scene(kind, scale, 32, 0)
    cam = sc.camera(w / h)
    d = sc.desc.contents
    g = rt.default_gconst(cam, w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
    g.refrence_mode = 0; g.enable_restir_gi = mode["gi"]; g.enable_restir_di = mode["di"]; g.environment = mode.get("env", 0)
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1; isp.enable_initial_visibility = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    bufs = [rt.BUF_DIFFUSE_F16, rt.BUF_SPECULAR_F16, rt.BUF_GBUFFER_DEPTH, rt.BUF_GBUFFER_NORMALS, rt.BUF_GBUFFER_DIFFUSE_ALBEDO, rt.BUF_GBUFFER_EMISSIVE]
    if mode["gi"]:
        bufs += [rt.BUF_GI_RESERVOIRS, rt.BUF_SECONDARY_GBUFFER]
    if mode["di"]:
        bufs += [rt.BUF_DI_RESERVOIRS]

    def run(world, band_rows):
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        lp = rt.LightPasses(ctx, w, h)
        rt.PrepareLightsTasks(ctx).execute()
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
        rays = 0
        for frame in (1, 2):
            rt.advance_frame(g, cam, w, h, frame)
            lp.update_uniform(g)
            for rank in range(world):
                lp.execute(frame, rt.LaunchParams(band_rows=band_rows if world > 1 else 0, band_rank=rank, band_world=world))
                rays += ctx.stats().rays_traced
        out = {b: ctx.read(b, np.uint8).copy() for b in bufs}
        ctx.close()
        return out, rays

    want, rays = run(1, 0)
    for world, band_rows in ((2, 8), (3, 16), (5, 4)):
        got, r = run(world, band_rows)
        assert r == rays, (world, r, rays)
        for b in bufs:
            assert np.array_equal(got[b], want[b]), f"buffer {b} differs with {world} ranks of {band_rows}-row bands"


def test_resampling_is_rejected_under_row_band_sharding():
    sc = rt.Scene(rt.SCENE_CORNELL)
    g = rt.default_gconst(sc.camera(1.0), 64, 64, sc.num_lights)
    g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_spatial_resampling = 1
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, 64, 64)
    lp.update_uniform(g)
    lp.execute(1)                                   # whole frame: fine
    with pytest.raises(rt.Rt2Error, match="halo"):
        lp.execute(1, rt.LaunchParams(band_rows=8, band_rank=0, band_world=2))
    ctx.close()
