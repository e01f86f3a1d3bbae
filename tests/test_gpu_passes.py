"""GPU parity of the small passes through the C ABI against the CPU oracle: K10 prepare_lights, K12
mip_pass (local-light r32f and environment r16f pyramids), K13 postprocess.  log2 / exp2 / pow are
pinned on both sides to "binary64 then round", so packed integers, fp32 fluxes, fp16 mips and the
RGBA8 image are compared bit for bit (a handful of 1-LSB differences are tolerated where a binary64
libm result straddles a rounding boundary; the bound is written at each assert)."""
import ctypes as C
import math

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

pytestmark = pytest.mark.gpu


def _setup(kind, scale, tex=16, seed=0):
    sc = rt.Scene(kind, scale, tex, seed)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    return sc, ctx


def _read_mips(ctx, base, w, h, mips, dtype):
    return [ctx.read(base + k, dtype).reshape(mh, mw) for k, (mw, mh) in enumerate(O.mip_dims(w, h, mips))]


@pytest.mark.parametrize("kind,scale", [(rt.SCENE_CORNELL, 1.0), (rt.SCENE_HALL, 0.125), (rt.SCENE_MANYLIGHT, 0.3), (rt.SCENE_MANYLIGHT, 1.0)])
def test_prepare_lights_and_local_mips_bit_exact(kind, scale):
    sc, ctx = _setup(kind, scale)
    rt.PrepareLightsTasks(ctx).execute()
    w, h, mips = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(w), C.byref(h), C.byref(mips))
    w, h, mips = w.value, h.value, mips.value
    osc = O.OracleScene(sc.desc)
    tasks, g2l, lights, pdf = osc.prepare_lights(w, h)
    assert np.array_equal(ctx.read(rt.BUF_LIGHT_TASKS, np.uint32).reshape(-1, 4), tasks)
    assert np.array_equal(ctx.read(rt.BUF_GEOM_TO_LIGHT, np.uint32), g2l)
    got = ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12)
    assert got.shape == lights.shape
    bad = np.flatnonzero((got != lights).any(axis=1))
    assert len(bad) <= max(1, len(lights) // 100000), f"{len(bad)} of {len(lights)} packed lights differ, first {bad[:4]}"
    got_pdf = ctx.read(rt.BUF_LOCAL_PDF_MIP0, np.float32).reshape(h, w)
    nbad = int((got_pdf.view(np.uint32) != pdf.view(np.uint32)).sum())
    assert nbad <= max(1, len(lights) // 100000), f"{nbad} PDF texels differ"
    # K12 on top of it: seed the oracle with the GPU's mip 0 so the chain itself is compared exactly
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(w, h, mips)), np.float32)
    packed[:w * h] = got_pdf.ravel()
    osc.generate_mips(0, packed, w, h, mips)
    ref = O.split_mips(packed, w, h, mips)
    for k, m in enumerate(_read_mips(ctx, rt.BUF_LOCAL_PDF_MIP0, w, h, mips, np.float32)):
        assert np.array_equal(m.view(np.uint32), ref[k].view(np.uint32)), f"local PDF mip {k}"
    if w == h and mips > 1:   # checksum of checksums: the top texel is the mean flux
        top = float(_read_mips(ctx, rt.BUF_LOCAL_PDF_MIP0, w, h, mips, np.float32)[-1][0, 0])
        assert abs(top - float(got_pdf.astype(np.float64).mean())) <= 1e-4 * top
    ctx.close()


def test_prepare_lights_requires_scene():
    ctx = rt.Renderer(0)
    with pytest.raises(rt.Rt2Error):
        rt.PrepareLightsTasks(ctx).execute()
    ctx.close()


@pytest.mark.parametrize("scale", [0.05, 1.0])
def test_environment_mips_bit_exact(scale):
    sc, ctx = _setup(rt.SCENE_INSTANCED, scale)
    d = sc.desc.contents
    w, h = d.sky_width, d.sky_height
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    mips = 0
    while ctx.buffer_size(rt.BUF_ENV_PDF_MIP0 + mips):
        mips += 1
    assert mips == int(math.ceil(math.log2(max(w, h)))) + 1
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(w, h, mips)), np.uint16)
    O.OracleScene(sc.desc).generate_mips(1, packed, w, h, mips)
    ref = O.split_mips(packed, w, h, mips)
    total_bad = 0
    for k, m in enumerate(_read_mips(ctx, rt.BUF_ENV_PDF_MIP0, w, h, mips, np.uint16)):
        total_bad += int((m != ref[k]).sum())
    assert total_bad == 0, f"{total_bad} environment PDF texels differ"
    ctx.close()


def test_environment_mips_without_sky_fails_loudly():
    sc, ctx = _setup(rt.SCENE_CORNELL, 1.0)
    with pytest.raises(rt.Rt2Error):
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    ctx.close()


@pytest.mark.parametrize("kind,scale,w,h", [(rt.SCENE_CORNELL, 1.0, 512, 512), (rt.SCENE_HALL, 0.125, 320, 180)])
def test_postprocess_reference_mode_bit_exact(kind, scale, w, h):
    sc, ctx = _setup(kind, scale, tex=64)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    lp.execute(1, rt.LaunchParams(spp=2, max_bounces=3))
    rt.PostProcessPass(ctx).execute(1)
    d16 = ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4)
    out = ctx.read(rt.BUF_OUTPUT_RGBA8, np.uint8).reshape(h, w, 4)
    ref, _ = O.OracleScene(sc.desc).postprocess(g, w, h, d16)
    diff = np.abs(out.astype(np.int32) - ref.astype(np.int32))
    assert diff.max() <= 1 and int((diff > 0).sum()) <= 4, f"{int((diff > 0).sum())} channels differ, max {diff.max()}"
    assert (out[:, :, 3] == 255).all() and out[:, :, :3].max() > 0
    ctx.close()


def test_postprocess_synthetic_radiance_sweep():
    """Full dynamic range through the tonemap: write a synthetic DiffuseLighting image, compare all pixels."""
    w, h = 256, 64
    sc, ctx = _setup(rt.SCENE_CORNELL, 1.0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    rng = np.random.default_rng(11)
    d = np.zeros((h, w, 4), np.float16)
    d[:, :, :3] = (10.0 ** rng.uniform(-5, 3, (h, w, 3))).astype(np.float16)
    d[0, :8, :3] = 0
    d[1, :8, :3] = np.float16(65504)
    ctx.write(rt.BUF_DIFFUSE_F16, d.view(np.uint16))
    rt.PostProcessPass(ctx).execute(1)
    out = ctx.read(rt.BUF_OUTPUT_RGBA8, np.uint8).reshape(h, w, 4)
    ref, _ = O.OracleScene(sc.desc).postprocess(g, w, h, d.view(np.uint16))
    diff = np.abs(out.astype(np.int32) - ref.astype(np.int32))
    assert diff.max() <= 1 and int((diff > 0).sum()) <= 4
    ctx.close()


# ---- K3 g_buffer / K9 di_fused_resampling ---------------------------------------------------------------
GB_KEYS = [("depth", rt.BUF_GBUFFER_DEPTH, np.float32, 1), ("normals", rt.BUF_GBUFFER_NORMALS, np.uint32, 1),
           ("geo_normals", rt.BUF_GBUFFER_GEO_NORMALS, np.uint32, 1), ("diffuse_albedo", rt.BUF_GBUFFER_DIFFUSE_ALBEDO, np.uint32, 1),
           ("specular_rough", rt.BUF_GBUFFER_SPECULAR_ROUGH, np.uint32, 1), ("emissive", rt.BUF_GBUFFER_EMISSIVE, np.uint16, 4),
           ("motion", rt.BUF_MOTION_VECTORS, np.float32, 4)]


def _restir_setup(kind, scale, w, h, tex=64):
    sc, ctx = _setup(kind, scale, tex=tex)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 0
    lp = rt.LightPasses(ctx, w, h)
    return sc, ctx, g, lp


def _read_gbuffer(ctx, w, h):
    return {k: (ctx.read(b, dt).reshape(h, w) if n == 1 else ctx.read(b, dt).reshape(h, w, n)) for k, b, dt, n in GB_KEYS}


@pytest.mark.parametrize("kind,scale,w,h,brute", [
    (rt.SCENE_CORNELL, 1.0, 512, 512, True), (rt.SCENE_HALL, 0.125, 320, 180, True), (rt.SCENE_HALL, 0.5, 480, 270, False),
    (rt.SCENE_INSTANCED, 0.05, 320, 180, True), (rt.SCENE_MANYLIGHT, 0.1, 320, 180, True)])
@pytest.mark.parametrize("frame", [1, 2])
def test_gbuffer_bit_exact(kind, scale, w, h, brute, frame):
    sc, ctx, g, lp = _restir_setup(kind, scale, w, h)
    lp.update_uniform(g)
    lp.execute(frame)
    got = _read_gbuffer(ctx, w, h)
    ref = O.OracleScene(sc.desc).gbuffer(g, w, h, flags=O.ORC_BRUTE_FORCE if brute else 0)
    st = ctx.stats()
    assert st.rays_traced == w * h and st.rays_hit == int((ref["depth"] != 100000.0).sum())
    for k in ("depth", "normals", "geo_normals", "diffuse_albedo", "emissive", "motion"):
        assert np.array_equal(got[k].view(np.uint32 if got[k].dtype.itemsize == 4 else np.uint16),
                              ref[k].view(np.uint32 if ref[k].dtype.itemsize == 4 else np.uint16)), k
    nbad = int((got["specular_rough"] != ref["specular_rough"]).sum())          # pow(x, 1/2.2) through binary64: <= 1e-6 of pixels may flip an LSB
    assert nbad <= max(1, w * h // 1000000), nbad
    assert (got["depth"] > 0).all()
    ctx.close()


def _di_run(kind, scale, w, h, frame, **di):
    sky = di.get("sky", 0)              # > 0: give the scene the procedural sky box of that width (the instanced scene has its own)
    if sky or kind == rt.SCENE_INSTANCED:
        sc = rt.Scene(kind, scale, 64, 0)
        if sky:
            sc.add_sky(sky)
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, sc)
        d = sc.desc.contents
        g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
        g.refrence_mode = 0; g.enable_restir_gi = 0; g.environment = 1
        lp = rt.LightPasses(ctx, w, h)
    else:
        sc, ctx, g, lp = _restir_setup(kind, scale, w, h)
    g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = di.get("local", 4); isp.num_primary_brdf_samples = di.get("brdf", 1)
    isp.enable_initial_visibility = di.get("initial_vis", 0); isp.brdf_cutoff = di.get("cutoff", 0.0)
    g.restir_di.shading_params.enable_final_visibility = di.get("final_vis", 0)
    g.enable_accumulation = di.get("accumulate", 0); g.blend_factor = 0.25
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    isp.local_light_sampling_mode = di.get("mode", 0)
    isp.num_primary_environment_samples = di.get("env", 0); isp.environment_map_importance_sampling = di.get("env_is", 0)
    lp.update_uniform(g)
    presampling = di.get("presampling", False)
    ris = ris_data = env_pdf = None
    env_dims = (0, 0, 0)
    has_sky = sc.desc.contents.sky_width != 0
    if presampling and has_sky and not di.get("skip_env_mips", False):
        rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
        sw, sh = sc.desc.contents.sky_width, sc.desc.contents.sky_height
        env_dims = (sw, sh, int(math.ceil(math.log2(max(sw, sh)))) + 1)
        env_pdf = np.concatenate([m.ravel() for m in _read_mips(ctx, rt.BUF_ENV_PDF_MIP0, *env_dims, np.uint16)])
    if presampling:                 # K11 fills the tiles the POWER_RIS / environment samplers draw from (its own parity: test_presample_*_bit_exact)
        lp.execute_presampeling(frame, env_pdf is not None)
        ris = ctx.read(rt.BUF_RIS, np.uint32).reshape(-1, 2)
        ris_data = ctx.read(rt.BUF_RIS_LIGHT_DATA, np.uint32).reshape(-1, 8)
    if di.get("accumulate", 0):     # seed prior lighting so the accumulation branch reads something
        rng = np.random.default_rng(4)
        prior = np.zeros((h, w, 4), np.float16); prior[:, :, :3] = rng.random((h, w, 3))
        ctx.write(rt.BUF_DIFFUSE_F16, prior.view(np.uint16)); ctx.write(rt.BUF_SPECULAR_F16, prior.view(np.uint16))
    else:
        prior = None
    lp.execute(frame, rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING) if presampling else None)
    gb = _read_gbuffer(ctx, w, h)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    pdf = np.concatenate([m.ravel() for m in _read_mips(ctx, rt.BUF_LOCAL_PDF_MIP0, pw.value, ph.value, pm.value, np.float32)])
    lights = ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12)
    g2l = ctx.read(rt.BUF_GEOM_TO_LIGHT, np.uint32)
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    oflags = di.get("oflags", O.ORC_BRUTE_FORCE) | (O.ORC_PRESAMPLING if presampling else 0)
    d_in = prior.view(np.uint16).copy() if prior is not None else None
    s_in = prior.view(np.uint16).copy() if prior is not None else None
    res, d16, s16, rays = O.OracleScene(sc.desc).di_fused_resampling(g, frame, w, h, gb, lights, g2l, pdf, pw.value, ph.value, pm.value, pitch,
                                                                     diffuse_f16=d_in, specular_f16=s_in, flags=oflags, ris=ris, ris_light_data=ris_data,
                                                                     env_pdf=env_pdf, env_dims=env_dims)
    got_res = ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).reshape(-1, 6)
    got_d = ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4)
    got_s = ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4)
    st = ctx.stats()
    ctx.close()
    return (got_res, got_d, got_s, st.rays_traced - w * h), (res, d16, s16, rays)


@pytest.mark.parametrize("kind,scale,w,h,di", [
    (rt.SCENE_CORNELL, 1.0, 512, 512, dict(local=4, brdf=1)),                      # BASELINE config 1 size
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict(local=0, brdf=1)),                      # the reference's default sample counts (src/main.rs:322-331)
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict(local=8, brdf=2, initial_vis=1, final_vis=1)),
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(local=3, brdf=1, cutoff=0.1, final_vis=1, accumulate=1)),
    (rt.SCENE_HALL, 0.125, 320, 180, dict(local=4, brdf=1, final_vis=1, oflags=0)),
    (rt.SCENE_MANYLIGHT, 0.2, 320, 180, dict(local=8, brdf=1, initial_vis=1, oflags=0)),     # config 3 shape: many emissive triangles
])
def test_di_reservoir_sampling_bit_exact(kind, scale, w, h, di):
    (res, d16, s16, rays), (rres, rd16, rs16, rrays) = _di_run(kind, scale, w, h, 1, **di)
    nvalid = int((rres[:, 0] != 0).sum())
    assert nvalid > 0 and rays == rrays
    tol = max(2, w * h // 50000)     # log2/exp2/pow via binary64 may round a last bit differently in <= 2e-5 of pixels
    bad = np.flatnonzero((res != rres).any(axis=1))
    assert len(bad) <= tol, f"{len(bad)} reservoirs differ (of {nvalid} valid), first {bad[:4]}"
    assert int((d16 != rd16).any(axis=2).sum()) <= tol and int((s16 != rs16).any(axis=2).sum()) <= tol
    f = d16.view(np.float16)[:, :, :3].astype(np.float32)
    assert np.isfinite(f).all() and f.max() > 0


@pytest.mark.parametrize("kind,scale,w,h,di", [
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict(local=4, brdf=1, presampling=True, mode=1)),
    (rt.SCENE_MANYLIGHT, 0.2, 320, 180, dict(local=8, brdf=1, initial_vis=1, final_vis=1, oflags=0, presampling=True, mode=1)),
    (rt.SCENE_MANYLIGHT, 0.2, 320, 180, dict(local=8, brdf=1, oflags=0, presampling=True, mode=0)),   # uniform selection, but the environment combine draw shifts the RNG
    (rt.SCENE_MANYLIGHT, 0.2, 200, 120, dict(local=2, brdf=2, final_vis=1, oflags=0, presampling=True, mode=1)),  # 2 BRDF samples: megakernel form
])
def test_di_power_ris_sampling_bit_exact(kind, scale, w, h, di):
    """the light sampler as compiled with RTXDI_ENABLE_PRESAMPLING 1 — POWER_RIS local lights drawn from K11's tiles."""
    (res, d16, s16, rays), (rres, rd16, rs16, rrays) = _di_run(kind, scale, w, h, 3, **di)
    nvalid = int((rres[:, 0] != 0).sum())
    assert nvalid > 0 and rays == rrays
    tol = max(2, w * h // 50000)
    bad = np.flatnonzero((res != rres).any(axis=1))
    assert len(bad) <= tol, f"{len(bad)} reservoirs differ (of {nvalid} valid), first {bad[:4]}"
    assert int((d16 != rd16).any(axis=2).sum()) <= tol and int((s16 != rs16).any(axis=2).sum()) <= tol
    # the presampled branch really ran: the same frame without it selects different lights
    (res0, _, _, _), _ = _di_run(kind, scale, w, h, 3, **{**di, "presampling": False, "mode": 0})
    ndiff = int((res0[:, 0] != res[:, 0]).sum())
    assert ndiff > (nvalid // 4 if di["mode"] == 1 else 0)      # mode 0: only the environment combine's extra draw differs


@pytest.mark.parametrize("kind,scale,w,h,di", [
    # the instanced scene with its sky on, numPrimaryEnvironmentSamples = 4 (no local lights in this scene)
    (rt.SCENE_INSTANCED, 0.05, 320, 180, dict(local=0, brdf=1, env=4, env_is=1, final_vis=1, oflags=0, presampling=True)),
    (rt.SCENE_INSTANCED, 0.05, 200, 120, dict(local=0, brdf=1, env=4, env_is=0, initial_vis=1, final_vis=1, oflags=0, presampling=True)),
    # lamp + sky: local, environment and BRDF reservoirs all populated (the open Cornell box with the procedural sky box)
    (rt.SCENE_CORNELL, 1.0, 256, 256, dict(local=4, brdf=1, env=4, env_is=1, final_vis=1, presampling=True, mode=1, sky=256)),
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(local=2, brdf=1, env=2, env_is=1, cutoff=0.1, initial_vis=1, final_vis=1, accumulate=1, presampling=True, sky=128)),
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(local=2, brdf=2, env=3, env_is=1, final_vis=1, presampling=True, sky=128)),     # 2 BRDF samples: megakernel form
    (rt.SCENE_CORNELL, 1.0, 200, 120, dict(local=4, brdf=1, env=0, env_is=1, final_vis=1, presampling=True, sky=128)),     # the sky only through BRDF misses
    (rt.SCENE_MANYLIGHT, 0.2, 320, 180, dict(local=8, brdf=1, env=4, env_is=1, final_vis=1, oflags=0, presampling=True, mode=1, sky=512)),
])
def test_di_environment_map_sampling_bit_exact(kind, scale, w, h, di):
    """SURVEY §8f N3, second half: RTXDI_SampleEnvironmentMap (InitialSamplingFunctions.hlsli:465-499) drawing from K11's environment
    tiles, the BRDF-miss environment branch (:556-564) and RAB_EvaluateEnvironmentMapSamplingPdf (bridge :397-418), with the
    environment light K10 writes at lights + 1 (Q7 pinned to intent)."""
    (res, d16, s16, rays), (rres, rd16, rs16, rrays) = _di_run(kind, scale, w, h, 3, **di)
    valid = rres[:, 0] != 0
    nvalid = int(valid.sum())
    assert nvalid > 0 and rays == rrays
    tol = max(2, w * h // 20000)     # sky lookups go through atan2 / asin in binary64 on both sides: a last-bit difference moves a uv by one ulp
    bad = np.flatnonzero((res != rres).any(axis=1))
    assert len(bad) <= tol, f"{len(bad)} reservoirs differ (of {nvalid} valid), first {bad[:4]}"
    assert int((d16 != rd16).any(axis=2).sum()) <= tol and int((s16 != rs16).any(axis=2).sum()) <= tol
    f = d16.view(np.float16)[:, :, :3].astype(np.float32)
    assert np.isfinite(f).all() and f.max() > 0
    # the environment light is among the selected lights
    n_lights = int(rres[valid, 0].max() & 0x7fffffff) - 1 if kind == rt.SCENE_INSTANCED else None
    env_index = (rres[valid, 0] & 0x7fffffff).max()
    n_env = int(((res[:, 0] & 0x7fffffff) == env_index).sum())
    assert n_env > nvalid // 20, (n_env, nvalid)
    if n_lights is not None:
        assert n_lights == 0


def test_environment_map_sampling_needs_its_inputs():
    w, h = 96, 64
    sc = rt.Scene(rt.SCENE_INSTANCED, 0.05, 16, 0)
    d = sc.desc.contents
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1; g.environment = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 0; isp.num_primary_brdf_samples = 1; isp.num_primary_environment_samples = 4
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    pre = rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING)
    with pytest.raises(rt.Rt2Error, match="compiled out"):
        lp.execute(1)                                   # the shipped build has no environment reservoir
    with pytest.raises(rt.Rt2Error, match="rt2_prepare_lights"):
        lp.execute(1, pre)                              # K10 writes the environment light
    rt.PrepareLightsTasks(ctx).execute()
    lights = ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12)
    assert lights.shape[0] == sc.num_lights + 2 and lights[-1, 3] >> 24 == 6
    ref = O.OracleScene(sc.desc).prepare_lights(0, 0, env=True)[2]
    assert np.array_equal(lights, ref)
    with pytest.raises(rt.Rt2Error, match="rt2_generate_mips"):
        lp.execute(1, pre)
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    with pytest.raises(rt.Rt2Error, match="rt2_presample"):
        lp.execute(1, pre)
    lp.execute_presampeling(1, True)
    lp.execute(1, pre)
    g.light_buffer_params.environment_light_params.light_index = sc.num_lights + 5
    lp.update_uniform(g)
    with pytest.raises(rt.Rt2Error, match="lights \\+ 1"):
        lp.execute(1, pre)
    # lightPresent == 0: no environment light, numPrimaryEnvironmentSamples is then inert (RTXDI_SampleEnvironmentMap returns at :477)
    g.light_buffer_params.environment_light_params.light_present = 0
    lp.update_uniform(g)
    lp.execute(1, pre)
    assert not ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).reshape(-1, 6)[:, 0].any()
    ctx.close()


def test_power_ris_needs_presampled_tiles():
    sc, ctx, g, lp = _restir_setup(rt.SCENE_CORNELL, 1.0, 64, 64)
    g.enable_restir_di = 1
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = 2
    g.restir_di.initial_sampling_params.local_light_sampling_mode = 1
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    lp.update_uniform(g)
    lp.execute(1)                                                   # the reference's build: the mode is ignored (presampling compiled out)
    with pytest.raises(rt.Rt2Error):
        lp.execute(1, rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING))   # rt2_presample has not run
    g.restir_di.initial_sampling_params.local_light_sampling_mode = 2
    lp.update_uniform(g)
    with pytest.raises(rt.Rt2Error):
        lp.execute(1, rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING))   # ReGIR is compiled out
    ctx.close()


def test_di_needs_light_list():
    sc, ctx, g, lp = _restir_setup(rt.SCENE_CORNELL, 1.0, 64, 64)
    g.enable_restir_di = 1
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = 2
    lp.update_uniform(g)
    with pytest.raises(rt.Rt2Error):
        lp.execute(1)               # rt2_prepare_lights has not run
    ctx.close()


# ---- K11 presampling ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,scale", [(rt.SCENE_CORNELL, 1.0), (rt.SCENE_MANYLIGHT, 0.3), (rt.SCENE_MANYLIGHT, 1.0)])
def test_presample_local_lights_bit_exact(kind, scale):
    sc, ctx = _setup(kind, scale)
    g = rt.default_gconst(sc.camera(16 / 9), 320, 180, sc.num_lights)
    lp = rt.LightPasses(ctx, 320, 180)
    lp.update_uniform(g)
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    assert lp.execute_presampeling(7, False) is True               # src/light_passes.rs:543-547
    ris = ctx.read(rt.BUF_RIS, np.uint32).reshape(-1, 2)
    data = ctx.read(rt.BUF_RIS_LIGHT_DATA, np.uint32).reshape(-1, 8)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    pdf = np.concatenate([m.ravel() for m in _read_mips(ctx, rt.BUF_LOCAL_PDF_MIP0, pw.value, ph.value, pm.value, np.float32)])
    lights = ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12)
    rris, rdata = O.presample(g, 7, 0, pdf, pw.value, ph.value, pm.value, lights, entries=len(ris))
    assert len(ris) == 128 * 1024
    assert np.array_equal(ris, rris) and np.array_equal(data, rdata)
    assert (ris[:, 1].view(np.float32) > 0).all()
    ctx.close()


def test_presample_environment_bit_exact():
    sc, ctx = _setup(rt.SCENE_INSTANCED, 0.125)
    d = sc.desc.contents
    w, h = d.sky_width, d.sky_height
    g = rt.default_gconst(sc.camera(16 / 9), 320, 180, sc.num_lights, env_size=(w, h))
    lp = rt.LightPasses(ctx, 320, 180)
    lp.update_uniform(g)
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    assert lp.execute_presampeling(2, True) is False
    ris = ctx.read(rt.BUF_RIS, np.uint32).reshape(-1, 2)
    mips = int(math.ceil(math.log2(max(w, h)))) + 1
    env = np.concatenate([m.ravel() for m in _read_mips(ctx, rt.BUF_ENV_PDF_MIP0, w, h, mips, np.uint16)])
    off = g.environment_light_risbuffer_segment_params.buffer_offset
    rris, _ = O.presample(g, 2, 1, env, w, h, mips, entries=len(ris))
    assert len(ris) == 2 * 128 * 1024
    assert np.array_equal(ris[off:], rris[off:])
    assert (ris[off:, 1].view(np.float32) > 0).all()
    ctx.close()


def test_presample_needs_its_inputs():
    sc, ctx = _setup(rt.SCENE_CORNELL, 1.0)
    g = rt.default_gconst(sc.camera(1.0), 64, 64, sc.num_lights)
    lp = rt.LightPasses(ctx, 64, 64)
    lp.update_uniform(g)
    with pytest.raises(rt.Rt2Error):
        lp.execute_presampeling(1, False)          # K10 / K12 have not run
    ctx.close()
