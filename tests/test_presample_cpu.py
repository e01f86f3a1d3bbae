"""CPU checks of the oracle's K11 presampling (rtxdi/PresamplingFunctions.hlsli:30-162): the hierarchical descent through the
PDF mip pyramid must draw light i with probability power_i / sum(power) and report exactly that probability's inverse."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O


def _local_inputs(kind, scale):
    sc = rt.Scene(kind, scale, 16, 0)
    g = rt.default_gconst(sc.camera(1.0), 64, 64, sc.num_lights)
    osc = O.OracleScene(sc.desc)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    tasks, g2l, lights, pdf0 = osc.prepare_lights(pw.value, ph.value)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(pw.value, ph.value, pm.value)), np.float32)
    packed[:pw.value * ph.value] = pdf0.ravel()
    osc.generate_mips(0, packed, pw.value, ph.value, pm.value)
    return sc, g, lights, pdf0, packed, (pw.value, ph.value, pm.value)


def test_local_light_presampling_follows_the_power_distribution():
    sc, g, lights, pdf0, packed, (w, h, m) = _local_inputs(rt.SCENE_MANYLIGHT, 0.3)
    assert g.local_lights_risbuffer_segment_params.tile_size == 1024 and g.local_lights_risbuffer_segment_params.tile_count == 128   # src/main.rs:305-310
    ris, data = O.presample(g, 3, 0, packed, w, h, m, lights)
    seg = ris[:128 * 1024]
    idx = seg[:, 0]
    inv = seg[:, 1].view(np.float32)
    assert (idx >> 31 == 1).all()                                   # every K10 light packs compactly (no spot shaping)
    li = idx & 0x7fffffff
    n = len(lights)
    assert li.max() < n and (inv > 0).all()
    # light i sits at Z-curve texel i, whose mip-0 value is its power
    L = O.lib()
    power = np.zeros(n)
    x, y = C.c_uint32(), C.c_uint32()
    for i in range(n):
        L.orc_linear_to_zcurve(i, C.byref(x), C.byref(y)); power[i] = pdf0[y.value, x.value]
    p = power / power.sum()
    # the reported inverse pdf is the product of the per-level choices == 1 / p_i (up to fp32 rounding through <= 13 levels)
    assert np.allclose(1.0 / inv.astype(np.float64), p[li], rtol=2e-5)
    # and the empirical histogram matches p (chi-square-ish bound on 131072 draws)
    counts = np.bincount(li, minlength=n)
    expect = p * len(li)
    big = expect > 20
    z = (counts[big] - expect[big]) / np.sqrt(expect[big])
    assert np.abs(z).max() < 6 and abs(z.mean()) < 0.3
    assert (counts[power == 0] == 0).all()
    # the compact record is the light itself
    sel = np.arange(0, len(seg), 997)
    for k in sel:
        assert np.array_equal(data[k, 0:3], lights[li[k], 0:3]) and data[k, 3] == lights[li[k], 3]
        assert np.array_equal(data[k, 4:8], lights[li[k], 4:8])
    # a different frame draws different samples; the same frame the same
    ris2, _ = O.presample(g, 3, 0, packed, w, h, m, lights)
    ris3, _ = O.presample(g, 4, 0, packed, w, h, m, lights)
    assert np.array_equal(ris, ris2) and not np.array_equal(ris, ris3)
    assert not ris[128 * 1024:].any()                               # the environment segment is untouched


def test_single_light_and_empty_pyramids():
    sc, g, lights, pdf0, packed, (w, h, m) = _local_inputs(rt.SCENE_CORNELL, 1.0)
    ris, data = O.presample(g, 1, 0, packed, w, h, m, lights)
    li = ris[:128 * 1024, 0] & 0x7fffffff
    assert set(np.unique(li)) <= set(range(len(lights))) and len(lights) == sc.num_lights
    # all-zero pyramid: nothing found -> (firstLightIndex, 0) entries, no compact bit
    zero = np.zeros_like(packed)
    ris0, _ = O.presample(g, 1, 0, zero, w, h, m, lights)
    assert (ris0[:128 * 1024, 1] == 0).all() and (ris0[:128 * 1024, 0] >> 31 == 0).all()


def test_environment_presampling_importance_samples_the_sky():
    sc = rt.Scene(rt.SCENE_INSTANCED, 0.125, 16, 0)         # 512 x 256 sky: the descent doubles the position per level, so it needs a power-of-two pyramid
    d = sc.desc.contents
    w, h = d.sky_width, d.sky_height
    assert (w, h) == (512, 256)
    g = rt.default_gconst(sc.camera(1.0), 64, 64, sc.num_lights, env_size=(w, h))
    osc = O.OracleScene(sc.desc)
    import math
    mips = int(math.ceil(math.log2(max(w, h)))) + 1
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(w, h, mips)), np.uint16)
    osc.generate_mips(1, packed, w, h, mips)
    ris, _ = O.presample(g, 2, 1, packed, w, h, mips)
    off = g.environment_light_risbuffer_segment_params.buffer_offset
    assert off == 128 * 1024 and not ris[:off].any()                # src/main.rs:299-304
    seg = ris[off:off + 128 * 1024]
    inv = seg[:, 1].view(np.float32)
    assert (inv > 0).all()
    u = (seg[:, 0] & 0xffff) / 65535.0; v = (seg[:, 0] >> 16) / 65535.0
    tx = np.minimum((u * w).astype(int), w - 1); ty = np.minimum((v * h).astype(int), h - 1)
    pdf0 = packed[:w * h].view(np.float16).astype(np.float64).reshape(h, w)
    # texels are drawn in proportion to their mip-0 weight: compare row marginals (rows are wide enough to have counts)
    rows = np.bincount(ty, minlength=h) / len(ty)
    want = pdf0.sum(axis=1) / pdf0.sum()
    assert np.abs(rows - want).max() < 0.01
    # 1 / inv == texel probability (the pyramid is r16f, so only to half precision)
    p = pdf0 / pdf0.sum()
    rel = np.abs(1.0 / inv.astype(np.float64) - p[ty, tx]) / p[ty, tx]
    assert np.median(rel) < 2e-3 and (rel < 0.03).mean() > 0.99      # uv is quantised to 16 bits: a sample on a texel edge can decode into the neighbour


def test_power_ris_light_sampling_is_an_unbiased_alternative_to_uniform_selection():
    """with the sampler's RTXDI_ENABLE_PRESAMPLING 1 branches (ORC_PRESAMPLING) and localLightSamplingMode
    POWER_RIS, local lights come from K11's tiles with 1/pdf = sum(power)/power_i.  The direct-lighting estimate must
    agree in expectation with uniform light selection (same surfaces, independent estimators), and the selected lights
    must lean towards the powerful ones."""
    w, h = 48, 27
    sc, g, lights, pdf0, packed, dims = _local_inputs(rt.SCENE_MANYLIGHT, 0.1)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    osc = O.OracleScene(sc.desc)
    gb = osc.gbuffer(g, w, h)
    _, g2l, _, _ = osc.prepare_lights(dims[0], dims[1])
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    frames = 40
    acc = {0: np.zeros((h, w, 3)), 1: np.zeros((h, w, 3))}
    chosen = {0: [], 1: []}
    for mode in (0, 1):
        isp.local_light_sampling_mode = mode
        for f in range(1, frames + 1):
            ris, data = O.presample(g, f, 0, packed, *dims, lights)
            res, d16, s16, rays = osc.di_fused_resampling(g, f, w, h, gb, lights, g2l, packed, *dims, pitch,
                                                          flags=O.ORC_PRESAMPLING, ris=ris, ris_light_data=data)
            acc[mode] += d16.view(np.float16)[:, :, :3].astype(np.float64)
            valid = res[:, 0] != 0
            chosen[mode].append(res[valid, 0] & 0x7fffffff)
    a, b = acc[0].mean() / frames, acc[1].mean() / frames
    assert a > 0 and abs(a - b) < 0.06 * a, (a, b)
    # mean power of the lights that end up in the reservoirs: higher with power-proportional candidates
    L = O.lib(); x, y = C.c_uint32(), C.c_uint32()
    power = np.zeros(len(lights))
    for i in range(len(lights)):
        L.orc_linear_to_zcurve(i, C.byref(x), C.byref(y)); power[i] = pdf0[y.value, x.value]
    p0 = power[np.concatenate(chosen[0])].mean(); p1 = power[np.concatenate(chosen[1])].mean()
    if power.std() > 0.05 * power.mean():
        assert p1 > p0, (p0, p1)
    # ORC_PRESAMPLING with uniform selection differs from the shipped build only by the environment combine's extra draw
    isp.local_light_sampling_mode = 0
    r_plain = osc.di_fused_resampling(g, 1, w, h, gb, lights, g2l, packed, *dims, pitch)[0]
    r_pre = osc.di_fused_resampling(g, 1, w, h, gb, lights, g2l, packed, *dims, pitch, flags=O.ORC_PRESAMPLING)[0]
    assert not np.array_equal(r_plain, r_pre)


def _env_inputs(sc, w, h):
    """gconst + oracle inputs of a scene with a sky: light list with the environment light at lights + 1, both PDF pyramids."""
    import math
    d = sc.desc.contents
    sw, sh = d.sky_width, d.sky_height
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(sw, sh))
    g.environment = 1
    osc = O.OracleScene(sc.desc)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    dims = (pw.value, ph.value, pm.value)
    tasks, g2l, lights, pdf0 = osc.prepare_lights(dims[0], dims[1], env=True)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(*dims)), np.float32)
    packed[:dims[0] * dims[1]] = pdf0.ravel()
    osc.generate_mips(0, packed, *dims)
    emips = int(math.ceil(math.log2(max(sw, sh)))) + 1
    epacked = np.zeros(sum(a * b for a, b in O.mip_dims(sw, sh, emips)), np.uint16)
    osc.generate_mips(1, epacked, sw, sh, emips)
    return g, osc, g2l, lights, packed, dims, epacked, (sw, sh, emips)


def test_environment_light_record():
    """The record K10 writes at lights + 1 (SURVEY Q7 pinned to intent): type kEnvironment, radiance scale 1 through packLightColor,
    rotation 0.25 turns | importance sampled, the PDF texture's size."""
    sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 16, 0).add_sky(256)
    g, osc, g2l, lights, packed, dims, epacked, edims = _env_inputs(sc, 32, 32)
    assert lights.shape == (sc.num_lights + 2, 12)
    assert g.light_buffer_params.environment_light_params.light_present == 1
    assert g.light_buffer_params.environment_light_params.light_index == sc.num_lights + 1          # src/main.rs:381-386
    env = lights[sc.num_lights + 1]
    assert not lights[sc.num_lights].any()
    assert env[3] >> 24 == 6 and env[3] & 0xffffff == 0xffffff                                       # kEnvironment, colour 1,1,1
    assert env[4] == 0 and env[5] == 256 | (128 << 16)
    assert env[6] == (1 << 16) | int(np.float16(0.25).view(np.uint16))
    scale = 2.0 ** ((int(env[7]) - 1) / 65534.0 * 48.0 - 8.0)                                        # unpackLightRadiance
    assert 1.0 <= scale < 1.001
    # a scene without a sky gets no environment record, whatever the capacity
    sc2 = rt.Scene(rt.SCENE_CORNELL, 1.0, 16, 0)
    l2 = O.OracleScene(sc2.desc).prepare_lights(dims[0], dims[1], env=True)[2]
    assert not l2[sc2.num_lights:].any()


@pytest.mark.parametrize("scene,local", [("instanced", 0), ("cornell+sky", 2)])
def test_environment_map_sampling_agrees_with_the_path_tracers_sky_lighting(scene, local):
    """SURVEY §8f N3, second half: RTXDI_SampleEnvironmentMap (InitialSamplingFunctions.hlsli:465-499) fed by K11's environment tiles,
    plus the BRDF-miss environment branch (:556-564), MIS-combined with the local-light and BRDF reservoirs.  E[DI diffuse] * albedo
    must equal the direct lighting (emitters + sky) the brute-force path tracer K1 sees: K1(2 bounces) - K1(1 bounce) — the same
    cross-check, with the same 5 % tolerance, as test_di_light_sampling_agrees_with_brute_force_path_tracer."""
    if scene == "instanced":
        w, h = 48, 27
        sc = rt.Scene(rt.SCENE_INSTANCED, 0.125, 16, 0)
    else:
        w = h = 40
        sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 16, 0).add_sky(256)
    g, osc, g2l, lights, packed, dims, epacked, edims = _env_inputs(sc, w, h)
    gb = osc.gbuffer(g, w, h)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = local; isp.num_primary_brdf_samples = 1; isp.num_primary_environment_samples = 4
    isp.environment_map_importance_sampling = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    frames = 32
    acc = np.zeros((h, w, 3)); n_env = n_valid = 0
    for f in range(1, frames + 1):
        ris, data = O.presample(g, f, 0, packed, *dims, lights)
        ris, data = O.presample(g, f, 1, epacked, *edims, ris=ris, ris_light_data=data)
        res, d16, s16, rays = osc.di_fused_resampling(g, f, w, h, gb, lights, g2l, packed, *dims, pitch, flags=O.ORC_PRESAMPLING,
                                                      ris=ris, ris_light_data=data, env_pdf=epacked, env_dims=edims)
        acc += d16.view(np.float16)[:, :, :3].astype(np.float64)
        valid = res[:, 0] != 0
        n_valid += int(valid.sum()); n_env += int(((res[valid, 0] & 0x7fffffff) == sc.num_lights + 1).sum())
    assert n_env > 0.3 * n_valid                          # the environment light really is selected
    if local:
        assert n_env < n_valid                            # ... and so are the lamp's triangles
    L = O.lib()
    albedo = np.zeros((h, w, 3), np.float32); tmp = np.zeros(3, np.float32)
    for y in range(h):
        for x in range(w):
            L.orc_unpack_r11g11b10(int(gb["diffuse_albedo"][y, x]), tmp.ctypes.data_as(C.c_void_p)); albedo[y, x] = tmp
    direct_di = acc / frames * albedo
    g.refrence_mode = 1
    k2 = np.zeros((h, w, 3)); k1 = np.zeros((h, w, 3))
    for f in range(1, 9):
        k2 += osc.render_reference(g, f, w, h, spp=64, max_bounces=2)[2][:, :, :3]
        k1 += osc.render_reference(g, f, w, h, spp=1, max_bounces=1)[2][:, :, :3]
    direct_pt = (k2 - k1) / 8
    m = (gb["emissive"].view(np.float16)[:, :, 0] == 0) & (gb["depth"] < 1e4)
    a, b = direct_di[m].mean(), direct_pt[m].mean()
    assert b > 0 and abs(a - b) < 0.05 * b, (a, b)


def test_environment_light_through_brdf_misses_only_and_rejections():
    """numPrimaryEnvironmentSamples == 0 (what K5 passes, shade_secondary_surfaces.rgen:69): the sky is still found by BRDF rays that miss
    (InitialSamplingFunctions.hlsli:556-564) and weighted by RAB_EvaluateEnvironmentMapSamplingPdf — same expectation, more noise."""
    w, h = 48, 27
    sc = rt.Scene(rt.SCENE_INSTANCED, 0.125, 16, 0)
    g, osc, g2l, lights, packed, dims, epacked, edims = _env_inputs(sc, w, h)
    gb = osc.gbuffer(g, w, h)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 0; isp.num_primary_brdf_samples = 1; isp.num_primary_environment_samples = 0
    g.restir_di.shading_params.enable_final_visibility = 1
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    run = lambda f, **kw: osc.di_fused_resampling(g, f, w, h, gb, lights, g2l, packed, *dims, pitch, **kw)
    acc = {0: np.zeros((h, w, 3)), 1: np.zeros((h, w, 3))}
    for imp in (0, 1):                 # environmentMapImportanceSampling only changes the source pdf BRDF samples report (1.0 vs the texel's)
        isp.environment_map_importance_sampling = imp
        for f in range(1, 41):
            res, d16, _, _ = run(f, flags=O.ORC_PRESAMPLING, env_pdf=epacked, env_dims=edims)
            acc[imp] += d16.view(np.float16)[:, :, :3].astype(np.float64)
            valid = res[:, 0] != 0
            assert valid.any() and ((res[valid, 0] & 0x7fffffff) == sc.num_lights + 1).all()
    a, b = acc[0].mean(), acc[1].mean()
    assert a > 0 and abs(a - b) < 0.05 * a, (a, b)          # one technique only: the MIS weight cancels whatever the source pdf is
    # as shipped (no ORC_PRESAMPLING) there is no environment light: nothing is selected in a scene without emitters (Q7)
    res, d16, _, _ = run(1)
    assert not res[:, 0].any() and not d16.any()
    # environment samples need the presampling build; the environment light needs its record and its PDF pyramid
    isp.num_primary_environment_samples = 2
    with pytest.raises(O.OracleError) as e:
        run(1)
    assert e.value.args[0] == -3
    with pytest.raises(O.OracleError) as e:
        run(1, flags=O.ORC_PRESAMPLING)                      # no pyramid
    assert e.value.args[0] == -5
    with pytest.raises(O.OracleError) as e:
        osc.di_fused_resampling(g, 1, w, h, gb, lights[:1], g2l, packed, *dims, pitch, flags=O.ORC_PRESAMPLING, env_pdf=epacked, env_dims=edims)
    assert e.value.args[0] == -5                             # light buffer too short for index lights + 1
    with pytest.raises(O.OracleError) as e:
        run(1, flags=O.ORC_PRESAMPLING, env_pdf=epacked, env_dims=edims)     # environment samples without K11's tiles
    assert e.value.args[0] == -4
