"""CPU checks of the oracle's K3 (g_buffer) and K9 (di_fused_resampling) restatements.  The reference ships
no vectors, so the reservoir light sampling is cross-validated against an independent estimator
of the same quantity: direct lighting from the brute-force path tracer K1 (BRDF sampling, no light list)."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O


def _scene(w, h, kind=rt.SCENE_CORNELL, scale=1.0):
    sc = rt.Scene(kind, scale, 16, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    return sc, g, O.OracleScene(sc.desc)


def test_reservoir_addressing_kat():
    """RTXDI_ReservoirPositionToPointer (RtxdiHelpers.hlsli:75-88) + calculate_reservoir_buffer_parameters (src/light_passes.rs:718-731)."""
    L = O.lib()
    for (w, h) in [(1920, 1080), (512, 512), (17, 5)]:
        row = ((w + 15) // 16) * 256
        arr = row * ((h + 15) // 16)
        if w == 1920:
            assert arr == 2088960
        seen = set()
        pts = [(x, y) for (x, y) in [(0, 0), (15, 15), (16, 0), (0, 16), (w - 1, h - 1), (w // 2, h // 3)] if x < w and y < h]
        for (x, y) in pts:
            for idx in (0, 1):
                want = idx * arr + (y // 16) * row + (x // 16) * 256 + (y % 16) * 16 + (x % 16)
                got = L.orc_reservoir_position_to_pointer(row, arr, x, y, idx)
                assert got == want and got < 2 * arr
                seen.add(got)
        assert len(seen) == 2 * len(set(pts))


def test_gbuffer_matches_primary_visibility_and_materials():
    w = h = 96
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h, flags=O.ORC_BRUTE_FORCE)
    hits = osc.trace_primary(g, w, h, flags=O.ORC_BRUTE_FORCE)
    hit = hits[:, :, 1] != 0xFFFFFFFF
    assert hit.mean() > 0.3                            # the box is open toward the camera: border pixels miss
    assert np.array_equal(gb["depth"] != 100000.0, hit)
    assert np.array_equal(gb["depth"][hit], hits[:, :, 3].view(np.float32)[hit])      # TMin 0 vs 0.001: same closest hit here
    assert np.array_equal(gb["normals"], gb["geo_normals"])                 # g_buffer.rgen:33-34 writes the shading normal twice
    L = O.lib()
    d = sc.desc.contents
    n = np.zeros(3, np.float32); rgb = np.zeros(3, np.float32); sr = np.zeros(4, np.float32)
    picks = []
    for geom in np.unique(hits[:, :, 1][hit]):          # one pixel per visible geometry
        ys, xs = np.nonzero(hits[:, :, 1] == geom)
        picks.append((int(ys[len(ys) // 2]), int(xs[len(xs) // 2])))
    assert len(picks) >= 5
    for (y, x) in picks:
        geom = int(hits[y, x, 1])
        L.orc_oct_unorm32_to_ndir(int(gb["normals"][y, x]), n.ctypes.data_as(C.c_void_p))
        assert abs(np.linalg.norm(n) - 1) < 1e-5
        L.orc_unpack_r11g11b10(int(gb["diffuse_albedo"][y, x]), rgb.ctypes.data_as(C.c_void_p))
        base = np.array(list(d.geometries[geom].base_color)[:3], np.float32)
        assert np.allclose(rgb, base, atol=1.0 / 1023)
        L.orc_unpack_rgba8_gamma(int(gb["specular_rough"][y, x]), sr.ctypes.data_as(C.c_void_p))
        assert sr[3] == 1.0 and np.allclose(sr[:3], base * d.geometries[geom].metallic_factor, atol=0.02)   # roughness == 1 (Hit.glsl:40)
        em = gb["emissive"][y, x].view(np.float16)[:3].astype(np.float32)
        assert np.allclose(em, np.array(list(d.geometries[geom].emission)[:3]) * 12.0, rtol=1e-3)
    assert (gb["motion"] == 0).all()                   # prev_view == view: a static camera has zero motion


def test_gbuffer_miss_writes_background():
    w, h = 64, 32
    sc = rt.Scene(rt.SCENE_INSTANCED, 0.05, 16, 0)     # open scene under a sky
    cam = sc.camera(w / h)
    cam.direction[0], cam.direction[1], cam.direction[2] = 0.0, 1.0, 0.0     # look straight up
    g = rt.default_gconst(cam, w, h, sc.num_lights)
    gb = O.OracleScene(sc.desc).gbuffer(g, w, h)
    miss = gb["depth"] == 100000.0
    assert miss.any()
    for k in ("normals", "geo_normals", "diffuse_albedo", "specular_rough"):
        assert (gb[k][miss] == 0).all()
    assert (gb["emissive"][miss] == 0).all() and (gb["motion"][miss] == 0).all()


def _di(osc, sc, g, w, h, frame, gb, lights, g2l, pdf, dims):
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    return osc.di_fused_resampling(g, frame, w, h, gb, lights, g2l, pdf, dims[0], dims[1], dims[2], pitch)


def test_di_light_sampling_agrees_with_brute_force_path_tracer():
    """E[DI diffuse] * albedo == direct lighting == K1(2 bounces) - K1(1 bounce), two independent estimators
    (uniform light-list sampling + BRDF MIS through reservoirs vs cosine-sampled rays hitting emitters)."""
    w = h = 40
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    tasks, g2l, lights, pdf0 = osc.prepare_lights(pw.value, ph.value)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(pw.value, ph.value, pm.value)), np.float32)
    packed[:pw.value * ph.value] = pdf0.ravel()
    osc.generate_mips(0, packed, pw.value, ph.value, pm.value)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    acc = np.zeros((h, w, 3), np.float64)
    frames = 48
    for f in range(1, frames + 1):
        res, d16, s16, rays = _di(osc, sc, g, w, h, f, gb, lights, g2l, packed, (pw.value, ph.value, pm.value))
        acc += d16.view(np.float16)[:, :, :3].astype(np.float64)
        assert rays > 0
    L = O.lib()
    albedo = np.zeros((h, w, 3), np.float32); tmp = np.zeros(3, np.float32)
    for y in range(h):
        for x in range(w):
            L.orc_unpack_r11g11b10(int(gb["diffuse_albedo"][y, x]), tmp.ctypes.data_as(C.c_void_p)); albedo[y, x] = tmp
    direct_di = acc / frames * albedo
    g.refrence_mode = 1
    k2 = np.zeros((h, w, 3), np.float64); k1 = np.zeros((h, w, 3), np.float64)
    for f in range(1, 9):
        k2 += osc.render_reference(g, f, w, h, spp=64, max_bounces=2)[2][:, :, :3]
        k1 += osc.render_reference(g, f, w, h, spp=1, max_bounces=1)[2][:, :, :3]
    direct_pt = (k2 - k1) / 8
    emitter = gb["emissive"].view(np.float16)[:, :, 0] > 0            # pixels that see the lamp itself: DI shades them like any surface
    m = ~emitter
    a, b = direct_di[m].mean(), direct_pt[m].mean()
    assert b > 0 and abs(a - b) < 0.05 * b, (a, b)
    # per-region agreement (ceiling is dark in both, floor bright in both)
    rows = np.arange(h)[:, None] * np.ones((1, w), int)
    for lo, hi in [(0, h // 3), (h // 3, 2 * h // 3), (2 * h // 3, h)]:
        sel = m & (rows >= lo) & (rows < hi)
        ra, rb = direct_di[sel].mean(), direct_pt[sel].mean()
        assert abs(ra - rb) < 0.12 * max(rb, 0.02), (lo, ra, rb)


def test_di_reservoir_invariants():
    w = h = 48
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    tasks, g2l, lights, pdf0 = osc.prepare_lights(pw.value, ph.value)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(pw.value, ph.value, pm.value)), np.float32)
    packed[:pw.value * ph.value] = pdf0.ravel()
    osc.generate_mips(0, packed, pw.value, ph.value, pm.value)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = 2
    res, d16, s16, rays = _di(osc, sc, g, w, h, 1, gb, lights, g2l, packed, (pw.value, ph.value, pm.value))
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    L = O.lib()
    row = g.restir_di.reservoir_buffer_params.reservoir_block_row_pitch
    nvalid = 0
    for y in range(h):
        for x in range(w):
            r = res[L.orc_reservoir_position_to_pointer(row, pitch, x, y, g.restir_di.buffer_indices.shading_input_buffer_index)]
            assert (r[2] >> 18) == 1                                   # M == 1 after RTXDI_SampleLightsForSurface (:655)
            if r[0]:
                nvalid += 1
                assert r[0] & 0x80000000 and (r[0] & 0x7FFFFFFF) < sc.num_lights
                assert r[4:6].view(np.float32)[0] > 0 and np.isfinite(r[4:6].view(np.float32)).all()   # targetPdf > 0, finite weight
            else:
                assert r[5] == 0                                       # empty reservoir: weight 0
    assert nvalid > w * h // 4
    # determinism: same frame -> same bits; different frame -> different samples
    res2 = _di(osc, sc, g, w, h, 1, gb, lights, g2l, packed, (pw.value, ph.value, pm.value))[0]
    res3 = _di(osc, sc, g, w, h, 2, gb, lights, g2l, packed, (pw.value, ph.value, pm.value))[0]
    assert np.array_equal(res, res2) and not np.array_equal(res, res3)
    # the other reservoir page is untouched
    other = 1 - g.restir_di.buffer_indices.shading_input_buffer_index
    assert (res[other * pitch:(other + 1) * pitch] == 0).all()
