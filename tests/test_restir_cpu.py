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


# ---- the ReSTIR-GI chain K4 -> K5 -> K8 (oracle/rt2_oracle_gi.inc) ---------------------------------------------
def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def test_logluv_codec_kat():
    """RTXDI_EncodeRGBToLogLuv / RTXDI_DecodeLogLuvToRGB (rtxdi/RtxdiMath.hlsli:233-296): 14-bit log2(Y) over [-20, 20) in
    steps of 1/409.6, 9-bit u and v scaled by 820; black is exact; luminance below 2^-20 decodes as zero."""
    L = O.lib()
    enc = lambda c: L.orc_encode_logluv(_vp(np.array(c, np.float32)))
    def dec(p):
        o = np.zeros(3, np.float32); L.orc_decode_logluv(p, _vp(o)); return o
    assert enc((0, 0, 0)) == 0 and (dec(0) == 0).all()
    assert enc((1e-7, 1e-7, 1e-7)) == 0                                   # Y < 2^-20 -> Le == 0
    assert enc((-1, -1, -1)) == 0                                         # log2 of a negative luminance is NaN -> 0 (pin P3)
    # white: Y = 1 -> Le = floor(409.6 * 20) = 8192; D65 chroma u' = 0.1978, v' = 0.4683 -> 162, 384
    p = enc((1, 1, 1))
    assert p >> 18 == 8192 and (p >> 9) & 0x1ff == 162 and p & 0x1ff == 384
    # doubling the radiance moves Le by 409.6 and leaves the chroma alone
    q = enc((2, 2, 2))
    assert (q >> 18) - (p >> 18) in (409, 410) and (q & 0x3ffff) == (p & 0x3ffff)
    rng = np.random.default_rng(11)
    for _ in range(300):
        c = (rng.random(3) ** 3 * 10.0 ** rng.uniform(-3, 3)).astype(np.float32)
        o = dec(enc(c))
        lum = lambda v: 0.2126 * v[0] + 0.7152 * v[1] + 0.0722 * v[2]
        tol = 0.002 if o.min() > 0 else 0.02                              # 0.17 % luminance steps; out-of-gamut chroma is clamped at 0 after decoding
        assert abs(lum(o) - lum(c)) <= tol * lum(c) + 1e-6
        if o.min() > 0:
            assert enc(o) >> 18 == enc(c) >> 18                           # decoding lands in the middle of the bucket: idempotent in Le
    assert enc((1e7, 1e7, 1e7)) >> 18 == 16383                            # clamps at the top of the range


def test_snorm_octahedral_normal_codec_kat():
    """RTXDI_EncodeNormalizedVectorToSnorm2x16 / Decode (rtxdi/RtxdiMath.hlsli:129-196)."""
    L = O.lib()
    enc = lambda n: L.orc_encode_snorm_oct(_vp(np.array(n, np.float32)))
    def dec(p):
        o = np.zeros(3, np.float32); L.orc_decode_snorm_oct(p, _vp(o)); return o
    assert enc((0, 0, 1)) == 0 and np.array_equal(dec(0), [0, 0, 1])
    assert enc((1, 0, 0)) == 0x00007fff and enc((0, 1, 0)) == 0x7fff0000
    assert enc((-1, 0, 0)) == 0x00008001 and enc((0, -1, 0)) == 0x80010000
    assert enc((0, 0, -1)) == 0x7fff7fff and np.array_equal(dec(0x7fff7fff), [0, 0, -1])
    assert enc((0, 0, 0)) == 0                                            # 0 * inf = NaN -> (0, 0), the empty reservoir's normal
    assert np.array_equal(dec(0x8000_8000), dec(0x8001_8001))             # -32768 clamps to -1
    rng = np.random.default_rng(5)
    v = rng.normal(size=(500, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True)
    for n in v.astype(np.float32):
        o = dec(enc(n))
        assert abs(np.linalg.norm(o) - 1) < 1e-6 and np.linalg.norm(o.astype(np.float64) - n) < 1e-4
        assert enc(o) == enc(n) or np.linalg.norm(dec(enc(o)).astype(np.float64) - o) < 1e-4


def _gi_frames(osc, g, w, h, gb, frames, light_inputs=None):
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    acc_d = np.zeros((h, w, 3)); acc_s = np.zeros((h, w, 3)); st = None
    for f in range(1, frames + 1):
        st = osc.gi_chain(g, f, w, h, gb, pitch, light_inputs=light_inputs)
        acc_d += st["diffuse"].view(np.float16)[:, :, :3].astype(np.float64)
        acc_s += st["specular"].view(np.float16)[:, :, :3].astype(np.float64)
    return acc_d / frames, acc_s / frames, st


def _albedo(gb, w, h):
    L = O.lib(); out = np.zeros((h, w, 3), np.float32); tmp = np.zeros(3, np.float32)
    for y in range(h):
        for x in range(w):
            L.orc_unpack_r11g11b10(int(gb["diffuse_albedo"][y, x]), _vp(tmp)); out[y, x] = tmp
    return out


def test_gi_default_frame_agrees_with_brute_force_path_tracer():
    """Default frame (DI off, GI on): K8's demodulated diffuse x albedo estimates Le(x1) + albedo(x1) Le(x2) seen from x0,
    i.e. K1 with 3 bounces minus K1 with 1 bounce (the emission of x0 itself) — two independent estimators."""
    w = h = 40
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    g.refrence_mode = 0
    assert g.enable_restir_gi == 1 and g.enable_restir_di == 0 and g.enable_brdf_indirect == 1     # src/main.rs:388-399
    d, s, st = _gi_frames(osc, g, w, h, gb, 96)
    hit = gb["depth"] != 100000.0
    r4, r5, r6, r7, r8 = st["rays"]
    assert 0 < r4 <= hit.sum() and r5 == w * h and r8 == 0               # K4 skips background + below-horizon; K5 traces every index; no final visibility
    gi = d * _albedo(gb, w, h)
    g.refrence_mode = 1
    k3 = np.zeros((h, w, 3)); k1 = np.zeros((h, w, 3))
    for f in range(1, 9):
        k3 += osc.render_reference(g, f, w, h, spp=64, max_bounces=3)[2][:, :, :3]
        k1 += osc.render_reference(g, f, w, h, spp=1, max_bounces=1)[2][:, :, :3]
    pt = (k3 - k1) / 8
    a, b = gi[hit].mean(), pt[hit].mean()
    assert b > 0 and abs(a - b) < 0.05 * b, (a, b)
    assert (d[~hit] == 0).all()                                           # background: no valid reservoir on a fresh buffer


def test_gi_reservoir_and_secondary_gbuffer_invariants():
    w = h = 48
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    g.refrence_mode = 0
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    row = g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch
    st = osc.gi_chain(g, 1, w, h, gb, pitch)
    L = O.lib()
    page = g.restir_gi.buffer_indices.secondary_surface_restirdi_output_buffer_index
    sec, res = st["secondary"], st["gi_reservoirs"]
    nvalid = 0
    for y in range(h):
        for x in range(w):
            s = sec[L.orc_reservoir_position_to_pointer(row, pitch, x, y, 0)]
            r = res[L.orc_reservoir_position_to_pointer(row, pitch, x, y, page)]
            if gb["depth"][y, x] == 100000.0:
                assert not s.any() and not r.any()                        # K4 returns early; K5 finds zero throughput -> empty reservoir
                continue
            flags = int(s[5]) >> 16
            assert flags & 2 == 0                                         # roughness == 1 everywhere (Hit.glsl:40): never a delta surface
            thr = np.array([s[4] & 0xffff, s[4] >> 16, s[5] & 0xffff], np.uint16).view(np.float16).astype(np.float32)
            M, age = int(r[3]) & 0xff, (int(r[3]) >> 8) & 0xff
            assert age == 0 and int(r[3]) >> 16 == 0 and r[7] == 0
            if (thr != 0).any():
                nvalid += 1
                assert M == 1
                pdf = s[11:12].view(np.float32)[0]
                wgt = r[5:6].view(np.float32)[0]
                assert pdf > 0 and wgt == np.float32(1.0) / pdf           # RTXDI_MakeGIReservoir: weightSum = 1 / pdf
                assert np.array_equal(r[0:3], s[0:3])                     # reservoir position == secondary world position
            else:
                assert M == 0 and not r.any()
    assert nvalid > (gb["depth"] != 100000.0).sum() * 0.8
    other = 1 - page
    assert not res[other * pitch:(other + 1) * pitch].any()              # the other reservoir page is untouched
    st2 = osc.gi_chain(g, 1, w, h, gb, pitch)
    assert all(np.array_equal(st[k], st2[k]) for k in ("secondary", "gi_reservoirs", "diffuse", "specular"))


def test_gi_with_light_sampling_at_the_secondary_hit():
    """DI on: K9 shades direct light at x0; K5 samples the light list at x1 (one BRDF ray, ShadeSurfaceWithLightSample).
    Sum of both demodulated outputs x albedo == K1 with 3 bounces minus emission at x0, up to the 1-sample estimators' noise."""
    w = h = 32
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    tasks, g2l, lights, pdf0 = osc.prepare_lights(pw.value, ph.value)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(pw.value, ph.value, pm.value)), np.float32)
    packed[:pw.value * ph.value] = pdf0.ravel()
    osc.generate_mips(0, packed, pw.value, ph.value, pm.value)
    li = {"lights": lights, "geom_to_light": g2l, "pdf_packed": packed, "pdf_dims": (pw.value, ph.value, pm.value)}
    g.refrence_mode = 0; g.enable_restir_di = 1
    g.restir_di.initial_sampling_params.num_primary_local_light_samples = 4
    g.restir_di.initial_sampling_params.num_primary_brdf_samples = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    acc = np.zeros((h, w, 3)); frames = 128
    for f in range(1, frames + 1):
        res, d16, s16, rays = osc.di_fused_resampling(g, f, w, h, gb, lights, g2l, packed, pw.value, ph.value, pm.value, g.restir_di.reservoir_buffer_params.reservoir_array_pitch)
        st = {"secondary": np.zeros((pitch, 12), np.uint32), "gi_reservoirs": np.zeros((pitch * 2, 8), np.uint32), "diffuse": d16, "specular": s16}
        st = osc.gi_chain(g, f, w, h, gb, pitch, light_inputs=li, state=st)
        acc += st["diffuse"].view(np.float16)[:, :, :3].astype(np.float64)      # K8 adds onto K9's image (isFirstPass false when DI is on)
    total = acc / frames * _albedo(gb, w, h)
    g.refrence_mode = 1
    k3 = np.zeros((h, w, 3)); k1 = np.zeros((h, w, 3))
    for f in range(1, 9):
        k3 += osc.render_reference(g, f, w, h, spp=64, max_bounces=3)[2][:, :, :3]
        k1 += osc.render_reference(g, f, w, h, spp=1, max_bounces=1)[2][:, :, :3]
    pt = (k3 - k1) / 8
    m = (gb["depth"] != 100000.0) & ~(gb["emissive"].view(np.float16)[:, :, 0] > 0)
    a, b = total[m].mean(), pt[m].mean()
    assert b > 0 and abs(a - b) < 0.08 * b, (a, b)
