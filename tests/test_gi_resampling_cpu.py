"""CPU checks of the oracle's K6 (GI temporal resampling) and K7 (GI spatial resampling) restatements
(oracle/rt2_oracle_gi.inc; lighting_passes/temporal_resampling.rgen, spatial_resampling.rgen,
rtxdi/GIResamplingFunctions.hlsli:25-553).  No reference vectors exist: the resampled estimator is
cross-validated against the brute-force path tracer, and the reservoir bookkeeping against its invariants."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O


def _scene(w, h):
    sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 16, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 0
    return sc, g, O.OracleScene(sc.desc)


def _albedo(gb, w, h):
    L = O.lib(); out = np.zeros((h, w, 3), np.float32); tmp = np.zeros(3, np.float32)
    for y in range(h):
        for x in range(w):
            L.orc_unpack_r11g11b10(int(gb["diffuse_albedo"][y, x]), tmp.ctypes.data_as(C.c_void_p)); out[y, x] = tmp
    return out


def test_neighbor_offset_table_kat():
    """fill_neighbor_offset_buffer (src/light_passes.rs:671-698): R2 sequence in f32, rejected outside the unit disc,
    scaled by 250 and truncated; bytes pinned to signed."""
    tab = O.fill_neighbor_offset_buffer(8192)
    assert tab.shape == (8192, 2) and tab.dtype == np.int8
    # independent restatement in numpy float32
    phi2 = np.float32(1.0) / np.float32(1.3247179572447)
    u = np.float32(0.5); v = np.float32(0.5); want = []
    while len(want) < 64:
        u = np.float32(u + phi2); v = np.float32(v + np.float32(phi2 * phi2))
        if u >= 1: u = np.float32(u - np.float32(1))
        if v >= 1: v = np.float32(v - np.float32(1))
        du, dv = np.float32(u - np.float32(0.5)), np.float32(v - np.float32(0.5))
        if np.float32(np.float32(du * du) + np.float32(dv * dv)) > np.float32(0.25):
            continue
        want.append((int(np.float32(du * np.float32(250))), int(np.float32(dv * np.float32(250)))))
    assert [tuple(int(c) for c in t) for t in tab[:64]] == want
    r = np.hypot(tab[:, 0].astype(float), tab[:, 1].astype(float))
    assert r.max() <= 125.0 and (tab < 0).any() and (tab > 0).any()
    assert abs(tab.astype(float).mean()) < 1.5                    # centred on the pixel
    q = [((tab[:, 0] >= 0) == sx) & ((tab[:, 1] >= 0) == sy) for sx in (0, 1) for sy in (0, 1)]
    assert all(1700 < m.sum() < 2400 for m in q)                  # low-discrepancy: the four quadrants are evenly filled


def _frames(osc, g, w, h, gb, n, offsets, start=1):
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    motion = np.zeros((h, w, 4), np.float32)
    st = None; acc = np.zeros((h, w, 3)); per_frame = []
    for f in range(start, start + n):
        st = osc.gi_chain(g, f, w, h, gb, pitch, state=st, gb_prev=gb, motion=motion, neighbor_offsets=offsets)
        st = {k: v for k, v in st.items()}
        acc += st["diffuse"].view(np.float16)[:, :, :3].astype(np.float64)
        per_frame.append(st["diffuse"].view(np.float16)[:, :, :3].astype(np.float64).copy())
    return acc / n, st, per_frame


def _path_traced_two_bounce(osc, g, w, h):
    g.refrence_mode = 1
    k3 = np.zeros((h, w, 3)); k1 = np.zeros((h, w, 3))
    for f in range(1, 9):
        k3 += osc.render_reference(g, f, w, h, spp=64, max_bounces=3)[2][:, :, :3]
        k1 += osc.render_reference(g, f, w, h, spp=1, max_bounces=1)[2][:, :, :3]
    g.refrence_mode = 0
    return (k3 - k1) / 8


def test_spatiotemporal_resampling_converges_to_the_path_tracer_and_reduces_variance():
    """Static camera, temporal + spatial reuse on (history needs both: K7 writes the page K6 reads next frame, src/main.rs:241-247).
    ReSTIR GI with basic bias correction and no final visibility stays close to the two-bounce estimate, with far less
    frame-to-frame variance than the unresampled chain."""
    w = h = 40
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    offsets = O.fill_neighbor_offset_buffer(8192)
    hit = gb["depth"] != 100000.0
    alb = _albedo(gb, w, h)
    base_mean, _, base_frames = _frames(osc, g, w, h, gb, 48, offsets)
    g.enable_temporal_resampling = 1; g.enable_spatial_resampling = 1
    res_mean, st, res_frames = _frames(osc, g, w, h, gb, 48, offsets)
    pt = _path_traced_two_bounce(osc, g, w, h)
    a, b, c = (res_mean * alb)[hit].mean(), (base_mean * alb)[hit].mean(), pt[hit].mean()
    assert abs(b - c) < 0.06 * c, (b, c)
    assert abs(a - c) < 0.15 * c, (a, c)                          # reuse without visibility is biased, but mildly in an open box
    var_base = np.var(np.stack(base_frames[8:]), axis=0)[hit].mean()
    var_res = np.var(np.stack(res_frames[8:]), axis=0)[hit].mean()
    assert var_res < 0.5 * var_base, (var_res, var_base)
    r4, r5, r6, r7, r8 = st["rays"]
    assert r6 == 0 and r7 == 0                                    # bias correction mode 2 (< RAY_TRACED): no visibility rays


def test_reservoir_bookkeeping_over_frames():
    w = h = 32
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    offsets = O.fill_neighbor_offset_buffer(8192)
    g.enable_temporal_resampling = 1; g.enable_spatial_resampling = 1
    g.restir_gi.temporal_resampling_params.max_history_length = 6
    g.restir_gi.temporal_resampling_params.max_reservoir_age = 40
    g.restir_gi.spatial_resampling_params.num_spatial_samples = 2
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    row = g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch
    L = O.lib()
    motion = np.zeros((h, w, 4), np.float32)
    st = None
    hit = gb["depth"] != 100000.0
    ptr = np.array([[L.orc_reservoir_position_to_pointer(row, pitch, x, y, 0) for x in range(w)] for y in range(h)])
    maxM_t = []
    for f in range(1, 9):
        st = osc.gi_chain(g, f, w, h, gb, pitch, state=st, gb_prev=gb, motion=motion, neighbor_offsets=offsets)
        res = st["gi_reservoirs"]
        t_out = res[ptr]                    # page 0: temporal output
        s_out = res[pitch + ptr]            # page 1: spatial output
        Mt, Ms = t_out[:, :, 3] & 0xff, s_out[:, :, 3] & 0xff
        age_t = (t_out[:, :, 3] >> 8) & 0xff
        assert (Mt[~hit] == 0).all() and (Ms[~hit] == 0).all()
        # temporal: M <= own sample + clamped history; spatial adds at most numSpatialSamples neighbours of that size
        assert Mt.max() <= 1 + 6 and Ms.max() <= (1 + 6) * 3
        assert age_t.max() <= f - 1                                # a sample can only be as old as the history
        w_t = t_out[:, :, 5].view(np.float32)
        assert np.isfinite(w_t).all() and (w_t >= 0).all()
        assert ((Mt > 0) | (w_t == 0)).all()
        maxM_t.append(int(Mt.max()))
    assert maxM_t[0] == 1 and maxM_t[1] > 1 and maxM_t[-1] == 7    # frame 1 has no history; then it builds up to the clamp
    # determinism
    st2 = None
    for f in range(1, 9):
        st2 = osc.gi_chain(g, f, w, h, gb, pitch, state=st2, gb_prev=gb, motion=motion, neighbor_offsets=offsets)
    assert np.array_equal(st["gi_reservoirs"], st2["gi_reservoirs"]) and np.array_equal(st["diffuse"], st2["diffuse"])


def test_ray_traced_bias_correction_traces_visibility_rays():
    w = h = 32
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    offsets = O.fill_neighbor_offset_buffer(8192)
    g.enable_temporal_resampling = 1; g.enable_spatial_resampling = 1
    g.restir_gi.temporal_resampling_params.temporal_bias_correction_mode = 3
    g.restir_gi.spatial_resampling_params.spatial_bias_correction_mode = 3
    g.restir_gi.spatial_resampling_params.num_spatial_samples = 3
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    motion = np.zeros((h, w, 4), np.float32)
    st = None
    for f in range(1, 4):
        st = osc.gi_chain(g, f, w, h, gb, pitch, state=st, gb_prev=gb, motion=motion, neighbor_offsets=offsets)
    r4, r5, r6, r7, r8 = st["rays"]
    nhit = int((gb["depth"] != 100000.0).sum())
    assert 0 < r6 <= nhit and 0 < r7 <= 3 * nhit


def test_temporal_reprojection_follows_the_motion_vectors():
    """A history shifted by (+3, 0) pixels is found again when the motion vectors say so, and is not without them."""
    w = h = 32
    sc, g, osc = _scene(w, h)
    gb = osc.gbuffer(g, w, h)
    offsets = O.fill_neighbor_offset_buffer(8192)
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    row = g.restir_gi.reservoir_buffer_params.reservoir_block_row_pitch
    L = O.lib()
    ptr = np.array([[L.orc_reservoir_position_to_pointer(row, pitch, x, y, 0) for x in range(w)] for y in range(h)])
    st = osc.gi_chain(g, 1, w, h, gb, pitch)                       # plain chain: page 0 holds fresh samples
    fresh = st["gi_reservoirs"][ptr].copy()
    # previous frame = this frame shifted right by 3 pixels (G-buffer and history alike)
    shift = 3
    gb_prev = {k: np.roll(v, shift, axis=1) for k, v in gb.items() if k in ("depth", "normals", "geo_normals", "diffuse_albedo", "specular_rough")}
    for k in gb_prev:
        gb_prev[k][:, :shift] = 100000.0 if k == "depth" else 0
    hist = np.zeros_like(st["gi_reservoirs"])
    rolled = np.roll(fresh, shift, axis=1); rolled[:, :shift] = 0
    hist[pitch + ptr] = rolled                                     # page 1 = temporal input
    g.enable_temporal_resampling = 1
    g.restir_gi.temporal_resampling_params.enable_fallback_sampling = 0
    found = {}
    for name, mv in (("with", float(shift)), ("without", 0.0)):
        motion = np.zeros((h, w, 4), np.float32); motion[:, :, 0] = mv
        res = hist.copy(); res[ptr] = fresh                        # page 0 = K5 output of "this" frame
        osc.gi_temporal_resampling(g, 2, w, h, gb, gb_prev, motion, res)
        M = res[ptr][:, :, 3] & 0xff
        found[name] = int((M == 2).sum())
    hit = gb["depth"] != 100000.0
    # prevView == view here, so a shifted G-buffer reprojects to slightly different world positions: the Jacobian test drops some
    assert found["with"] > 0.5 * hit[:, :w - shift].sum()
    assert found["with"] > 1.5 * found["without"]
