"""CPU tests of the oracle's K10 / K12 / K13 restatements against independent Python restatements written
here from the shader sources (the reference ships no vectors): packers, octahedral codecs,
triangle-light encode/decode, the light task scan, the mip reduction tree and the AgX tonemap."""
import ctypes as C
import math

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

f32 = np.float32
VP = C.c_void_p


def _p(a):
    return a.ctypes.data_as(VP)


# ---- packers (packing.glsl) ---------------------------------------------------------------------------
def py_pack_ufloat(r, size):
    mask = (1 << size) - 1
    return int(math.floor(float(f32(f32(r) * f32(mask)) + f32(0.5)))) & mask


def test_r11g11b10_kats_and_roundtrip():
    L = O.lib()
    rng = np.random.default_rng(1)
    for rgb in [(0, 0, 0), (1, 1, 1), (0.5, 0.25, 0.75), (1 / 3, 2 / 3, 0.1)] + [tuple(rng.random(3)) for _ in range(200)]:
        a = np.array(rgb, f32)
        want = py_pack_ufloat(a[0], 11) | (py_pack_ufloat(a[1], 11) << 11) | (py_pack_ufloat(a[2], 10) << 22)
        got = L.orc_pack_r11g11b10(_p(a))
        assert got == want
        back = np.zeros(3, f32); L.orc_unpack_r11g11b10(got, _p(back))
        assert abs(back[0] - a[0]) <= 0.5 / 2047 + 1e-6 and abs(back[1] - a[1]) <= 0.5 / 2047 + 1e-6 and abs(back[2] - a[2]) <= 0.5 / 1023 + 1e-6
    assert L.orc_pack_r11g11b10(_p(np.array([1, 1, 1], f32))) == 0xFFFFFFFF
    assert L.orc_pack_r11g11b10(_p(np.array([1, 0, 0], f32))) == 0x7FF


def test_rgba8_gamma_roundtrip():
    L = O.lib()
    for v in np.linspace(0, 1, 41):
        a = np.array([v, v * 0.5, 1 - v, 1.0], f32)
        p = L.orc_pack_rgba8_gamma(_p(a))
        want = [int(math.floor(float(f32(f32(float(x) ** (1 / 2.2)) * f32(255)) + f32(0.5)))) for x in a]
        got = [(p >> (8 * i)) & 255 for i in range(4)]
        assert all(abs(g - w) <= 1 for g, w in zip(got, want)), (got, want)     # pow rounding may move a tie
        back = np.zeros(4, f32); L.orc_unpack_rgba8_gamma(p, _p(back))
        assert np.allclose(back, a, atol=0.012)
    assert L.orc_pack_rgba8_gamma(_p(np.array([0, 0, 0, 0], f32))) == 0
    assert L.orc_pack_rgba8_gamma(_p(np.array([1, 1, 1, 1], f32))) == 0xFFFFFFFF
    assert L.orc_pack_rgba8_gamma(_p(np.array([2, -1, 1, 1], f32))) == 0xFFFF00FF   # clamped


def test_oct_normal_roundtrip_and_kats():
    L = O.lib()
    rng = np.random.default_rng(2)
    # +z maps to the centre of the unorm square: uint(0.5 * 65534) = 32767 on both axes
    assert L.orc_ndir_to_oct_unorm32(_p(np.array([0, 0, 1], f32))) == (32767 | (32767 << 16))
    assert L.orc_ndir_to_oct_unorm32(_p(np.array([1, 0, 0], f32))) == (65534 | (32767 << 16))
    assert L.orc_ndir_to_oct_unorm32(_p(np.array([0, 1, 0], f32))) == (32767 | (65534 << 16))
    worst = 0.0
    for _ in range(2000):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        if n[2] < 0 and (n[0] < 0) != (n[1] < 0):
            continue   # octWrap uses ONE sign for both lanes (Helpers.glsl:251-254): mixed-sign lower-hemisphere directions do not round-trip — literal
        a = n.astype(f32)
        p = L.orc_ndir_to_oct_unorm32(_p(a))
        back = np.zeros(3, f32); L.orc_oct_unorm32_to_ndir(p, _p(back))
        assert abs(np.linalg.norm(back) - 1) < 1e-5
        if n[2] >= 0 or (n[0] >= 0 and n[1] >= 0):
            worst = max(worst, float(np.linalg.norm(back.astype(np.float64) - n)))
    assert worst < 2e-4


def test_triangle_light_encode_decode():
    L = O.lib()
    base = np.array([0.5, 1.0, -2.0], f32); e1 = np.array([0.3, 0.0, 0.0], f32); e2 = np.array([0.0, 0.0, 0.4], f32)
    rad = np.array([12.0, 6.0, 3.0], f32)
    li = np.zeros(12, np.uint32); power = C.c_float(0)
    L.orc_store_triangle_light(_p(base), _p(e1), _p(e2), _p(rad), _p(li), C.byref(power))
    # logRadiance: min(uint(ceil(saturate((log2(12) + 8) / 48) * 65534)) + 1, 0xffff)   (PolymorphicLight.glsl:83-85)
    want_log = min(int(math.ceil(float(f32(f32((f32(math.log2(12.0)) - f32(-8)) / f32(48)) * f32(65534))))) + 1, 0xFFFF)
    assert li[7] == want_log
    assert (li[3] >> 24) == 4                       # kTriangle << 24
    assert np.allclose(li[:3].view(f32), base + (e1 + e2) / 3, atol=1e-6)
    b, a1, a2, r, n = (np.zeros(3, f32) for _ in range(5)); area = C.c_float(0)
    L.orc_create_triangle_light(_p(li), _p(b), _p(a1), _p(a2), _p(r), _p(n), C.byref(area))
    assert np.allclose(a1, e1, atol=2e-4) and np.allclose(a2, e2, atol=2e-4) and np.allclose(b, base, atol=2e-4)
    assert abs(area.value - 0.06) < 1e-4
    assert np.allclose(np.abs(n), [0, 1, 0], atol=1e-3)
    assert np.allclose(r, rad, rtol=0.01)           # RGB8 normalised colour x 16-bit log radiance
    lum = 0.299 * r[0] + 0.587 * r[1] + 0.114 * r[2]
    assert abs(power.value - area.value * math.pi * lum) < 1e-4 * power.value
    # zero radiance: colour and log-radiance stay 0, power 0
    L.orc_store_triangle_light(_p(base), _p(e1), _p(e2), _p(np.zeros(3, f32)), _p(li), C.byref(power))
    assert li[7] == 0 and (li[3] & 0xFFFFFF) == 0 and power.value == 0.0


# ---- K10 ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,scale", [(rt.SCENE_CORNELL, 1.0), (rt.SCENE_HALL, 0.125), (rt.SCENE_MANYLIGHT, 0.1)])
def test_prepare_lights_tasks_and_pdf(kind, scale):
    sc = rt.Scene(kind, scale, 16, 0)
    d = sc.desc.contents
    osc = O.OracleScene(sc.desc)
    w, h, mips = (C.c_uint32(), C.c_uint32(), C.c_uint32())
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(w), C.byref(h), C.byref(mips))
    tasks, g2l, lights, pdf = osc.prepare_lights(w.value, h.value)
    # src/prepare_lights.rs:182-201 restated
    off, want_tasks, want_g2l = 0, [], []
    for g in range(d.num_geometries):
        e = d.geometries[g].emission
        if e[0] == 0 and e[1] == 0 and e[2] == 0:
            want_g2l.append(0xFFFFFFFF); continue
        want_g2l.append(off); n = d.index_counts[g] // 3
        want_tasks.append((g, off, n, 0)); off += n
    assert off == sc.num_lights == len(lights) and off > 0
    assert [tuple(t) for t in tasks] == want_tasks and list(g2l) == want_g2l
    assert w.value * h.value >= off and (w.value & (w.value - 1)) == 0 and (h.value & (h.value - 1)) == 0
    # every light is a triangle light whose flux landed on its Z-curve texel; untouched texels stay 0
    assert ((lights[:, 3] >> 24) == 4).all()
    L = O.lib()
    xs = np.zeros(1, np.uint32); ys = np.zeros(1, np.uint32)
    total = 0.0
    for i in range(0, off, max(1, off // 64)):
        L.orc_linear_to_zcurve(i, xs.ctypes.data_as(C.POINTER(C.c_uint32)), ys.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert pdf[ys[0], xs[0]] > 0
    assert np.count_nonzero(pdf) == off
    assert np.isfinite(pdf).all()


# ---- K12 ----------------------------------------------------------------------------------------------
def np_mip_chain(m0, mips):
    """Independent restatement of mips.glsl's reduction tree for power-of-two sizes >= 32: the first level
    of every 5-level pass sums the quad as ((p00 + p01) + p10) + p11 with p01 = (x, y+1) (mips.glsl:95-103),
    the other four levels follow the Z-curve lane order ((p00 + p10) + p01) + p11 (:116-184)."""
    out = [m0.astype(f32)]
    for k in range(1, mips):
        s = out[-1]
        p00, p10, p01, p11 = s[0::2, 0::2], s[0::2, 1::2], s[1::2, 0::2], s[1::2, 1::2]
        if (k - 1) % 5 == 0:
            v = (((p00 + p01) + p10) + p11) * f32(0.25)
        else:
            v = (((p00 + p10) + p01) + p11) * f32(0.25)
        out.append(v.astype(f32))
    return out


@pytest.mark.parametrize("w,h", [(32, 32), (128, 128), (256, 128), (64, 256)])
def test_local_mip_chain_matches_numpy_tree(w, h):
    mips = int(math.log2(max(w, h))) + 1
    rng = np.random.default_rng(w * 7 + h)
    m0 = rng.random((h, w), dtype=f32) * f32(100)
    packed = np.zeros(sum(mw * mh for mw, mh in O.mip_dims(w, h, mips)), f32)
    packed[:w * h] = m0.ravel()
    sc = rt.Scene(rt.SCENE_CORNELL)
    O.OracleScene(sc.desc).generate_mips(0, packed, w, h, mips)
    got = O.split_mips(packed, w, h, mips)
    ref = np_mip_chain(m0, mips)
    square_levels = int(math.log2(min(w, h))) + 1      # below that one dimension is clamped to 1: out-of-range loads give 0
    for k in range(square_levels):
        assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), f"mip {k}"
    if w == h:
        assert abs(float(got[-1][0, 0]) - float(m0.astype(np.float64).mean())) < 1e-3 * float(m0.mean())


def test_mip_chain_small_and_ragged():
    sc = rt.Scene(rt.SCENE_CORNELL)
    osc = O.OracleScene(sc.desc)
    one = np.array([3.5], f32)
    assert osc.generate_mips(0, one.copy(), 1, 1, 1)[0] == f32(3.5)       # a single level: nothing to do
    packed = np.array([2.0, 4.0, 0.0], f32)                               # 2x1 -> 1x1: the missing row reads as 0 (literal)
    assert osc.generate_mips(0, packed, 2, 1, 2)[2] == f32(1.5)


def test_env_mip_chain_properties():
    sc = rt.Scene(rt.SCENE_INSTANCED, 0.05, 16, 0)
    d = sc.desc.contents
    w, h = d.sky_width, d.sky_height
    assert w and h
    mips = int(math.ceil(math.log2(max(w, h)))) + 1
    packed = np.zeros(sum(mw * mh for mw, mh in O.mip_dims(w, h, mips)), np.uint16)
    O.OracleScene(sc.desc).generate_mips(1, packed, w, h, mips)
    lv = [m.view(np.float16).astype(np.float64) for m in O.split_mips(packed, w, h, mips)]
    assert (lv[0] >= 0).all() and lv[0].max() <= 65504 and lv[0].max() > 0
    sky = np.ctypeslib.as_array(d.sky_rgba32f, shape=(h, w, 4)).astype(np.float64)
    # mip 0 texel (x,y) = lum(2x2 box ending at (x,y)) * cos(elevation(y)); check an interior texel
    x, y = w // 3, h // 3
    box = sky[y - 1:y + 1, x - 1:x + 1, :3].mean(axis=(0, 1))
    want = (0.299 * box[0] + 0.587 * box[1] + 0.114 * box[2]) * math.cos(((y + 0.5) / h - 0.5) * math.pi)
    assert abs(lv[0][y, x] - want) <= 2e-3 * max(want, 1e-3) + 1e-4
    # each level is the 2x2 box of the one below (registers stay fp32 inside a 5-level pass, storage is fp16)
    for k in range(1, 4):
        yy, xx = lv[k].shape[0] // 2, lv[k].shape[1] // 2
        want = lv[k - 1][2 * yy:2 * yy + 2, 2 * xx:2 * xx + 2].mean()
        assert abs(lv[k][yy, xx] - want) <= 4e-3 * want + 1e-5, k


# ---- K13 ----------------------------------------------------------------------------------------------
def py_tonemap(rgb):
    """post_processing.comp:47-124,180-190 in double precision."""
    m = np.array([[0.842479062253094, 0.0423282422610123, 0.0423756549057051], [0.0784335999999992, 0.878468636469772, 0.0784336],
                  [0.0792237451477643, 0.0791661274605434, 0.879142973793104]]).T
    mi = np.array([[1.19687900512017, -0.0528968517574562, -0.0529716355144438], [-0.0980208811401368, 1.15190312990417, -0.0980434501171241],
                   [-0.0990297440797205, -0.0989611768448433, 1.15107367264116]]).T
    v = m @ np.asarray(rgb, np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        v = np.clip(np.log2(v), -12.47393, 4.026069)
    v = (v + 12.47393) / (4.026069 + 12.47393)
    x2 = v * v; x4 = x2 * x2
    v = 15.5 * x4 * x2 - 40.14 * x4 * v + 31.96 * x4 - 6.868 * x2 * v + 0.4298 * x2 + 0.1191 * v - 0.00232
    if (v < 0).any():
        v = np.full(3, np.nan)          # pow(x < 0, 1.0) pinned to NaN; luma + (NaN - luma) poisons every channel
    v = mi @ v
    v = np.where(v < 0, np.nan, np.power(np.abs(v), 2.2))
    v = np.where(np.isnan(v), 1e-6, np.maximum(v, 1e-6))
    return [int(math.floor(min(max(c, 0), 1) * 255 + 0.5)) for c in v]


def test_tonemap_matches_double_precision_restatement():
    L = O.lib()
    rng = np.random.default_rng(5)
    cases = [(0, 0, 0), (1, 1, 1), (0.18, 0.18, 0.18), (12, 12, 12), (1e-5, 0, 0), (100, 0.1, 0.001), (0.5, 0.2, 0.9)]
    cases += [tuple(10 ** rng.uniform(-4, 2, 3)) for _ in range(300)]
    for c in cases:
        p = L.orc_tonemap(_p(np.array(c, f32)))
        got = [(p >> (8 * i)) & 255 for i in range(3)]
        want = py_tonemap(np.array(c, f32).astype(np.float64))
        assert (p >> 24) == 255
        assert all(abs(g - w) <= 1 for g, w in zip(got, want)), (c, got, want)
    p = L.orc_tonemap(_p(np.array([0, 0, 0], f32)))
    assert p == 0xFF000000                              # black stays black (NaN path -> max(.., 1e-6) -> 0)
    hi = L.orc_tonemap(_p(np.array([1000, 1000, 1000], f32)))
    assert (hi & 0xFF) >= 200                           # saturates toward white
    # monotone in exposure on the grey axis
    prev = -1
    for e in np.linspace(-10, 4, 60):
        v = L.orc_tonemap(_p(np.array([2.0 ** e] * 3, f32))) & 0xFF
        assert v >= prev; prev = v


def test_postprocess_reference_mode_reads_diffuse_only():
    w, h = 16, 8
    sc = rt.Scene(rt.SCENE_CORNELL)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    rng = np.random.default_rng(3)
    d = np.zeros((h, w, 4), np.float16); d[:, :, :3] = rng.random((h, w, 3)) * 4
    out, _ = O.OracleScene(sc.desc).postprocess(g, w, h, d.view(np.uint16))
    L = O.lib()
    for (y, x) in [(0, 0), (3, 7), (7, 15)]:
        p = L.orc_tonemap(_p(d[y, x, :3].astype(f32)))
        assert list(out[y, x]) == [(p >> (8 * i)) & 255 for i in range(4)]
