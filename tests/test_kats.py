"""Known-answer tests of the oracle's pure functions against independent pure-Python restatements
written here straight from the shader sources (the reference ships no vectors of its own),
plus the host-side helpers of src/light_passes.rs:671-731 against hand-evaluated values."""
import ctypes as C
import os
import math
import struct

import numpy as np
import pytest

import raytracer2_b200 as rt
from raytracer2_b200 import _ffi
from oracle import oracle_ffi as O

M32 = 0xFFFFFFFF


def py_explode(x):
    x = (x | (x << 8)) & 0x00FF00FF
    x = (x | (x << 4)) & 0x0F0F0F0F
    x = (x | (x << 2)) & 0x33333333
    x = (x | (x << 1)) & 0x55555555
    return x


def py_jenkins(a):  # rtxdi/RtxdiMath.hlsli:69-79
    a = ((a + 0x7ed55d16) + (a << 12)) & M32
    a = ((a ^ 0xc761c23c) ^ (a >> 19)) & M32
    a = ((a + 0x165667b1) + (a << 5)) & M32
    a = ((a + 0xd3a2646c) ^ (a << 9)) & M32
    a = ((a + 0xfd7046c5) + (a << 3)) & M32
    a = ((a ^ 0xb55a4f09) ^ (a >> 16)) & M32
    return a


def py_murmur(seed, index):  # Helpers.glsl:25-55
    rot = lambda x, y: ((x << y) | (x >> (32 - y))) & M32
    k = (index * 0xcc9e2d51) & M32
    k = rot(k, 15)
    k = (k * 0x1b873593) & M32
    h = seed ^ k
    h = (rot(h, 13) * 5 + 0xe6546b64) & M32
    h ^= 4
    h ^= h >> 16
    h = (h * 0x85ebca6b) & M32
    h ^= h >> 13
    h = (h * 0xc2b2ae35) & M32
    h ^= h >> 16
    return h


def test_zcurve_known_values():
    L = O.lib()
    # bit-interleave by hand: x=0b11 y=0b00 -> 0b0101 ; x=0b00,y=0b11 -> 0b1010 ; (5,3) -> x=101,y=011 -> 0b011011 = 27
    assert L.orc_zcurve_to_linear(3, 0) == 0b0101
    assert L.orc_zcurve_to_linear(0, 3) == 0b1010
    assert L.orc_zcurve_to_linear(5, 3) == 0b011011
    assert L.orc_zcurve_to_linear(0xFFFF, 0xFFFF) == 0xFFFFFFFF
    rng = np.random.default_rng(1)
    for x, y in rng.integers(0, 65536, size=(200, 2)):
        lin = L.orc_zcurve_to_linear(int(x), int(y))
        assert lin == (py_explode(int(x)) | (py_explode(int(y)) << 1))
        ox, oy = C.c_uint32(), C.c_uint32()
        L.orc_linear_to_zcurve(lin, C.byref(ox), C.byref(oy))
        assert (ox.value, oy.value) == (x, y)


def test_jenkins_and_rng_stream():
    L = O.lib()
    assert L.orc_jenkins_hash(0) == py_jenkins(0)
    for a in (1, 2, 0xdeadbeef, 0xFFFFFFFF, 12345):
        assert L.orc_jenkins_hash(a) == py_jenkins(a)
    for (px, py, frame, ps) in ((0, 0, 1, 1), (17, 33, 1, 1), (1919, 1079, 7, 1), (100, 200, 123456, 2)):
        n = 16
        raw = np.zeros(n, np.uint32)
        uni = np.zeros(n, np.float32)
        L.orc_rng_stream(px, py, frame, ps, n, raw.ctypes.data_as(C.c_void_p), uni.ctypes.data_as(C.c_void_p))
        seed = (py_jenkins(py_explode(px) | (py_explode(py) << 1)) + frame + ps * 13) & M32  # bridge :378-381, Helpers.glsl:13-23
        for i in range(n):
            h = py_murmur(seed, 1 + i)
            assert raw[i] == h
            f = struct.unpack("<f", struct.pack("<I", (h & 0x7FFFFF) | 0x3F800000))[0] - 1.0  # Helpers.glsl:57-63
            assert uni[i] == np.float32(f) and 0.0 <= uni[i] < 1.0


def test_f16_matches_numpy():
    L = O.lib()
    rng = np.random.default_rng(2)
    vals = np.concatenate([rng.standard_normal(2000).astype(np.float32) * 10,
                           np.array([0, -0.0, 1, 65504, 65519.9, 65520, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8, 6.1e-5, 1e9, -1e9, np.inf, -np.inf], np.float32),
                           (rng.random(2000).astype(np.float32) * 1e-4)])
    for v in vals:
        got = L.orc_f32_to_f16(float(v))
        with np.errstate(over="ignore"):
            want = int(np.float32(v).astype(np.float16).view(np.uint16))
        assert got == want, (v, hex(got), hex(want))
    for h in list(range(0, 0x7C01, 97)) + [0x8000, 0x8001, 0xFBFF, 0x0001, 0x03FF, 0x0400]:
        assert L.orc_f16_to_f32(h) == float(np.uint16(h).view(np.float16).astype(np.float32))


def test_pinned_sincos_accuracy():
    L = O.lib()
    s, c = C.c_float(), C.c_float()
    worst = 0.0
    for a in np.linspace(0, 2 * math.pi, 4001, dtype=np.float32):
        L.orc_sincos(float(a), 0, C.byref(s), C.byref(c))
        worst = max(worst, abs(s.value - math.sin(float(a))), abs(c.value - math.cos(float(a))))
    assert worst < 2e-7  # Vulkan asks 2^-11 absolute; the pinned polynomial is ~1 ulp


def test_ray_triangle_basics():
    L = O.lib()
    f3 = lambda *v: (C.c_float * 3)(*v)
    t, u, v = C.c_float(), C.c_float(), C.c_float()
    hit = L.orc_ray_triangle(f3(0.25, 0.25, 1), f3(0, 0, -1), 0.001, 1e5, f3(0, 0, 0), f3(1, 0, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v))
    assert hit == 1 and t.value == 1.0 and u.value == 0.25 and v.value == 0.25
    # back face counts (TRIANGLE_CULL_DISABLE, src/model.rs:67-70)
    assert L.orc_ray_triangle(f3(0.25, 0.25, -1), f3(0, 0, 1), 0.001, 1e5, f3(0, 0, 0), f3(1, 0, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v)) == 1
    # outside, behind tmin, beyond tmax
    assert L.orc_ray_triangle(f3(2, 2, 1), f3(0, 0, -1), 0.001, 1e5, f3(0, 0, 0), f3(1, 0, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v)) == 0
    assert L.orc_ray_triangle(f3(0.25, 0.25, 0.0005), f3(0, 0, -1), 0.001, 1e5, f3(0, 0, 0), f3(1, 0, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v)) == 0
    assert L.orc_ray_triangle(f3(0.25, 0.25, 1), f3(0, 0, -1), 0.001, 0.5, f3(0, 0, 0), f3(1, 0, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v)) == 0
    # shared edge: a ray through the diagonal of a quad must hit at least one of the two triangles
    rng = np.random.default_rng(3)
    for _ in range(500):
        s = float(rng.random())
        o = f3(s, s, 1.0)
        d = f3(0, 0, -1)
        h1 = L.orc_ray_triangle(o, d, 0.001, 1e5, f3(0, 0, 0), f3(1, 0, 0), f3(1, 1, 0), C.byref(t), C.byref(u), C.byref(v))
        h2 = L.orc_ray_triangle(o, d, 0.001, 1e5, f3(0, 0, 0), f3(1, 1, 0), f3(0, 1, 0), C.byref(t), C.byref(u), C.byref(v))
        assert h1 or h2


def test_light_pass_helpers_hand_values():
    H = _ffi.host_lib()
    w, h, m = C.c_uint32(), C.c_uint32(), C.c_uint32()
    # light_passes.rs:700-716 by hand: 10000 -> ceil(sqrt)=100 -> 128; ceil(10000/128)=79 -> 128; mips = log2(128)+1 = 8
    for items, want in ((10000, (128, 128, 8)), (2, (2, 1, 2)), (1, (1, 1, 1)), (0, (1, 1, 1)), (5, (4, 2, 3)), (1 << 20, (1024, 1024, 11)), (129, (16, 16, 5))):
        H.rt2h_compute_pdf_texture_size(items, C.byref(w), C.byref(h), C.byref(m))
        assert (w.value, h.value, m.value) == want, items
    p = _ffi.ReservoirBufferParams()
    # light_passes.rs:718-731: 1920x1080 -> 120 x 68 blocks of 256
    H.rt2h_calculate_reservoir_buffer_parameters(1920, 1080, C.byref(p))
    assert (p.reservoir_block_row_pitch, p.reservoir_array_pitch) == (120 * 256, 120 * 256 * 68)
    H.rt2h_calculate_reservoir_buffer_parameters(512, 512, C.byref(p))
    assert (p.reservoir_block_row_pitch, p.reservoir_array_pitch) == (32 * 256, 32 * 256 * 32)
    H.rt2h_calculate_reservoir_buffer_parameters(17, 1, C.byref(p))
    assert (p.reservoir_block_row_pitch, p.reservoir_array_pitch) == (512, 512)
    # light_passes.rs:671-698 restated in Python (f32 arithmetic, saturating `as u8`)
    n = 64
    buf = (C.c_uint8 * (2 * n))()
    assert H.rt2h_fill_neighbor_offset_buffer(n, buf) == 2 * n
    f = np.float32
    phi2 = f(1.0 / 1.3247179572447)
    u = v = f(0.5)
    want = []
    while len(want) < 2 * n:
        u = f(u + phi2)
        v = f(v + f(phi2 * phi2))
        if u >= 1:
            u = f(u - f(1))
        if v >= 1:
            v = f(v - f(1))
        r = f(f(f(u - f(0.5)) * f(u - f(0.5))) + f(f(v - f(0.5)) * f(v - f(0.5))))
        if r > f(0.25):
            continue
        for val in (f(f(u - f(0.5)) * f(250)), f(f(v - f(0.5)) * f(250))):
            want.append(0 if not val > 0 else min(255, int(val)))
    assert list(buf) == want


def test_gconst_layout_and_defaults():
    # src/shader_params.rs summed field by field: PlanarViewConstants 464, GConst 1424
    assert C.sizeof(_ffi.GConst) == 1424
    assert _ffi.GConst.prev_view.offset == 464 and _ffi.GConst.runtime_params.offset == 928
    assert _ffi.GConst.restir_gi.offset == 960 and _ffi.GConst.restir_di.offset == 1104
    assert _ffi.GConst.enable_restir_di.offset == 1296 and _ffi.GConst.blend_factor.offset == 1312
    assert _ffi.GConst.light_buffer_params.offset == 1328 and _ffi.GConst.environment_pdf_texture_size.offset == 1408
    sc = rt.Scene(rt.SCENE_CORNELL)
    g = rt.default_gconst(sc.camera(16 / 9), 1920, 1080, sc.num_lights)
    # src/main.rs:237-400
    assert g.enable_restir_gi == 1 and g.enable_restir_di == 0 and g.refrence_mode == 0 and g.textures == 1
    assert abs(g.blend_factor - 0.1) < 1e-7 and g.runtime_params.neighbor_offset_mask == 8191
    assert g.restir_di.initial_sampling_params.num_primary_brdf_samples == 1
    assert g.light_buffer_params.local_light_buffer_region.num_lights == 2
    assert g.light_buffer_params.environment_light_params.light_index == 3
    assert g.restir_gi.reservoir_buffer_params.reservoir_array_pitch == 120 * 256 * 68
    assert tuple(g.view.viewport_size) == (1920.0, 1080.0)


def test_camera_matrices_are_consistent():
    sc = rt.Scene(rt.SCENE_CORNELL)
    cam = sc.camera(1.0)
    v = rt.planar_view_constants(cam, 512, 512)
    m = lambda a: np.array(list(a), np.float64).reshape(4, 4).T  # column-major -> numpy
    assert np.allclose(m(v.mat_world_to_view) @ m(v.mat_view_to_world), np.eye(4), atol=1e-5)
    assert np.allclose(m(v.mat_view_to_clip) @ m(v.mat_clip_to_view), np.eye(4), atol=1e-4)
    # perspective(): camera.rs:146-157 puts 1.0 at column 2 row 3 (w = +z_view)
    P = m(v.mat_view_to_clip)
    assert P[3, 2] == 1.0 and P[3, 3] == 0.0
    # centre ray leaves along -direction (view +z is -f for look_at_rh): Cornell camera looks down -z
    o = (C.c_float * 3)()
    d = (C.c_float * 3)()
    g = rt.default_gconst(cam, 512, 512, 2)
    O.lib().orc_primary_ray(C.cast(C.byref(g), C.c_void_p), 256, 256, o, d)
    assert tuple(o) == (0.0, 0.0, pytest.approx(3.4))
    assert d[2] < -0.99


def test_rt2_sys_declares_the_abi():
    """bindings/rt2-sys/src/lib.rs (the Rust -sys crate of SURVEY §8f N4) cannot be compiled here (no cargo): check it against
    include/rt2_abi.h textually — same function set, same argument counts, same buffer ids, flags and struct sizes."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "rt2_abi.h")).read()
    rs = open(os.path.join(root, "bindings", "rt2-sys", "src", "lib.rs")).read()
    code = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    c_fns = {m.group(1): (0 if m.group(2).strip() == "void" else m.group(2).count(",") + 1)
             for m in re.finditer(r"\b(rt2_[a-z_0-9]+)\s*\(([^)]*)\)\s*;", code)}
    rs_fns = {m.group(1): (0 if not m.group(2).strip() else m.group(2).count(":"))
              for m in re.finditer(r"pub fn (rt2_[a-z_0-9]+)\(([^)]*)\)", rs)}
    assert c_fns == rs_fns, set(c_fns.items()) ^ set(rs_fns.items())
    c_bufs = {m.group(1): int(m.group(2)) for m in re.finditer(r"\b(RT2_BUF_[A-Z0-9_]+)\s*=\s*(\d+)", code)}
    rs_bufs = {m.group(1): int(m.group(2)) for m in re.finditer(r"pub const (RT2_BUF_[A-Z0-9_]+): c_int = (\d+);", rs)}
    assert c_bufs == rs_bufs
    c_defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define (RT2_(?:LAUNCH|SECONDARY|FILTER|E)_[A-Z_]+|RT2_OK|RT2_ABI_VERSION|RT2_GCONST_SIZE)\s+(-?\d+)u?\b", code)}
    rs_defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"pub const (RT2_[A-Z_]+): (?:u32|c_int|usize) = (-?\d+);", rs) if not m.group(1).startswith("RT2_BUF_")}
    rs_defs.pop("RT2_NEIGHBOR_OFFSET_COUNT", None)
    assert c_defs == rs_defs, set(c_defs.items()) ^ set(rs_defs.items())
    sizes = {m.group(1): int(m.group(2)) for m in re.finditer(r"size_of::<(rt2_[a-z_]+)>\(\) == (\d+)", rs)}
    ct = {"rt2_vertex": _ffi.Vertex, "rt2_geometry_info": _ffi.GeometryInfo, "rt2_hit_record": _ffi.HitRecord, "rt2_launch_params": _ffi.LaunchParams,
          "rt2_stats": _ffi.Stats}
    for name, t in ct.items():
        assert sizes[name] == C.sizeof(t), name
    assert sizes["rt2_light_info"] == 48 and sizes["rt2_packed_di_reservoir"] == 24 and sizes["rt2_packed_gi_reservoir"] == 32 and sizes["rt2_secondary_gbuffer_data"] == 48
    # every struct field of the header appears, in order, in the Rust struct of the same name
    for sm in re.finditer(r"typedef struct (rt2_[a-z_]+) \{(.*?)\} \1;", code, flags=re.S):
        name, body = sm.group(1), sm.group(2)
        c_fields = [re.sub(r"\[.*", "", f.strip().split()[-1].lstrip("*")) for f in body.split(";") if f.strip()]
        rm = re.search(r"pub struct %s \{(.*?)\n\}" % name, rs, flags=re.S)
        assert rm, name
        rs_fields = re.findall(r"pub ([a-z_0-9]+):", rm.group(1))
        assert c_fields == rs_fields, (name, c_fields, rs_fields)
