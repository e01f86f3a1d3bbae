"""Committed fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py with the oracle's brute-force intersector).
CPU: the oracle, walking its BVH this time, still reproduces them bit for bit (drift guard + BVH-vs-brute-force check).
GPU: the CUDA path reproduces them through the C ABI without the oracle library in the loop."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
import make_golden as MG   # noqa: E402  (case table + input builders shared with the generator)


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


# ---- CPU: oracle vs fixtures -----------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["cornell", "hall", "instanced"])
def test_oracle_k1_reproduces_golden(name):
    z = _load("k1_" + name)
    sc, g, osc, w, h = MG.scene(name)
    g.refrence_mode = 1; g.environment = int(z["environment"])
    assert np.array_equal(osc.trace_primary(g, w, h), z["hit_ids"])                      # BVH walk == brute force of the fixture
    d16, s16, d32, rays = osc.render_reference(g, int(z["frame"]), w, h, spp=int(z["spp"]), max_bounces=int(z["bounces"]))
    assert rays == int(z["rays"])
    assert np.array_equal(d16, z["diffuse_f16"]) and np.array_equal(d32.view(np.uint32), z["diffuse_f32_bits"])


def test_oracle_k9_reproduces_golden():
    z = _load("k9_cornell")
    sc, g, osc, w, h = MG.scene("cornell")
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = int(z["local"]); isp.num_primary_brdf_samples = int(z["brdf"]); isp.enable_initial_visibility = int(z["initial_vis"])
    g.restir_di.shading_params.enable_final_visibility = int(z["final_vis"])
    gb = osc.gbuffer(g, w, h)
    assert np.array_equal(gb["depth"].view(np.uint32), z["depth_bits"]) and np.array_equal(gb["normals"], z["normals"])
    lights, g2l, packed, dims = MG.light_inputs(sc, osc)
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    res, d16, s16, rays = osc.di_fused_resampling(g, int(z["frame"]), w, h, gb, lights, g2l, packed, *dims, pitch)
    assert rays == int(z["rays"]) and np.array_equal(res, z["reservoirs"])
    assert np.array_equal(d16, z["diffuse_f16"]) and np.array_equal(s16, z["specular_f16"])


def _ris_matches(ris, z):
    """K11's tiles against the fixture (2 MB of tiles: the fixture keeps both ends and a CRC-32 of the whole)."""
    import zlib
    return np.array_equal(ris[:1024], z["ris_head"]) and np.array_equal(ris[-1024:], z["ris_tail"]) and zlib.crc32(np.ascontiguousarray(ris).tobytes()) == int(z["ris_crc"])


def _k9_env_gconst(z):
    w, h = int(z["width"]), int(z["height"])
    sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 64, 0).add_sky(int(z["sky"]))
    d = sc.desc.contents
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1; g.environment = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = int(z["local"]); isp.num_primary_brdf_samples = 1; isp.num_primary_environment_samples = int(z["env"])
    isp.environment_map_importance_sampling = 1; isp.local_light_sampling_mode = 1; isp.enable_initial_visibility = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    return sc, g, w, h


def test_oracle_k9_environment_sampling_reproduces_golden():
    """The presampling build of the sampler under a sky (environment light at lights + 1, K11 tiles, BRDF-miss branch): the oracle's BVH walk
    against the fixture its brute-force intersector wrote."""
    z = _load("k9_env_cornell_sky")
    sc, g, w, h = _k9_env_gconst(z)
    osc = O.OracleScene(sc.desc)
    gb = osc.gbuffer(g, w, h)
    import ctypes as C
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    dims = (pw.value, ph.value, pm.value)
    tasks, g2l, lights, pdf0 = osc.prepare_lights(dims[0], dims[1], env=True)
    assert np.array_equal(lights, z["lights"])
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(*dims)), np.float32)
    packed[:dims[0] * dims[1]] = pdf0.ravel()
    osc.generate_mips(0, packed, *dims)
    epacked, edims, ris, data = MG.env_inputs(sc, osc, g, int(z["frame"]), lights, packed, dims)
    assert np.array_equal(epacked[:edims[0] * edims[1]], z["env_pdf_mip0"]) and _ris_matches(ris, z)
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    res, d16, s16, rays = osc.di_fused_resampling(g, int(z["frame"]), w, h, gb, lights, g2l, packed, *dims, pitch, flags=O.ORC_PRESAMPLING,
                                                  ris=ris, ris_light_data=data, env_pdf=epacked, env_dims=edims)
    assert rays == int(z["rays"]) and np.array_equal(res, z["reservoirs"])
    assert np.array_equal(d16, z["diffuse_f16"]) and np.array_equal(s16, z["specular_f16"])
    chosen = res[res[:, 0] != 0, 0] & 0x7fffffff
    assert (chosen == sc.num_lights + 1).any() and (chosen < sc.num_lights).any()      # both the sky and the lamp are selected somewhere


@pytest.mark.parametrize("name", ["cornell", "hall"])
def test_oracle_gi_chain_reproduces_golden(name):
    z = _load("gi_" + name)
    sc, g, osc, w, h = MG.scene(name)
    g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 0; g.enable_temporal_resampling = 0; g.enable_spatial_resampling = 0
    gb = osc.gbuffer(g, w, h)
    st = osc.gi_chain(g, int(z["frame"]), w, h, gb, g.restir_gi.reservoir_buffer_params.reservoir_array_pitch)
    assert tuple(int(r) for r in st["rays"]) == tuple(int(r) for r in z["rays"])
    for k, key in (("secondary", "secondary"), ("gi_reservoirs", "gi_reservoirs"), ("diffuse_f16", "diffuse"), ("specular_f16", "specular")):
        assert np.array_equal(st[key], z[k]), k


def test_oracle_codecs_reproduce_golden():
    z = _load("codecs")
    L = O.lib()
    L.orc_encode_logluv.restype = C.c_uint32; L.orc_encode_snorm_oct.restype = C.c_uint32
    for i, (x, y) in enumerate(z["positions"]):
        assert L.orc_reservoir_position_to_pointer(240, 240 * 135, int(x), int(y), i & 1) == int(z["pointers"][i])
    rgb = np.ascontiguousarray(z["rgb"]); n = np.ascontiguousarray(z["normals"])
    assert [L.orc_encode_logluv(c.ctypes.data_as(C.c_void_p)) for c in rgb] == z["logluv"].tolist()
    assert [L.orc_encode_snorm_oct(v.ctypes.data_as(C.c_void_p)) for v in n] == z["snorm_oct"].tolist()
    x, y = C.c_uint32(), C.c_uint32()
    for i in range(256):
        L.orc_linear_to_zcurve(i * 257, C.byref(x), C.byref(y))
        assert (x.value, y.value) == tuple(z["zcurve"][i])


# ---- GPU: CUDA path vs fixtures (no oracle call) ------------------------------------------------------------------
def _gpu(name):
    kind, scale, seed, tex, w, h = MG.CASES[name]
    sc = rt.Scene(kind, scale, tex, seed)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    return sc, g, ctx, rt.LightPasses(ctx, w, h), w, h


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell", "hall", "instanced"])
def test_gpu_k1_reproduces_golden(name):
    z = _load("k1_" + name)
    sc, g, ctx, lp, w, h = _gpu(name)
    g.refrence_mode = 1; g.environment = int(z["environment"])
    lp.update_uniform(g)
    lp.trace_primary()
    assert np.array_equal(ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8), z["hit_ids"])
    lp.execute(int(z["frame"]), rt.LaunchParams(spp=int(z["spp"]), max_bounces=int(z["bounces"])))
    assert ctx.stats().rays_traced == int(z["rays"])
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(h, w, 4)
    assert int((d32 != z["diffuse_f32_bits"]).any(axis=2).sum()) <= 2
    assert int((ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4) != z["diffuse_f16"]).any(axis=2).sum()) <= 2
    ctx.close()


@pytest.mark.gpu
def test_gpu_k9_reproduces_golden():
    z = _load("k9_cornell")
    sc, g, ctx, lp, w, h = _gpu("cornell")
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = int(z["local"]); isp.num_primary_brdf_samples = int(z["brdf"]); isp.enable_initial_visibility = int(z["initial_vis"])
    g.restir_di.shading_params.enable_final_visibility = int(z["final_vis"])
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    lp.update_uniform(g)
    lp.execute(int(z["frame"]))
    assert np.array_equal(ctx.read(rt.BUF_GBUFFER_DEPTH, np.uint32).reshape(h, w), z["depth_bits"])
    assert ctx.stats().rays_traced - w * h == int(z["rays"])
    res = ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).reshape(-1, 6)
    assert int((res != z["reservoirs"]).any(axis=1).sum()) <= 2
    assert int((ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4) != z["diffuse_f16"]).any(axis=2).sum()) <= 2
    assert int((ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4) != z["specular_f16"]).any(axis=2).sum()) <= 2
    ctx.close()


@pytest.mark.gpu
def test_gpu_k9_environment_sampling_reproduces_golden():
    z = _load("k9_env_cornell_sky")
    sc, g, w, h = _k9_env_gconst(z)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    rt.PrepareLightsTasks(ctx).execute()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.LOCAL_LIGHTS).execute_mip_generation()
    rt.GenerateMipsPass(ctx, rt.GenerateMipsPass.ENVIRONMENT).execute_mip_generation()
    lp.update_uniform(g)
    lp.execute_presampeling(int(z["frame"]), True)
    assert np.array_equal(ctx.read(rt.BUF_LIGHT_INFO, np.uint32).reshape(-1, 12), z["lights"])
    assert _ris_matches(ctx.read(rt.BUF_RIS, np.uint32).reshape(-1, 2), z)
    lp.execute(int(z["frame"]), rt.LaunchParams(flags=rt.LAUNCH_PRESAMPLING))
    assert ctx.stats().rays_traced - w * h == int(z["rays"])
    res = ctx.read(rt.BUF_DI_RESERVOIRS, np.uint32).reshape(-1, 6)
    assert int((res != z["reservoirs"]).any(axis=1).sum()) <= 2
    assert int((ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4) != z["diffuse_f16"]).any(axis=2).sum()) <= 2
    assert int((ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4) != z["specular_f16"]).any(axis=2).sum()) <= 2
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell", "hall"])
def test_gpu_gi_chain_reproduces_golden(name):
    z = _load("gi_" + name)
    sc, g, ctx, lp, w, h = _gpu(name)
    g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 0; g.enable_temporal_resampling = 0; g.enable_spatial_resampling = 0
    lp.update_uniform(g)
    lp.execute(int(z["frame"]))
    assert ctx.stats().rays_traced - w * h == int(sum(int(r) for r in z["rays"]))
    sec = ctx.read(rt.BUF_SECONDARY_GBUFFER, np.uint32).reshape(-1, 12)
    gi = ctx.read(rt.BUF_GI_RESERVOIRS, np.uint32).reshape(-1, 8)
    assert int((sec != z["secondary"]).any(axis=1).sum()) <= 2
    assert int((gi != z["gi_reservoirs"]).any(axis=1).sum()) <= 2
    assert int((ctx.read(rt.BUF_DIFFUSE_F16, np.uint16).reshape(h, w, 4) != z["diffuse_f16"]).any(axis=2).sum()) <= 2
    assert int((ctx.read(rt.BUF_SPECULAR_F16, np.uint16).reshape(h, w, 4) != z["specular_f16"]).any(axis=2).sum()) <= 2
    ctx.close()
