"""Generates the fixtures in this directory.  Run from the repo root:  python tests/golden/make_golden.py

The reference ships no tests, golden images or known-answer vectors for its shaders, and its GLSL cannot be
run on this box (no graphics stack), so these vectors come from the CPU oracle (oracle/, brute-force intersection where
the scene is small enough: no BVH of either implementation is involved).  They pin the oracle against drift — any change
to oracle/ that alters a bit of these outputs fails tests/test_golden.py — and give the CUDA path a fixed target that
does not depend on the oracle library being rebuilt identically on the GPU box.  They are NOT outputs of the reference:
parity with the reference itself stays "unpinned" as DESIGN.md states.

Each .npz holds the inputs' description (scene kind/scale/seed/texture size, frame, launch parameters) next to the outputs.
"""
import ctypes as C
import os
import sys

import zlib
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import raytracer2_b200 as rt          # noqa: E402  (host library only: scene generators, GConst defaults)
from oracle import oracle_ffi as O    # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

# name -> (scene kind, scale, seed, texture size, width, height)
CASES = {
    "cornell": (rt.SCENE_CORNELL, 1.0, 0, 64, 64, 64),
    "hall": (rt.SCENE_HALL, 0.125, 0, 64, 80, 45),
    "instanced": (rt.SCENE_INSTANCED, 0.05, 0, 64, 80, 45),
}


def scene(name):
    kind, scale, seed, tex, w, h = CASES[name]
    sc = rt.Scene(kind, scale, tex, seed)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    return sc, g, O.OracleScene(sc.desc), w, h


def light_inputs(sc, osc):
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    tasks, g2l, lights, pdf0 = osc.prepare_lights(pw.value, ph.value)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(pw.value, ph.value, pm.value)), np.float32)
    packed[:pw.value * ph.value] = pdf0.ravel()
    osc.generate_mips(0, packed, pw.value, ph.value, pm.value)
    return lights, g2l, packed, (pw.value, ph.value, pm.value)


def k1(name, frame, spp, bounces, env=0):
    """K1 reference path tracer + primary hit ids."""
    sc, g, osc, w, h = scene(name)
    g.refrence_mode = 1; g.environment = env
    hits = osc.trace_primary(g, w, h, flags=O.ORC_BRUTE_FORCE)
    d16, s16, d32, rays = osc.render_reference(g, frame, w, h, spp=spp, max_bounces=bounces, flags=O.ORC_BRUTE_FORCE)
    np.savez_compressed(os.path.join(OUT, f"k1_{name}.npz"), case=name, frame=frame, spp=spp, bounces=bounces, environment=env,
                        hit_ids=hits, diffuse_f16=d16, diffuse_f32_bits=d32.view(np.uint32), rays=np.uint64(rays))


def k9(name, frame):
    """K3 G-buffer -> K9 direct lighting through DI reservoirs (uniform light selection, the shipped build)."""
    sc, g, osc, w, h = scene(name)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = 4; isp.num_primary_brdf_samples = 1; isp.enable_initial_visibility = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    gb = osc.gbuffer(g, w, h, flags=O.ORC_BRUTE_FORCE)
    lights, g2l, packed, dims = light_inputs(sc, osc)
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    res, d16, s16, rays = osc.di_fused_resampling(g, frame, w, h, gb, lights, g2l, packed, *dims, pitch, flags=O.ORC_BRUTE_FORCE)
    np.savez_compressed(os.path.join(OUT, f"k9_{name}.npz"), case=name, frame=frame, local=4, brdf=1, initial_vis=1, final_vis=1,
                        depth_bits=gb["depth"].view(np.uint32), normals=gb["normals"], diffuse_albedo=gb["diffuse_albedo"],
                        reservoirs=res, diffuse_f16=d16, specular_f16=s16, rays=np.uint64(rays))


def env_inputs(sc, osc, g, frame, lights, packed, dims):
    """K12 environment pyramid + K11 tiles of both segments (the presampling build's inputs)."""
    import math
    d = sc.desc.contents
    sw, sh = d.sky_width, d.sky_height
    emips = int(math.ceil(math.log2(max(sw, sh)))) + 1
    epacked = np.zeros(sum(a * b for a, b in O.mip_dims(sw, sh, emips)), np.uint16)
    osc.generate_mips(1, epacked, sw, sh, emips)
    ris, data = O.presample(g, frame, 0, packed, *dims, lights)
    ris, data = O.presample(g, frame, 1, epacked, sw, sh, emips, ris=ris, ris_light_data=data)
    return epacked, (sw, sh, emips), ris, data


def k9_env(frame, sky=128, local=2, env=3, w=64, h=64):
    """K9 as compiled with RTXDI_ENABLE_PRESAMPLING 1 under a sky: POWER_RIS local lights, environment-map importance sampling from
    K11's environment tiles, BRDF rays that miss select the environment light K10 writes at lights + 1 (SURVEY Q7 pinned to intent)."""
    sc = rt.Scene(rt.SCENE_CORNELL, 1.0, 64, 0).add_sky(sky)
    d = sc.desc.contents
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(d.sky_width, d.sky_height))
    osc = O.OracleScene(sc.desc)
    g.refrence_mode = 0; g.enable_restir_gi = 0; g.enable_restir_di = 1; g.environment = 1
    isp = g.restir_di.initial_sampling_params
    isp.num_primary_local_light_samples = local; isp.num_primary_brdf_samples = 1; isp.num_primary_environment_samples = env
    isp.environment_map_importance_sampling = 1; isp.local_light_sampling_mode = 1; isp.enable_initial_visibility = 1
    g.restir_di.shading_params.enable_final_visibility = 1
    gb = osc.gbuffer(g, w, h, flags=O.ORC_BRUTE_FORCE)
    pw, ph, pm = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rt._ffi.host_lib().rt2h_compute_pdf_texture_size(sc.num_lights, C.byref(pw), C.byref(ph), C.byref(pm))
    dims = (pw.value, ph.value, pm.value)
    tasks, g2l, lights, pdf0 = osc.prepare_lights(dims[0], dims[1], env=True)
    packed = np.zeros(sum(a * b for a, b in O.mip_dims(*dims)), np.float32)
    packed[:dims[0] * dims[1]] = pdf0.ravel()
    osc.generate_mips(0, packed, *dims)
    epacked, edims, ris, data = env_inputs(sc, osc, g, frame, lights, packed, dims)
    pitch = g.restir_di.reservoir_buffer_params.reservoir_array_pitch
    res, d16, s16, rays = osc.di_fused_resampling(g, frame, w, h, gb, lights, g2l, packed, *dims, pitch, flags=O.ORC_BRUTE_FORCE | O.ORC_PRESAMPLING,
                                                  ris=ris, ris_light_data=data, env_pdf=epacked, env_dims=edims)
    np.savez_compressed(os.path.join(OUT, "k9_env_cornell_sky.npz"), frame=frame, sky=sky, local=local, env=env, width=w, height=h, lights=lights,
                        env_pdf_mip0=epacked[:edims[0] * edims[1]], ris_head=ris[:1024], ris_tail=ris[-1024:], ris_crc=np.uint32(zlib.crc32(ris.tobytes())), reservoirs=res, diffuse_f16=d16, specular_f16=s16, rays=np.uint64(rays))
    return sc, g, osc, gb, lights, g2l, packed, dims, epacked, edims, ris, data


def gi(name, frame):
    """The default frame K3 -> K4 -> K5 -> K8 (src/main.rs:391-394: ReSTIR GI on, DI off, no resampling)."""
    sc, g, osc, w, h = scene(name)
    g.refrence_mode = 0; g.enable_restir_gi = 1; g.enable_restir_di = 0
    g.enable_temporal_resampling = 0; g.enable_spatial_resampling = 0
    gb = osc.gbuffer(g, w, h, flags=O.ORC_BRUTE_FORCE)
    pitch = g.restir_gi.reservoir_buffer_params.reservoir_array_pitch
    st = osc.gi_chain(g, frame, w, h, gb, pitch, flags=O.ORC_BRUTE_FORCE)
    np.savez_compressed(os.path.join(OUT, f"gi_{name}.npz"), case=name, frame=frame, secondary=st["secondary"], gi_reservoirs=st["gi_reservoirs"],
                        diffuse_f16=st["diffuse"], specular_f16=st["specular"], rays=np.array(st["rays"], np.uint64))


def codecs():
    """Packing known-answers: RTXDI_ReservoirPositionToPointer, LogLuv, snorm-octahedral, Z-curve."""
    L = O.lib()
    rng = np.random.default_rng(20240611)
    pos = rng.integers(0, 4096, size=(256, 2), dtype=np.uint32)
    ptr = np.array([L.orc_reservoir_position_to_pointer(240, 240 * 135, int(x), int(y), int(i & 1)) for i, (x, y) in enumerate(pos)], np.uint32)
    rgb = (rng.random((256, 3)) * np.array([4.0, 2.0, 8.0])).astype(np.float32)
    rgb[:4] = [[0, 0, 0], [1, 1, 1], [100, 0.5, 0.01], [1e-4, 1e-4, 1e-4]]
    L.orc_encode_logluv.restype = C.c_uint32
    logluv = np.array([L.orc_encode_logluv(c.ctypes.data_as(C.c_void_p)) for c in np.ascontiguousarray(rgb)], np.uint32)
    n = rng.normal(size=(256, 3)).astype(np.float32); n /= np.linalg.norm(n, axis=1, keepdims=True)
    n[:3] = [[0, 0, 1], [0, 0, -1], [1, 0, 0]]
    L.orc_encode_snorm_oct.restype = C.c_uint32
    oct_ = np.array([L.orc_encode_snorm_oct(v.ctypes.data_as(C.c_void_p)) for v in np.ascontiguousarray(n)], np.uint32)
    zc = np.zeros((256, 2), np.uint32); x, y = C.c_uint32(), C.c_uint32()
    for i in range(256):
        L.orc_linear_to_zcurve(i * 257, C.byref(x), C.byref(y)); zc[i] = (x.value, y.value)
    np.savez_compressed(os.path.join(OUT, "codecs.npz"), positions=pos, pointers=ptr, rgb=rgb, logluv=logluv, normals=n, snorm_oct=oct_, zcurve=zc)


if __name__ == "__main__":
    k1("cornell", 1, 1, 2)          # BASELINE config C1's launch at fixture size
    k1("hall", 3, 2, 3)
    k1("instanced", 2, 2, 3, env=1)
    k9("cornell", 1)
    k9_env(2)
    gi("cornell", 1)
    gi("hall", 2)
    codecs()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
