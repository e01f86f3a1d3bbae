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
