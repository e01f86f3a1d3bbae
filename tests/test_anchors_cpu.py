"""The oracle against anchors that do not come from the oracle (tests/anchors.py): the emissive-quad form factor, and a numpy
restatement of the sRGB decode + Vulkan LOD-0 nearest / bilinear REPEAT filters.  CPU only."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O
import anchors as A


def primary_ray(g, x, y):
    o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
    O.lib().orc_primary_ray(C.cast(C.byref(g), C.c_void_p), x, y, o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p))
    return o.astype(np.float64), d.astype(np.float64)


def test_emissive_quad_irradiance_closed_form(tmp_path):
    """one square light over a diffuse plane.  K1 with 2 bounces estimates albedo x 12 e x F(p), F the
    differential-element-to-rectangle form factor; the mean over the image must agree with the closed form within the
    Monte-Carlo error (sigma ~ 0.1 % here), and no pixel may exceed the brightest possible value albedo x 12 e."""
    w, h, spp = 48, 32, 512
    sc, g = A.emissive_quad_scene(tmp_path, w, h)
    osc = O.OracleScene(sc.desc)
    hits = osc.trace_primary(g, w, h, flags=O.ORC_BRUTE_FORCE)
    assert (hits[:, :, 1] == 0).all(), "every primary ray must land on the receiver"
    want = A.quad_expected_radiance(hits, g, w, h, primary_ray)
    d32 = osc.render_reference(g, 1, w, h, spp=spp, max_bounces=2)[2][:, :, 0].astype(np.float64)
    assert want.min() > 0.05 and want.max() < A.QUAD_ALBEDO * 12 * A.QUAD_EMISSION
    assert d32.max() <= A.QUAD_ALBEDO * 12 * A.QUAD_EMISSION * (1 + 1e-6)
    rel = (d32.mean() - want.mean()) / want.mean()
    assert abs(rel) < 5e-3, rel
    # per pixel: binomial noise, sigma = sqrt(F (1 - F) / spp) x scale; allow 5 sigma
    F = want / (A.QUAD_ALBEDO * 12 * A.QUAD_EMISSION)
    sigma = np.sqrt(F * (1 - F) / spp) * (A.QUAD_ALBEDO * 12 * A.QUAD_EMISSION)
    assert (np.abs(d32 - want) < 5 * sigma + 1e-6).mean() > 0.999


def test_srgb_decode_table_matches_iec_61966():
    """Every one of the 256 codes, through a NEAREST 256x1 texture: the oracle's LUT is float32(IEC formula in binary64)."""
    from asset_writers import GltfBuilder, write_png
    tex = np.zeros((1, 256, 4), np.uint8); tex[0, :, 0] = np.arange(256); tex[0, :, 1] = np.arange(256)[::-1]; tex[0, :, 2] = 128; tex[0, :, 3] = 255
    b = GltfBuilder()
    b.add_image_png(write_png(tex, 6))
    b.j["samplers"] = [{"magFilter": 9728, "minFilter": 9728}]
    b.j["textures"] = [{"source": 0, "sampler": 0}]
    A._add_mesh(b, [[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]], uv=[[0, 0], [1, 0], [0, 1]], material={"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}})
    import tempfile, os
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "lut.glb"); open(p, "wb").write(b.glb())
        sc = rt.Scene.from_gltf(p)
    osc = O.OracleScene(sc.desc)
    out = np.zeros(3, np.float32)
    for i in range(256):
        O.lib().orc_sample_texture(osc._p, 0, (i + 0.5) / 256.0, 0.5, out.ctypes.data_as(C.c_void_p))
        want = A.srgb_decode_iec([i, 255 - i, 128]).astype(np.float32)
        assert np.array_equal(out, want), (i, out, want)


@pytest.mark.parametrize("nearest", [True, False])
def test_texture_filter_matches_numpy_restatement(tmp_path, nearest):
    """REPEAT addressing with negative and > 1 coordinates, half-texel offset, weights: the oracle's sample_texture against the
    numpy restatement of the Vulkan rules.  NEAREST bit-exact; LINEAR within 2 fp32 ulps of the binary64 evaluation."""
    sc, g, tex, _ = A.textured_quad_scene(tmp_path, 32, 32, nearest)
    osc = O.OracleScene(sc.desc)
    rng = np.random.default_rng(1)
    uvs = np.concatenate([rng.uniform(-3.0, 4.0, (400, 2)), [[0.0, 0.0], [1.0, 1.0], [-1.0, 2.0], [0.5 / 7, 0.5 / 5], [-0.5 / 7, 1 - 0.5 / 5]]])
    out = np.zeros(3, np.float32)
    for u, v in uvs.astype(np.float32):
        O.lib().orc_sample_texture(osc._p, 0, float(u), float(v), out.ctypes.data_as(C.c_void_p))
        if nearest:
            assert np.array_equal(out, A.sample_nearest_repeat(tex, u, v).astype(np.float32)), (u, v)
        else:
            want = A.sample_bilinear_repeat(tex, u, v)
            assert np.allclose(out, want, rtol=3e-7, atol=1e-7), (u, v, out, want)
