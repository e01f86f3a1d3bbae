"""The CUDA path (through the C ABI) against anchors that do not come from the oracle: the furnace series, the emissive-quad form
factor, and the numpy restatement of sRGB decode + nearest / bilinear REPEAT filtering (tests/anchors.py).  No oracle call
decides pass / fail here."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from raytracer2_b200 import _ffi
import anchors as A

pytestmark = pytest.mark.gpu


def _ctx(sc, g, w, h):
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    return ctx, lp


def _primary_ray_host(g):
    """setupPrimaryRay restated in float64 from the view constants (GBufferHelpers.glsl:12-27): pixel centre -> clip at depth 1
    -> view through matClipToView, normalised, -> world through matViewToWorld; origin = camera position.  mat4 are
    column-major (glam)."""
    v = g.view
    m_cv = np.array(list(v.mat_clip_to_view), np.float64).reshape(4, 4).T
    m_vw = np.array(list(v.mat_view_to_world), np.float64).reshape(4, 4).T
    size = np.array(list(v.viewport_size), np.float64)
    origin = np.array(list(v.camera_direction_or_position), np.float64)[:3]

    def ray(_, x, y):
        uv = np.array([x + 0.5, y + 0.5]) / size
        target = m_cv @ np.array([uv[0] * 2 - 1, uv[1] * 2 - 1, 1.0, 1.0])
        nt = target[:3] / np.linalg.norm(target[:3])
        d = (m_vw @ np.array([nt[0], nt[1], nt[2], 0.0]))[:3]
        return origin, d
    return ray


def test_furnace_closed_form_on_the_gpu():
    """Closed box, albedo a, emission e: K1 returns sum_{k<B} a^k 12 e per sample whatever the directions are."""
    w = h = 64
    sc = rt.Scene(rt.SCENE_FURNACE, 1, 16, 0)
    g = rt.default_gconst(sc.camera(1.0), w, h, sc.num_lights)
    g.refrence_mode = 1
    ctx, lp = _ctx(sc, g, w, h)
    for bounces in (1, 2, 5):
        lp.execute(1, rt.LaunchParams(spp=3, max_bounces=bounces))
        d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)
        a, e = np.float32(0.5), np.float32(0.25) * np.float32(12.0)
        acc = np.float32(0)
        for _ in range(3):
            t = np.float32(1)
            for _ in range(bounces):
                acc = np.float32(acc + np.float32(t * e))
                t = np.float32(t * a)
        want = np.float32(acc / np.float32(3))
        rays = ctx.stats().rays_traced
        assert 0.995 * w * h * 3 * bounces <= rays <= w * h * 3 * bounces
        exact = np.all(d32[:, :, :3] == want, axis=2)
        assert exact.mean() > 0.98 and np.all(d32[:, :, :3] <= want), (bounces, d32[0, 0], want)
        if bounces == 1:
            assert exact.all()
    ctx.close()


def test_emissive_quad_irradiance_closed_form_on_the_gpu(tmp_path):
    w, h, spp = 96, 64, 2048
    sc, g = A.emissive_quad_scene(tmp_path, w, h)
    ctx, lp = _ctx(sc, g, w, h)
    lp.trace_primary()
    hits = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8)
    assert (hits[:, :, 1] == 0).all()
    lp.execute(1, rt.LaunchParams(spp=spp, max_bounces=2))
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)[:, :, 0].astype(np.float64)
    ctx.close()
    want = A.quad_expected_radiance(hits, g, w, h, _primary_ray_host(g))
    scale = A.QUAD_ALBEDO * 12 * A.QUAD_EMISSION
    assert d32.max() <= scale * (1 + 1e-6)
    rel = (d32.mean() - want.mean()) / want.mean()
    assert abs(rel) < 2e-3, rel
    F = want / scale
    sigma = np.sqrt(F * (1 - F) / spp) * scale
    assert (np.abs(d32 - want) < 5 * sigma + 1e-6).mean() > 0.999


@pytest.mark.parametrize("nearest", [True, False])
def test_texture_fetch_on_the_gpu_matches_numpy_restatement(tmp_path, nearest):
    """A textured quad (uv from -1.6 to 2.3: REPEAT of negative and > 1 coordinates) inside a box emitting 12 e = 3: after one
    bounce a quad pixel reads fp32(texture colour x 3).  Colour from the numpy restatement at the uv the hit record gives."""
    w, h = 160, 120
    sc, g, tex, uv = A.textured_quad_scene(tmp_path, w, h, nearest)
    ctx, lp = _ctx(sc, g, w, h)
    lp.trace_primary()
    hits = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8)
    lp.execute(1, rt.LaunchParams(spp=1, max_bounces=2))
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.float32).reshape(h, w, 4)[:, :, :3]
    ctx.close()
    on_quad = hits[:, :, 1] == 0
    assert on_quad.mean() > 0.3
    checked = exact = 0
    H, W = tex.shape[:2]
    for y in range(0, h, 3):
        for x in range(0, w, 3):
            if not on_quad[y, x]:
                continue
            u, v = A.quad_uv_at(hits, x, y, uv)
            if nearest:
                # skip pixels whose uv sits within 1e-4 texels of a texel edge (the shader interpolates uv in fp32)
                fu, fv = u * W, v * H
                if min(abs(fu - round(fu)), abs(fv - round(fv))) < 1e-3:
                    continue
                want = (A.sample_nearest_repeat(tex, u, v).astype(np.float32) * np.float32(3.0)).astype(np.float32)
                # the interpolated vertex colour (1, 1, 1) is (1-u-v) + u + v in fp32: up to 2 ulps off 1.0
                assert np.allclose(d32[y, x], want, rtol=4e-7, atol=0), (x, y, u, v, d32[y, x], want)
                exact += 1
            else:
                want = A.sample_bilinear_repeat(tex, u, v) * 3.0
                assert np.allclose(d32[y, x], want, rtol=0, atol=2e-4), (x, y, u, v, d32[y, x], want)
            checked += 1
    assert checked > 300
