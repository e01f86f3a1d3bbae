"""rt2_trace_primary_noalloc (no counterpart in the reference): the primary-visibility
probe for callers that cannot call cudaMalloc.  Called through the C ABI, two calls in a row on one context, and compared with
the existing entry point rt2_trace_primary — hit records bit-exact (the tolerance every hit-id test here uses), hit count equal.
Collected last on purpose (pytest -x must not lose other tests to a failure here)."""
import ctypes as C

import numpy as np
import pytest
import torch

import raytracer2_b200 as rt

pytestmark = pytest.mark.gpu

POISON = 0xA5A5A5A5


def _setup(kind, scale, w, h):
    sc = rt.Scene(kind, scale, 32, 0)
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights)
    g.refrence_mode = 1
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, w, h)
    lp.update_uniform(g)
    return sc, g, ctx, lp


def _reference(ctx, lp, w, h, params):
    lp.trace_primary(params)
    return ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(h, w, 8).copy(), ctx.stats().rays_hit


@pytest.mark.parametrize("kind,scale", [(rt.SCENE_HALL, 0.125), (rt.SCENE_INSTANCED, 0.05)])
def test_trace_primary_noalloc_matches_trace_primary_on_two_calls_in_a_row(kind, scale):
    w, h = 200, 150                          # ragged 8x4 tiles and a ragged last band
    sc, g, ctx, lp = _setup(kind, scale, w, h)
    full = None
    bands = rt.LaunchParams(band_rows=16, band_rank=1, band_world=3)
    want_full, hits_full = _reference(ctx, lp, w, h, full)
    ctx.write(rt.BUF_HIT_IDS, np.full((h, w, 8), POISON, np.uint32))
    want_band, hits_band = _reference(ctx, lp, w, h, bands)
    assert 0 < hits_band < hits_full

    # caller-owned device memory, allocated before the calls
    out = torch.from_numpy(np.full(h * w * 8, POISON, np.uint32).view(np.int32)).cuda()
    count = torch.zeros(1, dtype=torch.int64, device="cuda")
    nbytes = h * w * 32
    rt.trace_primary_noalloc(ctx, out.data_ptr(), nbytes, count.data_ptr(), full)          # warm-up: loads the kernels (lazy module loading)
    ctx.sync()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]

    # call 1: whole frame; call 2, straight after it on the same stream: one rank's bands into a re-poisoned buffer.
    # The static scratch still holds call 1's per-warp counts when call 2 runs.
    out.copy_(torch.from_numpy(np.full(h * w * 8, POISON, np.uint32).view(np.int32)))
    torch.cuda.synchronize()
    rt.trace_primary_noalloc(ctx, out.data_ptr(), nbytes, count.data_ptr(), full)
    ctx.sync()
    got_full = out.cpu().numpy().view(np.uint32).reshape(h, w, 8).copy()
    n_full = int(count.item())
    out.copy_(torch.from_numpy(np.full(h * w * 8, POISON, np.uint32).view(np.int32)))
    torch.cuda.synchronize()
    rt.trace_primary_noalloc(ctx, out.data_ptr(), nbytes, count.data_ptr(), bands)
    rt.trace_primary_noalloc(ctx, out.data_ptr(), nbytes, None, bands)                      # and again without the count word
    ctx.sync()
    got_band = out.cpu().numpy().view(np.uint32).reshape(h, w, 8)
    n_band = int(count.item())
    free1 = torch.cuda.mem_get_info()[0]

    assert np.array_equal(got_full, want_full), f"{int((got_full != want_full).any(axis=2).sum())} hit records differ (whole frame)"
    assert n_full == hits_full
    owned = np.array([(y // 16) % 3 == 1 for y in range(h)])
    assert np.array_equal(got_band[owned], want_band[owned]), "owned rows differ (band call)"
    assert (got_band[~owned] == POISON).all(), "the band call wrote rows it does not own"
    assert n_band == hits_band
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across the no-alloc calls"
    ctx.close()


def test_trace_primary_noalloc_rejects_bad_arguments_and_oversize_frames():
    w, h = 64, 48
    sc, g, ctx, lp = _setup(rt.SCENE_CORNELL, 1.0, w, h)
    out = torch.zeros(h * w * 8, dtype=torch.int32, device="cuda")
    with pytest.raises(rt.Rt2Error, match="size mismatch"):
        rt.trace_primary_noalloc(ctx, out.data_ptr(), h * w * 32 - 32)
    with pytest.raises(rt.Rt2Error, match="NULL"):
        rt.trace_primary_noalloc(ctx, 0, h * w * 32)
    with pytest.raises(rt.Rt2Error, match="band"):
        rt.trace_primary_noalloc(ctx, out.data_ptr(), h * w * 32, 0, rt.LaunchParams(band_rows=8, band_rank=3, band_world=3))
    # beyond the static scratch: 4096 x 2304 = 9 437 184 padded pixels > 2^23
    W2, H2 = 4096, 2304
    ctx.resize(W2, H2)
    g2 = rt.default_gconst(sc.camera(W2 / H2), W2, H2, sc.num_lights)
    g2.refrence_mode = 1
    lp.update_uniform(g2)
    big = torch.zeros(H2 * W2 * 8, dtype=torch.int32, device="cuda")
    with pytest.raises(rt.Rt2Error, match="exceed the static scratch"):
        rt.trace_primary_noalloc(ctx, big.data_ptr(), H2 * W2 * 32)
    # ... while one rank's share of the same frame fits, and the existing entry point still takes the whole frame
    rt.trace_primary_noalloc(ctx, big.data_ptr(), H2 * W2 * 32, 0, rt.LaunchParams(band_rows=8, band_rank=0, band_world=2))
    lp.trace_primary()
    ctx.sync()
    want = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(H2, W2, 8)
    got = big.cpu().numpy().view(np.uint32).reshape(H2, W2, 8)
    owned = np.array([(y // 8) % 2 == 0 for y in range(H2)])
    assert np.array_equal(got[owned], want[owned])
    ctx.close()
