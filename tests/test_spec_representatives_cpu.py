"""The index rule behind K1's primary-hit reuse in the speculative schedule (rt2_render.cu, k_spec_gen / k_spec_gen_cand / k_spec_shade): the
paths of a pixel are consecutive threads of a 256-thread block iteration, alive as a prefix of the pixel's group, and block_append_index hands
out consecutive places in thread order — so the path `back = min(j, threadIdx.x)` places before a path is an alive path of the SAME pixel whose
own `back` is 0 (the representative that is queued for the first walk).  A numpy restatement of the append, checked over group sizes that do
and do not divide the block, ragged alive prefixes and dead pixels."""
import numpy as np
import pytest


def _append(group, alive_prefix, block=256):
    """group: paths per pixel; alive_prefix[k]: how many of pixel k's paths are alive (0 = dead pixel).  Returns per item: place p (or -1), back."""
    n_items = group * len(alive_prefix)
    i = np.arange(n_items)
    k, j = i // group, i % group
    alive = j < np.asarray(alive_prefix)[k]
    t = i % block                                        # threadIdx.x (grid-stride: every block iteration covers 256 consecutive items)
    back = np.minimum(j, t)
    p = np.full(n_items, -1, np.int64)
    base = 0
    for b0 in range(0, n_items, block):                  # any order of the block iterations gives the same property; places are per iteration
        sel = slice(b0, min(b0 + block, n_items))
        a = alive[sel]
        p[sel] = np.where(a, base + np.cumsum(a) - 1, -1)
        base += int(a.sum())
    return k, j, alive, back, p


@pytest.mark.parametrize("group", [2, 3, 4, 5, 12, 13, 37, 300])
def test_every_path_finds_a_representative_of_its_own_pixel(group):
    rng = np.random.default_rng(group)
    n_pixels = 1500 if group < 100 else 40
    prefix = rng.integers(0, group + 1, n_pixels)
    prefix[rng.random(n_pixels) < 0.5] = group          # most pixels have every path alive (round 0)
    k, j, alive, back, p = _append(group, prefix)
    pix_of_place = np.full(int(alive.sum()), -1, np.int64)
    back_of_place = np.full(int(alive.sum()), -1, np.int64)
    pix_of_place[p[alive]] = k[alive]; back_of_place[p[alive]] = back[alive]
    assert (np.sort(p[alive]) == np.arange(alive.sum())).all()          # the places are a permutation: every path has its own slot
    rep = p[alive] - back[alive]
    assert (rep >= 0).all()
    assert (pix_of_place[rep] == k[alive]).all(), "the representative belongs to the same pixel"
    assert (back_of_place[rep] == 0).all(), "and is itself queued for the walk"
    # walks saved: one representative per (pixel, block part) instead of one per path
    reps = int((back[alive] == 0).sum())
    parts = len({(int(kk), int(ii) // 256) for kk, ii, a in zip(k, np.arange(len(k)), alive) if a})
    assert reps == parts
