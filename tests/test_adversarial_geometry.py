"""Edge cases of the acceleration structure and the triangle test on one hand-built GLB (through the C++ glTF reader):
degenerate triangles (repeated index, collinear, zero size), two geometries with identical coplanar triangles (the closest
hit's tie-break (t, instance, geometry, primitive) must pick the lower geometry whatever order the BVH visits them in),
a 1e4-unit triangle next to a 1e-3-unit one (quantised child boxes over six orders of magnitude), a needle of aspect 1e5,
a rotated + non-uniformly scaled node, and a 5x3 NEAREST-filtered texture addressed outside [0, 1) (REPEAT wrap of negative
coordinates, src/model.rs:165-183).  The product's builder + traversal logic on the CPU (tests/hostemu) and the CUDA path
must both reproduce the oracle's brute-force intersector bit for bit."""
import ctypes as C

import numpy as np
import pytest

import raytracer2_b200 as rt
from raytracer2_b200 import _ffi
from oracle import oracle_ffi as O
from asset_writers import GltfBuilder, write_png

W, H = 192, 128


def _add_mesh(b, pos, idx, uv=None, node=None, material=None):
    pos = np.asarray(pos, np.float32)
    nrm = np.tile(np.array([[0.0, 0.0, 1.0]], np.float32), (len(pos), 1))
    attrs = {"POSITION": b.accessor(pos, "VEC3"), "NORMAL": b.accessor(nrm, "VEC3")}
    if uv is not None:
        attrs["TEXCOORD_0"] = b.accessor(np.asarray(uv, np.float32), "VEC2")
    prim = {"attributes": attrs, "indices": b.accessor(np.asarray(idx, np.uint16).reshape(-1), "SCALAR")}
    if material is not None:
        b.j["materials"].append(material); prim["material"] = len(b.j["materials"]) - 1
    b.j["meshes"].append({"primitives": [prim]})
    b.j["nodes"].append(dict(mesh=len(b.j["meshes"]) - 1, **(node or {})))
    b.j["scenes"][0]["nodes"].append(len(b.j["nodes"]) - 1)


def build_scene(tmp_path):
    b = GltfBuilder()
    # G0 emissive back wall
    _add_mesh(b, [[-3, -3, -2], [3, -3, -2], [3, 3, -2], [-3, 3, -2]], [[0, 1, 2], [0, 2, 3]], material={"emissiveFactor": [1.0, 0.9, 0.8]})
    # G1 8x8 grid of quads at z = 0 sharing every edge
    n = 9
    gx, gy = np.meshgrid(np.linspace(-1, 1, n), np.linspace(-1, 1, n))
    grid = np.stack([gx.ravel(), gy.ravel(), np.zeros(n * n)], 1)
    quads = [[j * n + i, j * n + i + 1, (j + 1) * n + i + 1, j * n + i, (j + 1) * n + i + 1, (j + 1) * n + i] for j in range(n - 1) for i in range(n - 1)]
    gidx = np.array(quads).reshape(-1, 3)
    _add_mesh(b, grid, gidx, material={"pbrMetallicRoughness": {"baseColorFactor": [0.8, 0.2, 0.2, 1.0]}})
    # G2 the first 40 triangles of G1 again: exact ties in t between geometry 1 and geometry 2
    _add_mesh(b, grid, gidx[:40], material={"pbrMetallicRoughness": {"baseColorFactor": [0.2, 0.8, 0.2, 1.0]}})
    # G3 degenerate triangles in view: repeated index, collinear, zero size, and one healthy triangle
    _add_mesh(b, [[-2, 1.5, 0.5], [-1.5, 1.5, 0.5], [-1, 1.5, 0.5], [-1.5, 2.0, 0.5], [-1.7, 1.2, 0.5]],
              [[0, 0, 1], [0, 1, 2], [4, 4, 4], [0, 2, 3]], material={"pbrMetallicRoughness": {"baseColorFactor": [0.2, 0.2, 0.9, 1.0]}})
    # G4 six orders of magnitude in one geometry
    _add_mesh(b, [[-1e4, -1e4, -2.5], [1e4, -1e4, -2.5], [0, 1e4, -2.5], [0.0, 0.0, 2.0], [1e-3, 0.0, 2.0], [0.0, 1e-3, 2.0]], [[0, 1, 2], [3, 4, 5]],
              material={"pbrMetallicRoughness": {"baseColorFactor": [0.7, 0.7, 0.1, 1.0]}})
    # G5 NEAREST 5x3 texture, uv far outside [0,1) and negative
    rng = np.random.default_rng(5)
    tex = rng.integers(0, 256, (3, 5, 4), dtype=np.uint8); tex[..., 3] = 255
    b.add_image_png(write_png(tex, 6))
    b.j["samplers"] = [{"magFilter": 9728, "minFilter": 9728}]
    b.j["textures"] = [{"source": 0, "sampler": 0}]
    _add_mesh(b, [[1.2, -1, 0.3], [2.6, -1, 0.3], [2.6, 1, 0.3], [1.2, 1, 0.3]], [[0, 1, 2], [0, 2, 3]], uv=[[-2.3, -1.7], [3.1, -1.7], [3.1, 2.9], [-2.3, 2.9]],
              material={"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}})
    # G6 needle, aspect 1e5, under a rotated + non-uniformly scaled node
    q = np.array([0.2, -0.1, 0.3, 0.9]); q /= np.linalg.norm(q)
    _add_mesh(b, [[-3, 0, 0], [3, 0, 0], [0, 6e-5, 0], [-1, -1, 0], [1, -1, 0], [0, -0.4, 0.2]], [[0, 1, 2], [3, 4, 5]],
              node={"translation": [0.3, -1.6, 0.8], "rotation": [float(x) for x in q], "scale": [1.0, 2.5, 0.5]},
              material={"pbrMetallicRoughness": {"baseColorFactor": [0.9, 0.5, 0.9, 1.0]}})
    p = tmp_path / "adversarial.glb"
    p.write_bytes(b.glb())
    sc = rt.Scene.from_gltf(str(p))
    cam = rt.Camera()
    pos, dirv = (C.c_float * 3)(0.1, 0.05, 5.0), (C.c_float * 3)(0.0, 0.0, 1.0)
    _ffi.host_lib().rt2h_camera_new(C.byref(cam), pos, dirv, 65.0, W / H, 0.1, 1000.0)
    g = rt.default_gconst(cam, W, H, sc.num_lights)
    g.refrence_mode = 1
    return sc, g


def test_oracle_bvh_and_product_traversal_match_brute_force(tmp_path, hostemu):
    sc, g = build_scene(tmp_path)
    d = sc.desc.contents
    assert d.num_geometries == 7 and d.num_textures == 1 and d.samplers[d.textures[0].sampler_index].mag_filter == rt.FILTER_NEAREST
    osc = O.OracleScene(sc.desc)
    ref = osc.trace_primary(g, W, H, flags=O.ORC_BRUTE_FORCE)
    assert np.array_equal(osc.trace_primary(g, W, H, flags=0), ref)
    geoms = ref[:, :, 1]
    assert {0, 1, 3, 4, 5, 6} <= set(np.unique(geoms).tolist())                   # every healthy geometry is seen
    assert 2 not in set(np.unique(geoms).tolist())                                # the duplicate never wins a tie against geometry 1
    hs = hostemu.hostemu_build(C.cast(sc.desc, C.c_void_p))
    assert hs
    out = np.zeros((H, W, 8), np.uint32)
    hostemu.hostemu_trace_primary(hs, C.cast(C.byref(g), C.c_void_p), W, H, out.ctypes.data_as(C.c_void_p), None)
    assert np.array_equal(out, ref)
    # Bounce rays leaving the 1e4-unit triangle: the triangle test's sheared coordinates carry ~|vertex| * 2^-23 of error, so
    # it reports the triangle the ray starts on again at t ~ 1e-3 > TMin, outside that triangle's exact box.  The definition is
    # the brute-force result; both BVH walks must be conservative enough to reproduce it (the oracle's was not until its
    # triangle boxes were padded by |coordinate| * 2^-16; the product's quantised boxes keep 2^-10 grid cells of slack).
    rng = np.random.default_rng(1)
    n = 40000
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0] = rng.uniform(-7, 7, n); rays[:, 1] = rng.uniform(-5, 5, n); rays[:, 2] = -2.5
    d = rng.standard_normal((n, 3)); d[:, 2] = np.abs(d[:, 2]); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 4:7] = d; rays[:, 3] = 0.001; rays[:, 7] = 1e5
    brute = osc.trace_rays(rays, O.ORC_BRUTE_FORCE)
    assert ((brute[:, 1] == 4) & (brute[:, 3].view(np.float32) < 0.01)).sum() > 20          # the self-hits exist
    assert np.array_equal(osc.trace_rays(rays, 0), brute)
    out = np.zeros((n, 8), np.uint32)
    od = np.ascontiguousarray(rays[:, [0, 1, 2, 4, 5, 6]])
    hostemu.hostemu_trace_rays(hs, od.ctypes.data_as(C.c_void_p), n, 0.001, 1e5, out.ctypes.data_as(C.c_void_p), None)
    hostemu.hostemu_destroy(hs)
    assert np.array_equal(out, brute)
    a = osc.render_reference(g, 3, W, H, spp=2, max_bounces=4)[2]
    b = osc.render_reference(g, 3, W, H, spp=2, max_bounces=4, flags=O.ORC_BRUTE_FORCE)[2]
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.gpu
def test_gpu_matches_brute_force_on_adversarial_geometry(tmp_path):
    sc, g = build_scene(tmp_path)
    osc = O.OracleScene(sc.desc)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, W, H)
    lp.update_uniform(g)
    lp.trace_primary()
    got = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(H, W, 8)
    assert np.array_equal(got, osc.trace_primary(g, W, H, flags=O.ORC_BRUTE_FORCE))
    lp.execute(3, rt.LaunchParams(spp=3, max_bounces=4))
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(H, W, 4)
    ref = osc.render_reference(g, 3, W, H, spp=3, max_bounces=4)
    assert np.array_equal(d32, ref[2].view(np.uint32))
    assert (ref[2][:, :, :3] > 0).mean() > 0.1
    ctx.close()


def _soup_scene(tmp_path, k, seed):
    """k random triangles in front of the camera (some overlapping, some sharing a vertex), one geometry"""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1.5, 1.5, (k, 1, 3)) * np.array([1.0, 1.0, 0.6])
    pos = (c + rng.normal(0, 0.5, (k, 3, 3))).reshape(-1, 3)
    if k >= 2:
        pos[3] = pos[0]                                        # two triangles share a vertex
    b = GltfBuilder()
    _add_mesh(b, pos, np.arange(3 * k).reshape(-1, 3), material={"pbrMetallicRoughness": {"baseColorFactor": [0.7, 0.6, 0.5, 1.0]}})
    p = tmp_path / f"soup{k}.glb"
    p.write_bytes(b.glb())
    sc = rt.Scene.from_gltf(str(p))
    cam = rt.Camera()
    posv, dirv = (C.c_float * 3)(0.0, 0.0, 6.0), (C.c_float * 3)(0.0, 0.0, 1.0)
    _ffi.host_lib().rt2h_camera_new(C.byref(cam), posv, dirv, 50.0, 1.0, 0.1, 1000.0)
    g = rt.default_gconst(cam, 96, 96, sc.num_lights)
    g.refrence_mode = 1
    return sc, g


TINY = [1, 2, 3, 4, 8, 15, 16, 17, 31, 33, 100]        # around RT2_LEAF_MAX, one wide node, and the 16-leaf threshold of the reinsertion rounds


@pytest.mark.parametrize("k", TINY)
def test_tiny_trees_on_the_cpu_emulation(tmp_path, hostemu, k):
    """The builder's passes (PLOC, reinsertion from 16 leaves, SAH-optimal collapse) on trees of a handful of leaves: the product's
    builder + traversal functions on the CPU reproduce the brute-force intersector's hit records bit for bit."""
    sc, g = _soup_scene(tmp_path, k, 100 + k)
    ids = O.OracleScene(sc.desc).trace_primary(g, 96, 96, flags=O.ORC_BRUTE_FORCE)
    hs = hostemu.hostemu_build(C.cast(sc.desc, C.c_void_p))
    assert hs
    out = np.zeros((96 * 96, 8), np.uint32)
    hostemu.hostemu_trace_primary(hs, C.cast(C.byref(g), C.c_void_p), 96, 96, out.ctypes.data_as(C.c_void_p), None)
    hostemu.hostemu_destroy(hs)
    assert np.array_equal(out.reshape(96, 96, 8), ids)
    assert (ids[:, :, 1] != 0xFFFFFFFF).any()


@pytest.mark.gpu
@pytest.mark.parametrize("k", TINY)
def test_tiny_trees_on_the_gpu(tmp_path, k):
    sc, g = _soup_scene(tmp_path, k, 100 + k)
    ids = O.OracleScene(sc.desc).trace_primary(g, 96, 96, flags=O.ORC_BRUTE_FORCE)
    ctx = rt.Renderer(0)
    rt.Model.from_scene(ctx, sc)
    lp = rt.LightPasses(ctx, 96, 96)
    lp.update_uniform(g)
    lp.trace_primary()
    got = ctx.read(rt.BUF_HIT_IDS, np.uint32).reshape(96, 96, 8)
    assert np.array_equal(got, ids)
    # and a short path-traced frame through the same tree
    lp.execute(1, rt.LaunchParams(spp=2, max_bounces=3))
    d32 = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(96, 96, 4)
    ref = O.OracleScene(sc.desc).render_reference(g, 1, 96, 96, spp=2, max_bounces=3)[2].view(np.uint32)
    assert np.array_equal(d32, ref)
    ctx.close()
