"""the C++ asset readers (raytracer2_b200/host/rt2_ingest.cpp) that stand where the reference uses the
`gltf` and `image` crates (src/gltf/mod.rs:50-168, src/main.rs:62-63).  The reference's own assets are git-ignored
upstream, so the readers are checked on files written by tests/asset_writers.py: a procedural scene exported to GLB
must come back as the very same upload buffers (and render the very same image), and each format feature the reference's
loader handles (strided / normalised accessors, TRS nodes, shared meshes, PNG colour types + filters, EXR compressions)
has its own case."""
import ctypes as C
import os

import numpy as np
import pytest

import raytracer2_b200 as rt
from oracle import oracle_ffi as O
from asset_writers import GltfBuilder, write_exr, write_png


def desc_arrays(sc):
    d = sc.desc.contents
    out = {
        "vertices": np.ctypeslib.as_array(C.cast(d.vertices, C.POINTER(C.c_float)), shape=(d.num_vertices, 16)).copy(),
        "indices": np.ctypeslib.as_array(d.indices, shape=(d.num_indices,)).copy(),
        "geoms": np.ctypeslib.as_array(C.cast(d.geometries, C.POINTER(C.c_uint32)), shape=(d.num_geometries, 32)).copy(),
        "index_counts": np.ctypeslib.as_array(d.index_counts, shape=(d.num_geometries,)).copy(),
        "images": [np.ctypeslib.as_array(d.images[i].rgba8, shape=(d.images[i].height, d.images[i].width, 4)).copy() for i in range(d.num_images)],
        "samplers": [(d.samplers[i].mag_filter, d.samplers[i].min_filter) for i in range(d.num_samplers)],
        "textures": [(d.textures[i].image_index, d.textures[i].sampler_index) for i in range(d.num_textures)],
    }
    return out


def export_glb(sc, index_dtype=np.uint32, embed_images=True, stride_pad=0):
    """One glTF mesh + node per geometry of a procedural scene, attributes as the reference's reader expects them."""
    a = desc_arrays(sc)
    b = GltfBuilder()
    g = a["geoms"]
    for gi in range(len(g)):
        gf = g[gi].view(np.float32)
        voff, ioff, icount = int(g[gi][23]), int(g[gi][22]), int(a["index_counts"][gi])
        idx = a["indices"][ioff:ioff + icount]
        nv = int(idx.max()) + 1
        v = a["vertices"][voff:voff + nv]
        prim = {"attributes": {"POSITION": b.accessor(v[:, 0:3], "VEC3", stride_pad=stride_pad), "NORMAL": b.accessor(v[:, 4:7], "VEC3"),
                               "COLOR_0": b.accessor(v[:, 8:12], "VEC4"), "TEXCOORD_0": b.accessor(v[:, 12:14], "VEC2", byte_offset=8 if stride_pad else 0)},
                "indices": b.accessor(idx.astype(index_dtype), "SCALAR"), "material": gi}
        mat = {"emissiveFactor": [float(x) for x in gf[24:27]],
               "pbrMetallicRoughness": {"baseColorFactor": [float(x) for x in gf[16:20]], "metallicFactor": float(gf[21]), "roughnessFactor": float(gf[28])}}
        tex = int(g[gi].view(np.int32)[20])
        if tex >= 0:
            mat["pbrMetallicRoughness"]["baseColorTexture"] = {"index": tex}
        b.j["materials"].append(mat)
        b.j["meshes"].append({"primitives": [prim]})
        b.j["nodes"].append({"mesh": gi, "matrix": [float(x) for x in gf[0:16]]})
        b.j["scenes"][0]["nodes"].append(gi)
    for im in a["images"]:
        b.add_image_png(write_png(im, 6), embed=embed_images)
    if a["samplers"]:
        assert a["samplers"][0] == (rt.FILTER_LINEAR, rt.FILTER_LINEAR)
        b.j["samplers"] = [{"magFilter": 9729 if m == rt.FILTER_LINEAR else 9728, "minFilter": 9987 if n == rt.FILTER_LINEAR else 9984} for (m, n) in a["samplers"][1:]]
    if a["textures"]:
        b.j["textures"] = [dict(source=i, **({"sampler": s - 1} if s else {})) for (i, s) in a["textures"]]
    return b, a


def assert_same_scene(a, l):
    assert np.array_equal(a["vertices"].view(np.uint32), l["vertices"].view(np.uint32))
    assert np.array_equal(a["indices"], l["indices"]) and np.array_equal(a["index_counts"], l["index_counts"])
    assert np.array_equal(a["geoms"][:, :29], l["geoms"][:, :29])         # 116 bytes of fields; the tail is padding
    assert len(a["images"]) == len(l["images"]) and all(np.array_equal(x, y) for x, y in zip(a["images"], l["images"]))
    assert a["textures"] == l["textures"]
    if a["textures"]:
        assert a["samplers"] == l["samplers"]


@pytest.mark.parametrize("kind,scale,tex,index_dtype,pad", [
    (rt.SCENE_CORNELL, 1.0, 0, np.uint8, 0), (rt.SCENE_HALL, 0.05, 16, np.uint16, 4), (rt.SCENE_MANYLIGHT, 0.05, 0, np.uint32, 0)])
def test_glb_round_trip_reproduces_the_upload_buffers(tmp_path, kind, scale, tex, index_dtype, pad):
    sc = rt.Scene(kind, scale, tex, 0)
    b, a = export_glb(sc, index_dtype, stride_pad=pad)
    if index_dtype != np.uint32 and max(int(c) for c in a["index_counts"]) and a["vertices"].shape[0] >= np.iinfo(index_dtype).max:
        big = [gi for gi in range(len(a["geoms"])) if a["indices"][int(a["geoms"][gi][22]):int(a["geoms"][gi][22]) + int(a["index_counts"][gi])].max() > np.iinfo(index_dtype).max]
        assert not big, "scene too large for this index type"
    p = tmp_path / "scene.glb"
    p.write_bytes(b.glb())
    loaded = rt.Scene.from_gltf(str(p))
    assert_same_scene(a, desc_arrays(loaded))
    assert loaded.num_lights == sc.num_lights and loaded.num_triangles == sc.num_triangles
    cam = loaded.camera(16 / 9)                                          # src/main.rs:114-121
    assert tuple(cam.position) == (0.0, 0.0, 10.0) and cam.fov == 65.0 and abs(cam.z_near - 0.1) < 1e-7 and cam.z_far == 1000.0


def test_loaded_scene_renders_like_the_procedural_one(tmp_path):
    w = h = 48
    sc = rt.Scene(rt.SCENE_HALL, 0.05, 16, 0)
    b, _ = export_glb(sc, np.uint32, embed_images=False)
    js, bin_ = b.gltf("scene.bin")                                       # .gltf + external .bin + data-URI images
    (tmp_path / "scene.gltf").write_bytes(js); (tmp_path / "scene.bin").write_bytes(bin_)
    loaded = rt.Scene.from_gltf(str(tmp_path / "scene.gltf"))
    g = rt.default_gconst(sc.camera(1.0), w, h, sc.num_lights); g.refrence_mode = 1
    ref = O.OracleScene(sc.desc).render_reference(g, 1, w, h, spp=2, max_bounces=3)
    got = O.OracleScene(loaded.desc).render_reference(g, 1, w, h, spp=2, max_bounces=3)
    assert np.array_equal(ref[2].view(np.uint32), got[2].view(np.uint32)) and ref[3] == got[3]
    assert (ref[2][:, :, :3] > 0).any()


def _tri_builder(node, color=None, uv=None, normalized=True, material=True):
    b = GltfBuilder()
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    nrm = np.array([[0, 0, 1]] * 3, np.float32)
    attrs = {"POSITION": b.accessor(pos, "VEC3"), "NORMAL": b.accessor(nrm, "VEC3")}
    if color is not None:
        attrs["COLOR_0"] = b.accessor(color, "VEC3" if color.shape[1] == 3 else "VEC4", normalized=normalized and color.dtype != np.float32)
    if uv is not None:
        attrs["TEXCOORD_0"] = b.accessor(uv, "VEC2", normalized=normalized and uv.dtype != np.float32)
    prim = {"attributes": attrs, "indices": b.accessor(np.array([0, 1, 2], np.uint16), "SCALAR")}
    if material:
        b.j["materials"].append({"emissiveFactor": [1.0, 0.5, 0.25]}); prim["material"] = 0
    b.j["meshes"].append({"primitives": [prim]})
    b.j["nodes"].append(dict(mesh=0, **node))
    return b


def test_node_trs_is_composed_like_the_gltf_crate(tmp_path):
    q = np.array([0.1, -0.3, 0.2, 0.9], np.float64); q /= np.linalg.norm(q)
    t, s = [1.5, -2.0, 0.25], [2.0, 0.5, 3.0]
    b = _tri_builder({"translation": t, "rotation": [float(x) for x in q], "scale": s})
    b.j["nodes"].append({"mesh": 0})                                     # same mesh again, default transform: two geometries sharing vertices
    b.j["nodes"].append({"name": "empty"})                               # no mesh: skipped
    p = tmp_path / "trs.glb"; p.write_bytes(b.glb())
    l = desc_arrays(rt.Scene.from_gltf(str(p)))
    assert len(l["geoms"]) == 2 and len(l["vertices"]) == 3 and len(l["indices"]) == 3      # mesh data stored once (mesh_index_redirect)
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    M = np.eye(4); M[:3, :3] = R @ np.diag(s); M[:3, 3] = t
    got = l["geoms"][0][:16].view(np.float32).reshape(4, 4).T            # column-major in memory
    assert np.allclose(got, M, atol=1e-6)
    assert np.array_equal(l["geoms"][1][:16].view(np.float32).reshape(4, 4), np.eye(4, dtype=np.float32))
    gf = l["geoms"][0].view(np.float32)
    assert tuple(gf[24:28]) == (1.0, 0.5, 0.25, 1.0) and tuple(gf[16:20]) == (1.0, 1.0, 1.0, 1.0)      # emission.w = 1 (src/model.rs:404-409); default base colour
    assert gf[21] == 1.0 and gf[28] == 1.0 and l["geoms"][0].view(np.int32)[20] == -1                    # metallic / roughness defaults, no texture
    assert np.array_equal(l["vertices"][:, 8:12], np.ones((3, 4), np.float32)) and not l["vertices"][:, 12:14].any()   # Vec4::ONE, Vec2::ZERO


def test_normalised_integer_attributes_convert_like_into_f32(tmp_path):
    col = np.array([[0, 128, 255], [255, 0, 64], [1, 2, 3]], np.uint8)
    uv = np.array([[0, 65535], [32768, 1], [65535, 0]], np.uint16)
    b = _tri_builder({}, color=col, uv=uv, material=False)
    p = tmp_path / "norm.glb"; p.write_bytes(b.glb())
    l = desc_arrays(rt.Scene.from_gltf(str(p)))
    assert np.array_equal(l["vertices"][:, 8:11], col.astype(np.float32) / np.float32(255.0))
    assert np.array_equal(l["vertices"][:, 11], np.ones(3, np.float32))                              # RGB colour: alpha 1
    assert np.array_equal(l["vertices"][:, 12:14], uv.astype(np.float32) / np.float32(65535.0))
    gf = l["geoms"][0].view(np.float32)
    assert not gf[24:27].any()                                                                        # no material: the crate's default material


def test_unsupported_primitives_are_skipped_and_errors_are_reported(tmp_path):
    b = _tri_builder({})
    del b.j["meshes"][0]["primitives"][0]["attributes"]["NORMAL"]          # is_primitive_supported (mod.rs:170-174)
    p = tmp_path / "nonormal.glb"; p.write_bytes(b.glb())
    with pytest.raises(rt.Rt2Error, match="no node references a supported"):
        rt.Scene.from_gltf(str(p))
    b = _tri_builder({})
    b.add_image_png(b"\xff\xd8\xff\xe0" + b"\0" * 64)
    p = tmp_path / "jpeg.glb"; p.write_bytes(b.glb())
    with pytest.raises(rt.Rt2Error, match="JPEG"):
        rt.Scene.from_gltf(str(p))
    b = _tri_builder({})
    b.j["accessors"][0]["count"] = 1000                                    # reads past the buffer
    p = tmp_path / "oob.glb"; p.write_bytes(b.glb())
    with pytest.raises(rt.Rt2Error, match="past the end"):
        rt.Scene.from_gltf(str(p))
    (tmp_path / "bad.gltf").write_text('{"asset": {"version": "2.0"}, "meshes": [')
    with pytest.raises(rt.Rt2Error, match="JSON"):
        rt.Scene.from_gltf(str(tmp_path / "bad.gltf"))
    glb = bytearray(_tri_builder({}).glb()); glb[4] = 1
    (tmp_path / "v1.glb").write_bytes(bytes(glb))
    with pytest.raises(rt.Rt2Error, match="GLB"):
        rt.Scene.from_gltf(str(tmp_path / "v1.glb"))


@pytest.mark.parametrize("color_type,samples", [(0, 1), (4, 2), (2, 3), (6, 4)])
def test_png_colour_types_expand_like_the_reference(tmp_path, color_type, samples):
    """src/gltf/image.rs:57-109: R8 -> (r,0,0,255), R8G8 -> (r,g,0,255), RGB -> (r,g,b,255), RGBA verbatim; all five row filters."""
    rng = np.random.default_rng(color_type)
    px = rng.integers(0, 256, size=(13, 9, samples), dtype=np.uint8)
    b = _tri_builder({})
    b.add_image_png(write_png(px, color_type))
    b.j["textures"] = [{"source": 0}]
    p = tmp_path / "img.glb"; p.write_bytes(b.glb())
    l = desc_arrays(rt.Scene.from_gltf(str(p)))
    want = np.zeros((13, 9, 4), np.uint8); want[:, :, 3] = 255
    want[:, :, :min(samples, 3) if samples != 2 else 2] = px[:, :, :min(samples, 3) if samples != 2 else 2]
    if samples == 4:
        want[:, :, 3] = px[:, :, 3]
    assert np.array_equal(l["images"][0], want)
    assert l["textures"] == [(0, 0)] and l["samplers"][0] == (rt.FILTER_LINEAR, rt.FILTER_LINEAR)


def test_png_palette_low_bit_depth_and_rejections(tmp_path):
    rng = np.random.default_rng(5)
    pal = rng.integers(0, 256, size=(16, 3), dtype=np.uint8)
    idx = rng.integers(0, 16, size=(7, 11), dtype=np.uint8)
    for trns in (None, rng.integers(0, 256, size=10, dtype=np.uint8)):
        b = _tri_builder({})
        b.add_image_png(write_png(idx, 3, bit_depth=4, palette=pal, trns=trns))
        p = tmp_path / "pal.glb"; p.write_bytes(b.glb())
        img = desc_arrays(rt.Scene.from_gltf(str(p)))["images"][0]
        assert np.array_equal(img[:, :, :3], pal[idx])
        alpha = np.full(16, 255, np.uint8)
        if trns is not None:
            alpha[:10] = trns
        assert np.array_equal(img[:, :, 3], alpha[idx])
    grey = rng.integers(0, 4, size=(5, 6), dtype=np.uint8)                # 2-bit grey -> L8 -> red channel
    b = _tri_builder({}); b.add_image_png(write_png(grey, 0, bit_depth=2))
    p = tmp_path / "g2.glb"; p.write_bytes(b.glb())
    img = desc_arrays(rt.Scene.from_gltf(str(p)))["images"][0]
    assert np.array_equal(img[:, :, 0], grey * 85) and not img[:, :, 1:3].any()
    wide = rng.integers(0, 65536, size=(4, 4, 3)).astype(np.uint16)
    b = _tri_builder({}); b.add_image_png(write_png(wide, 2, bit_depth=16))
    p = tmp_path / "wide.glb"; p.write_bytes(b.glb())
    with pytest.raises(rt.Rt2Error, match="16 bytes images"):             # the reference's own message (image.rs:38)
        rt.Scene.from_gltf(str(p))


@pytest.mark.parametrize("compression", [0, 1, 2, 3])
@pytest.mark.parametrize("dtype,alpha", [(np.float16, False), (np.float32, True)])
def test_exr_sky_box_reader(tmp_path, compression, dtype, alpha):
    rng = np.random.default_rng(compression * 7 + (1 if alpha else 0))
    h, w = 37, 50                                                         # not a multiple of the 16-line ZIP block
    ch = {n: (rng.random((h, w)) * 8).astype(dtype) for n in ("R", "G", "B")}
    ch["R"][3:9, :] = dtype(0.5)                                          # flat runs: RLE / ZIP actually compress
    if alpha:
        ch["A"] = rng.random((h, w)).astype(dtype)
    p = tmp_path / "sky.exr"
    p.write_bytes(write_exr(ch, compression, line_order=compression & 1))
    img = rt.load_exr(str(p))
    assert img.shape == (h, w, 4) and img.dtype == np.float32
    for i, n in enumerate("RGB"):
        assert np.array_equal(img[:, :, i], ch[n].astype(np.float32)), n
    assert np.array_equal(img[:, :, 3], ch["A"].astype(np.float32) if alpha else np.ones((h, w), np.float32))   # to_rgba32f: alpha 1


def test_exr_rejections_and_scene_with_sky(tmp_path):
    ch = {n: np.full((8, 8), v, np.float16) for n, v in (("R", 1.0), ("G", 2.0), ("B", 3.0))}
    good = write_exr(ch, 3)
    bad = bytearray(good); k = bad.index(b"compression\0compression\0") + len(b"compression\0compression\0") + 4; bad[k] = 5     # PXR24 (PIZ = 4 is read since round 2: test_ingest_independent_cpu.py)
    (tmp_path / "pxr24.exr").write_bytes(bytes(bad))
    with pytest.raises(rt.Rt2Error, match="compression 5"):
        rt.load_exr(str(tmp_path / "pxr24.exr"))
    (tmp_path / "junk.exr").write_bytes(b"not an exr file at all")
    with pytest.raises(rt.Rt2Error, match="OpenEXR"):
        rt.load_exr(str(tmp_path / "junk.exr"))
    (tmp_path / "sky.exr").write_bytes(good)
    (tmp_path / "tri.glb").write_bytes(_tri_builder({}).glb())
    sc = rt.Scene.from_gltf(str(tmp_path / "tri.glb"), str(tmp_path / "sky.exr"))
    d = sc.desc.contents
    assert (d.sky_width, d.sky_height) == (8, 8)
    sky = np.ctypeslib.as_array(d.sky_rgba32f, shape=(8, 8, 4))
    assert np.array_equal(sky[0, 0], np.array([1, 2, 3, 1], np.float32))
    # the loaded sky lights the scene: a ray that misses the triangle returns it (environment = 1)
    g = rt.default_gconst(sc.camera(1.0), 8, 8, 0, env_size=(8, 8)); g.refrence_mode = 1; g.environment = 1
    out = O.OracleScene(sc.desc).render_reference(g, 1, 8, 8, spp=1, max_bounces=1)[2]
    assert np.allclose(out[0, 0, :3], [1, 2, 3], rtol=1e-3)


@pytest.mark.gpu
def test_gpu_renders_a_loaded_glb_like_the_procedural_scene(tmp_path):
    """The whole asset path on the device: GLB + EXR -> rt2_scene_desc -> rt2_upload_scene -> BVH build -> K1, against the
    same scene handed over from the generator, and against the oracle on the loaded buffers."""
    w, h = 160, 90
    sc = rt.Scene(rt.SCENE_HALL, 0.125, 32, 0)
    b, _ = export_glb(sc, np.uint32)
    (tmp_path / "hall.glb").write_bytes(b.glb())
    rng = np.random.default_rng(3)
    sky = {n: (rng.random((16, 32)) * 2).astype(np.float16) for n in "RGB"}
    (tmp_path / "sky.exr").write_bytes(write_exr(sky, 3))
    loaded = rt.Scene.from_gltf(str(tmp_path / "hall.glb"), str(tmp_path / "sky.exr"))
    g = rt.default_gconst(sc.camera(w / h), w, h, sc.num_lights, env_size=(32, 16)); g.refrence_mode = 1

    def render(s, env):
        g.environment = env
        ctx = rt.Renderer(0)
        rt.Model.from_scene(ctx, s)
        lp = rt.LightPasses(ctx, w, h); lp.update_uniform(g)
        lp.execute(2, rt.LaunchParams(spp=2, max_bounces=3))
        out = ctx.read(rt.BUF_DIFFUSE_F32, np.uint32).reshape(h, w, 4)
        ctx.close()
        return out

    # without the sky the loaded scene is the generator's scene
    assert np.array_equal(render(sc, 0), render(loaded, 0))
    # with it (the scaled-down hall leaks a few rays to the sky), the device agrees with the oracle on the loaded buffers
    got = render(loaded, 1)
    ref = O.OracleScene(loaded.desc).render_reference(g, 2, w, h, spp=2, max_bounces=3)[2].view(np.uint32)
    assert int((got != ref).any(axis=2).sum()) <= 2
