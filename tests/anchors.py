"""Independent anchors for the oracle and the CUDA path (the reference ships no vectors):

* closed forms — the single emissive quad's irradiance (differential-area-to-rectangle form factor) and the furnace series;
* the texture path — an IEC 61966-2-1 sRGB decode and the Vulkan §16 LOD-0 nearest / bilinear REPEAT filters restated here in
  numpy (float64), compared with the oracle's sample_texture and with what the CUDA path renders through a textured quad.

Scene builders only; the tests live in test_anchors_cpu.py (oracle) and test_gpu_anchors.py (CUDA)."""
import ctypes as C

import numpy as np

import raytracer2_b200 as rt
from raytracer2_b200 import _ffi
from asset_writers import GltfBuilder, write_png


def _add_mesh(b, pos, idx, uv=None, normal=(0.0, 0.0, 1.0), material=None):
    pos = np.asarray(pos, np.float32)
    nrm = np.tile(np.array([normal], np.float32), (len(pos), 1))
    attrs = {"POSITION": b.accessor(pos, "VEC3"), "NORMAL": b.accessor(nrm, "VEC3")}
    if uv is not None:
        attrs["TEXCOORD_0"] = b.accessor(np.asarray(uv, np.float32), "VEC2")
    prim = {"attributes": attrs, "indices": b.accessor(np.asarray(idx, np.uint16).reshape(-1), "SCALAR")}
    if material is not None:
        b.j["materials"].append(material); prim["material"] = len(b.j["materials"]) - 1
    b.j["meshes"].append({"primitives": [prim]})
    b.j["nodes"].append(dict(mesh=len(b.j["meshes"]) - 1))
    b.j["scenes"][0]["nodes"].append(len(b.j["nodes"]) - 1)


def _camera(pos, w, h, fov=65.0):
    cam = rt.Camera()
    p, d = (C.c_float * 3)(*pos), (C.c_float * 3)(0.0, 0.0, 1.0)      # looks down -z (same convention as test_adversarial_geometry)
    _ffi.host_lib().rt2h_camera_new(C.byref(cam), p, d, fov, w / h, 0.1, 1000.0)
    return cam


# ---- single emissive quad over a diffuse receiver ----------------------------------------------------------------
QUAD_ALBEDO = 0.75
QUAD_EMISSION = 0.5          # emissiveFactor; the shader multiplies by 12 (Hit.glsl:41)
QUAD_HALF = 1.0              # light = [-1, 1]^2 at z = QUAD_HEIGHT
QUAD_HEIGHT = 3.0
QUAD_CAMERA_Z = 2.0          # between receiver (z = 0) and light: the light is behind the camera


def emissive_quad_scene(tmp_path, w, h):
    b = GltfBuilder()
    R = 40.0
    _add_mesh(b, [[-R, -R, 0], [R, -R, 0], [R, R, 0], [-R, R, 0]], [[0, 1, 2], [0, 2, 3]],
              material={"pbrMetallicRoughness": {"baseColorFactor": [QUAD_ALBEDO] * 3 + [1.0]}})
    s, z = QUAD_HALF, QUAD_HEIGHT
    _add_mesh(b, [[-s, -s, z], [s, -s, z], [s, s, z], [-s, s, z]], [[0, 1, 2], [0, 2, 3]], normal=(0.0, 0.0, -1.0),
              material={"emissiveFactor": [QUAD_EMISSION] * 3, "pbrMetallicRoughness": {"baseColorFactor": [0.0, 0.0, 0.0, 1.0]}})
    p = tmp_path / "emissive_quad.glb"
    p.write_bytes(b.glb())
    sc = rt.Scene.from_gltf(str(p))
    g = rt.default_gconst(_camera((0.0, 0.0, QUAD_CAMERA_Z), w, h), w, h, sc.num_lights)
    g.refrence_mode = 1
    return sc, g


def _corner_factor(a, b, h):
    """Form factor from a differential element to a parallel a x b rectangle whose corner sits on the element's normal at
    distance h (Howell's catalogue, B-4): 1/(2 pi) [a/sqrt(a^2+h^2) atan(b/sqrt(a^2+h^2)) + b/sqrt(b^2+h^2) atan(a/sqrt(b^2+h^2))]."""
    ra, rb = np.sqrt(a * a + h * h), np.sqrt(b * b + h * h)
    return (a / ra * np.arctan(b / ra) + b / rb * np.arctan(a / rb)) / (2.0 * np.pi)


def quad_form_factor(px, py):
    """Cosine-weighted fraction of the hemisphere above (px, py, 0) covered by the light: inclusion-exclusion over corners."""
    def G(a, b):
        return np.sign(a) * np.sign(b) * _corner_factor(np.abs(a), np.abs(b), QUAD_HEIGHT)
    s = QUAD_HALF
    return G(s - px, s - py) - G(-s - px, s - py) - G(s - px, -s - py) + G(-s - px, -s - py)


def quad_expected_radiance(hit_records, g, w, h, primary_ray):
    """Per pixel: albedo x 12 e x F(hit point) for pixels whose primary ray lands on the receiver (geometry 0), NaN elsewhere.
    K1 with 2 bounces and no sky is exactly this estimator (refrence.rgen:28-46: emission at the second hit x throughput)."""
    out = np.full((h, w), np.nan)
    for y in range(h):
        for x in range(w):
            if hit_records[y, x, 1] != 0:
                continue
            o, d = primary_ray(g, x, y)
            t = float(hit_records[y, x, 3:4].view(np.float32)[0])
            p = o + t * d
            out[y, x] = QUAD_ALBEDO * 12.0 * QUAD_EMISSION * quad_form_factor(p[0], p[1])
    return out


# ---- textured quad inside an emissive box ----------------------------------------------------------------------
TEX_BOX_EMISSION = 0.25      # x 12 = 3.0 exactly: a pixel on the quad reads fp32(texture colour x 3) after one bounce


def srgb_decode_iec(c8):
    """IEC 61966-2-1: C_lin = C/12.92 for C <= 0.04045 else ((C + 0.055)/1.055)^2.4, C = c8/255."""
    c = np.asarray(c8, np.float64) / 255.0
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def sample_nearest_repeat(tex, u, v):
    """Vulkan §16.5.9 (nearest, LOD 0): texel index floor(u * W) wrapped with REPEAT (mod W, non-negative)."""
    H, W = tex.shape[:2]
    x = int(np.floor(np.float32(u) * np.float32(W))) % W
    y = int(np.floor(np.float32(v) * np.float32(H))) % H
    return srgb_decode_iec(tex[y, x, :3])


def sample_bilinear_repeat(tex, u, v):
    """Vulkan §16.5.9 / §16.9.3 (linear, LOD 0): unnormalised coordinate u*W - 0.5, i0 = floor, weights = fraction, both texel
    indices wrapped with REPEAT; filtered in linear light (sRGB decode before the weights)."""
    H, W = tex.shape[:2]
    uu = np.float64(np.float32(np.float32(u) * np.float32(W)) - np.float32(0.5))
    vv = np.float64(np.float32(np.float32(v) * np.float32(H)) - np.float32(0.5))
    x0, y0 = int(np.floor(uu)), int(np.floor(vv))
    a, b = uu - x0, vv - y0
    t = lambda x, y: srgb_decode_iec(tex[y % H, x % W, :3])
    return (t(x0, y0) * (1 - a) + t(x0 + 1, y0) * a) * (1 - b) + (t(x0, y0 + 1) * (1 - a) + t(x0 + 1, y0 + 1) * a) * b


def textured_quad_scene(tmp_path, w, h, nearest, tex_w=7, tex_h=5, seed=9):
    """A quad with uv spanning [-1.6, 2.3] x [-0.7, 1.9] (negative and > 1: REPEAT) in front of the camera, inside a closed
    emissive box.  Returns (scene, gconst, texture uint8 [H, W, 4], uv corners)."""
    rng = np.random.default_rng(seed)
    tex = rng.integers(0, 256, (tex_h, tex_w, 4), dtype=np.uint8); tex[..., 3] = 255
    b = GltfBuilder()
    b.add_image_png(write_png(tex, 6))
    b.j["samplers"] = [{"magFilter": 9728 if nearest else 9729, "minFilter": 9728 if nearest else 9729}]
    b.j["textures"] = [{"source": 0, "sampler": 0}]
    uv = [[-1.6, -0.7], [2.3, -0.7], [2.3, 1.9], [-1.6, 1.9]]
    _add_mesh(b, [[-1.5, -1, 0], [1.5, -1, 0], [1.5, 1, 0], [-1.5, 1, 0]], [[0, 1, 2], [0, 2, 3]], uv=uv,
              material={"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}}})
    # closed emissive box around everything (6 faces, 12 triangles), black albedo so nothing but its emission matters
    B = 6.0
    c = np.array([[-B, -B, -B], [B, -B, -B], [B, B, -B], [-B, B, -B], [-B, -B, B], [B, -B, B], [B, B, B], [-B, B, B]], np.float32)
    faces = [[0, 1, 2], [0, 2, 3], [4, 6, 5], [4, 7, 6], [0, 4, 5], [0, 5, 1], [3, 2, 6], [3, 6, 7], [0, 3, 7], [0, 7, 4], [1, 5, 6], [1, 6, 2]]
    _add_mesh(b, c, faces, material={"emissiveFactor": [TEX_BOX_EMISSION] * 3, "pbrMetallicRoughness": {"baseColorFactor": [0.0, 0.0, 0.0, 1.0]}})
    p = tmp_path / ("textured_quad_%s.glb" % ("nearest" if nearest else "linear"))
    p.write_bytes(b.glb())
    sc = rt.Scene.from_gltf(str(p))
    g = rt.default_gconst(_camera((0.0, 0.0, 2.2), w, h), w, h, sc.num_lights)
    g.refrence_mode = 1
    return sc, g, tex, np.array(uv, np.float64)


def quad_uv_at(hit_records, x, y, uv):
    """uv of the primary hit on the textured quad from the hit record's barycentrics (weights of vertices 1 and 2)."""
    prim = int(hit_records[y, x, 2])
    bu, bv = (float(f) for f in hit_records[y, x, 4:6].view(np.float32))
    tri = ([0, 1, 2], [0, 2, 3])[prim]
    return uv[tri[0]] * (1.0 - bu - bv) + uv[tri[1]] * bu + uv[tri[2]] * bv
