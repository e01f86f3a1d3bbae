"""The asset readers against files written by INDEPENDENT encoders in this image: PIL (libjpeg-turbo, zlib-ng) for
JPEG / PNG, OpenCV (the OpenEXR library) for EXR in every compression the reader supports, PIZ included.  Until round 2 the
readers had only ever seen files written by tests/asset_writers.py."""
import ctypes as C
import io
import os

import numpy as np
import pytest
from PIL import Image

import raytracer2_b200 as rt
from raytracer2_b200 import _ffi
from asset_writers import GltfBuilder

os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"


def decode(data: bytes):
    h = _ffi.host_lib()
    w_, h_, c_ = C.c_uint32(), C.c_uint32(), C.c_uint32()
    err = C.create_string_buffer(256)
    p = h.rt2h_decode_image(data, len(data), C.byref(w_), C.byref(h_), C.byref(c_), err, len(err))
    if not p:
        raise rt.Rt2Error(err.value.decode())
    try:
        n = w_.value * h_.value * c_.value
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n,)).reshape(h_.value, w_.value, c_.value).copy()
    finally:
        h.rt2h_free(p)


def _photo(w, h, seed):
    """smooth gradients + edges + noise: exercises every coefficient band without being pure noise"""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.stack([127 + 120 * np.sin(x / 7.0 + seed) * np.cos(y / 11.0), 255.0 * x / max(w - 1, 1), 255.0 * ((x // 8 + y // 8) % 2)], axis=2)
    img += rng.normal(0, 12, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


@pytest.mark.parametrize("w,h", [(64, 48), (37, 53), (17, 9), (8, 8), (1, 1), (3, 130)])
@pytest.mark.parametrize("subsampling,quality", [(0, 90), (1, 75), (2, 75), (2, 30)])
def test_baseline_jpeg_matches_pil(w, h, subsampling, quality):
    """4:4:4, 4:2:2 and 4:2:0 baseline files written by PIL decode to what PIL itself decodes (libjpeg-turbo: islow IDCT, fancy
    upsampling, fixed-point colour) — stated tolerance +-1 LSB, observed 0."""
    buf = io.BytesIO()
    Image.fromarray(_photo(w, h, w + h)).save(buf, "JPEG", quality=quality, subsampling=subsampling)
    want = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
    got = decode(buf.getvalue())
    assert got.shape == want.shape
    diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
    assert diff.max() <= 1, (diff.max(), int((diff > 0).sum()))


def test_jpeg_greyscale_restart_markers_and_rejections():
    grey = _photo(70, 41, 3)[:, :, 0]
    buf = io.BytesIO(); Image.fromarray(grey, "L").save(buf, "JPEG", quality=85)
    got = decode(buf.getvalue())
    assert got.shape == (41, 70, 1) and np.abs(got[:, :, 0].astype(int) - np.asarray(Image.open(io.BytesIO(buf.getvalue()))).astype(int)).max() <= 1
    for kw in (dict(restart_marker_blocks=3), dict(restart_marker_rows=1)):
        buf = io.BytesIO(); Image.fromarray(_photo(90, 50, 5)).save(buf, "JPEG", quality=80, subsampling=2, **kw)
        assert b"\xff\xdd" in buf.getvalue()            # DRI present
        want = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
        assert np.abs(decode(buf.getvalue()).astype(int) - want.astype(int)).max() <= 1
    buf = io.BytesIO(); Image.fromarray(_photo(64, 64, 1)).save(buf, "JPEG", progressive=True)
    with pytest.raises(rt.Rt2Error, match="progressive"):
        decode(buf.getvalue())
    buf = io.BytesIO(); Image.fromarray(_photo(32, 32, 1)).convert("CMYK").save(buf, "JPEG")
    with pytest.raises(rt.Rt2Error, match="CMYK|components"):
        decode(buf.getvalue())
    with pytest.raises(rt.Rt2Error):
        decode(b"\xff\xd8\xff\xe0\x00\x10JFIF")          # truncated


@pytest.mark.parametrize("mode", ["RGB", "RGBA", "L", "LA", "P"])
@pytest.mark.parametrize("optimize", [False, True])
def test_pil_written_png_matches_pil(mode, optimize):
    """someone else's zlib (zlib-ng) and filter choices, not asset_writers.write_png's"""
    img = Image.fromarray(_photo(61, 47, 9))
    if mode == "RGBA":
        img.putalpha(Image.fromarray(_photo(61, 47, 2)[:, :, 0]))
    elif mode in ("L", "LA"):
        img = img.convert(mode)
    elif mode == "P":
        img = img.quantize(37)
    buf = io.BytesIO(); img.save(buf, "PNG", optimize=optimize)
    got = decode(buf.getvalue())
    want = np.asarray(img.convert("RGB") if mode == "P" else img)
    if want.ndim == 2:
        want = want[:, :, None]
    assert got.shape == want.shape and np.array_equal(got, want)


def test_glb_with_pil_written_jpeg_and_png_textures(tmp_path):
    """the real path: gltf::import's image decode inside the GLB loader (src/gltf/mod.rs:50-52, src/gltf/image.rs:57-109)"""
    jpg, png = io.BytesIO(), io.BytesIO()
    a, b_ = _photo(40, 24, 4), _photo(16, 16, 6)
    Image.fromarray(a).save(jpg, "JPEG", quality=88, subsampling=2)
    Image.fromarray(b_).save(png, "PNG")
    b = GltfBuilder()
    b.add_image_png(jpg.getvalue()); b.add_image_png(png.getvalue())
    b.j["images"][0]["mimeType"] = "image/jpeg"
    b.j["textures"] = [{"source": 0}, {"source": 1}]
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    for t in (0, 1):
        attrs = {"POSITION": b.accessor(pos, "VEC3"), "NORMAL": b.accessor(np.tile([[0, 0, 1.0]], (3, 1)).astype(np.float32), "VEC3"),
                 "TEXCOORD_0": b.accessor(pos[:, :2].copy(), "VEC2")}
        b.j["materials"].append({"pbrMetallicRoughness": {"baseColorTexture": {"index": t}}})
        b.j["meshes"].append({"primitives": [{"attributes": attrs, "indices": b.accessor(np.arange(3, dtype=np.uint16), "SCALAR"), "material": t}]})
        b.j["nodes"].append({"mesh": t}); b.j["scenes"][0]["nodes"].append(t)
    p = tmp_path / "textured.glb"; p.write_bytes(b.glb())
    sc = rt.Scene.from_gltf(str(p))
    d = sc.desc.contents
    assert d.num_images == 2
    want = [np.asarray(Image.open(io.BytesIO(jpg.getvalue())).convert("RGB")), b_]
    for i in range(2):
        im = d.images[i]
        got = np.ctypeslib.as_array(im.rgba8, shape=(im.height, im.width, 4))
        assert (got[:, :, 3] == 255).all()
        assert np.abs(got[:, :, :3].astype(int) - want[i].astype(int)).max() <= (1 if i == 0 else 0)


@pytest.mark.parametrize("compression", ["NO", "RLE", "ZIPS", "ZIP", "PIZ"])
@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("w,h", [(53, 37), (200, 70), (16, 33)])
def test_exr_written_by_opencv(tmp_path, compression, half, w, h):
    """scan-line EXR files from the OpenEXR library itself (through OpenCV): every supported compression, HALF and FLOAT, sizes that
    leave ragged last blocks (PIZ: 32-line blocks) and odd widths (PIZ's wavelet has edge cases for odd sizes).  Lossless => exact."""
    import cv2
    rng = np.random.default_rng(w * h)
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([np.sin(x / 9) * 3 + 3, y / h * 100, (rng.random((h, w)) * 8).astype(np.float32)], axis=2).astype(np.float32)   # RGB
    img[0, 0] = [0, 65000.0, 1e-5]
    path = str(tmp_path / f"sky_{compression}.exr")
    ok = cv2.imwrite(path, img[:, :, ::-1].copy(), [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_HALF if half else cv2.IMWRITE_EXR_TYPE_FLOAT,
                                                     cv2.IMWRITE_EXR_COMPRESSION, getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + compression)])
    assert ok
    got = rt.load_exr(path)
    assert got.shape == (h, w, 4) and (got[:, :, 3] == 1.0).all()
    want = img.astype(np.float16).astype(np.float32) if half else img
    assert np.array_equal(got[:, :, :3], want)


def test_exr_unsupported_compressions_are_named(tmp_path):
    import cv2
    img = np.ones((8, 8, 3), np.float32)
    for name in ("PXR24", "B44", "DWAA"):
        path = str(tmp_path / f"x_{name}.exr")
        assert cv2.imwrite(path, img, [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_HALF, cv2.IMWRITE_EXR_COMPRESSION, getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + name)])
        with pytest.raises(rt.Rt2Error, match="not supported"):
            rt.load_exr(path)
