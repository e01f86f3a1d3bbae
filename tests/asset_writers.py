"""Test-side writers for the asset formats the host readers ingest (raytracer2_b200/host/rt2_ingest.cpp): GLB / glTF,
PNG and OpenEXR.  Independent of the readers (numpy + zlib + struct only), so a round trip checks the reader, not itself."""
import base64
import json
import struct
import zlib

import numpy as np


# ---- PNG ------------------------------------------------------------------------------------------------------------
def _png_chunk(tag: bytes, body: bytes) -> bytes:
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def write_png(pixels: np.ndarray, color_type: int, bit_depth: int = 8, palette=None, trns=None, filters=(0, 1, 2, 3, 4)) -> bytes:
    """pixels: uint8 [h, w, samples] (samples = 1, 2, 3, 4 for colour types 0/3, 4, 2, 6) or packed rows for bit_depth < 8
    given as [h, w] values < 2**bit_depth.  Row y uses filter type filters[y % len(filters)]."""
    h, w = pixels.shape[:2]
    if bit_depth == 8:
        rows = pixels.reshape(h, -1).astype(np.uint8)
        bpp = rows.shape[1] // w
    elif bit_depth == 16:
        rows = pixels.reshape(h, -1).astype(">u2").view(np.uint8).reshape(h, -1)
        bpp = rows.shape[1] // w
    else:
        per = 8 // bit_depth
        packed = np.zeros((h, (w + per - 1) // per), np.uint8)
        for x in range(w):
            packed[:, x // per] |= (pixels[:, x].astype(np.uint8) << (8 - bit_depth - (x % per) * bit_depth))
        rows, bpp = packed, 1
    raw = bytearray()
    prev = np.zeros(rows.shape[1], np.int32)
    for y in range(h):
        ft = filters[y % len(filters)]
        cur = rows[y].astype(np.int32)
        out = np.zeros_like(cur)
        for i in range(len(cur)):
            a = cur[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = (0, a, b, (a + b) >> 1, _paeth(int(a), int(b), int(c)))[ft]
            out[i] = (cur[i] - pred) & 0xFF
        raw.append(ft); raw.extend(out.astype(np.uint8).tobytes())
        prev = cur
    body = b"\x89PNG\r\n\x1a\n" + _png_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, bit_depth, color_type, 0, 0, 0))
    if palette is not None:
        body += _png_chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        body += _png_chunk(b"tRNS", np.asarray(trns, np.uint8).tobytes())
    comp = zlib.compress(bytes(raw), 6)
    half = len(comp) // 2                       # two IDAT chunks: readers must concatenate them
    return body + _png_chunk(b"IDAT", comp[:half]) + _png_chunk(b"IDAT", comp[half:]) + _png_chunk(b"IEND", b"")


# ---- glTF / GLB -------------------------------------------------------------------------------------------------------
class GltfBuilder:
    def __init__(self):
        self.bin = bytearray()
        self.j = {"asset": {"version": "2.0"}, "buffers": [{}], "bufferViews": [], "accessors": [], "meshes": [], "nodes": [],
                  "materials": [], "scenes": [{"nodes": []}], "scene": 0}

    def view(self, data: bytes, stride: int = 0) -> int:
        while len(self.bin) % 4:
            self.bin.append(0)
        bv = {"buffer": 0, "byteOffset": len(self.bin), "byteLength": len(data)}
        if stride:
            bv["byteStride"] = stride
        self.bin.extend(data)
        self.j["bufferViews"].append(bv)
        return len(self.j["bufferViews"]) - 1

    def accessor(self, arr: np.ndarray, type_: str, normalized=False, stride_pad: int = 0, byte_offset: int = 0) -> int:
        comp = {np.dtype(np.float32): 5126, np.dtype(np.uint8): 5121, np.dtype(np.uint16): 5123, np.dtype(np.uint32): 5125,
                np.dtype(np.int8): 5120, np.dtype(np.int16): 5122}[arr.dtype]
        a = np.ascontiguousarray(arr)
        elem = a.dtype.itemsize * (1 if a.ndim == 1 else a.shape[1])
        if stride_pad or byte_offset:                                       # interleave padding to exercise byteStride / accessor byteOffset
            stride = elem + stride_pad
            buf = bytearray(byte_offset + stride * len(a))
            for i in range(len(a)):
                buf[byte_offset + i * stride:byte_offset + i * stride + elem] = a[i].tobytes()
            v = self.view(bytes(buf), stride)
        else:
            v = self.view(a.tobytes())
        acc = {"bufferView": v, "componentType": comp, "count": len(a), "type": type_}
        if byte_offset:
            acc["byteOffset"] = byte_offset
        if normalized:
            acc["normalized"] = True
        if type_ == "VEC3" and comp == 5126:
            acc["min"] = [float(x) for x in a.min(axis=0)]; acc["max"] = [float(x) for x in a.max(axis=0)]
        self.j["accessors"].append(acc)
        return len(self.j["accessors"]) - 1

    def add_image_png(self, png: bytes, embed=True) -> int:
        self.j.setdefault("images", [])
        if embed:
            self.j["images"].append({"bufferView": self.view(png), "mimeType": "image/png"})
        else:
            self.j["images"].append({"uri": "data:image/png;base64," + base64.b64encode(png).decode()})
        return len(self.j["images"]) - 1

    def glb(self) -> bytes:
        self.j["buffers"][0] = {"byteLength": len(self.bin)}
        js = json.dumps(self.j, separators=(",", ":")).encode()
        js += b" " * (-len(js) % 4)
        b = bytes(self.bin) + b"\0" * (-len(self.bin) % 4)
        total = 12 + 8 + len(js) + 8 + len(b)
        return struct.pack("<4sII", b"glTF", 2, total) + struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(b), 0x004E4942) + b

    def gltf(self, bin_uri: str) -> tuple[bytes, bytes]:
        """(.gltf JSON bytes, .bin bytes); bin_uri = file name to reference, or 'data' for an embedded base64 buffer."""
        uri = "data:application/octet-stream;base64," + base64.b64encode(bytes(self.bin)).decode() if bin_uri == "data" else bin_uri
        self.j["buffers"][0] = {"byteLength": len(self.bin), "uri": uri}
        return json.dumps(self.j, indent=1).encode(), bytes(self.bin)


# ---- OpenEXR ------------------------------------------------------------------------------------------------------------
def _exr_attr(name: str, type_: str, data: bytes) -> bytes:
    return name.encode() + b"\0" + type_.encode() + b"\0" + struct.pack("<I", len(data)) + data


def _exr_pack_block(raw: bytes, compression: int) -> bytes:
    if compression == 0:
        return raw
    a = np.frombuffer(raw, np.uint8)
    n = len(a)
    re = np.concatenate([a[0::2], a[1::2]])                                    # reorder: even bytes, then odd bytes
    d = re.astype(np.int32)
    pred = np.empty(n, np.uint8)
    pred[0] = re[0]
    pred[1:] = ((d[1:] - d[:-1] + 128 + 256) & 0xFF).astype(np.uint8)          # delta predictor
    if compression in (2, 3):
        out = zlib.compress(pred.tobytes(), 6)
    else:                                                                      # RLE (compression 1)
        out = bytearray(); i = 0; p = pred.tobytes()
        while i < n:
            run = 1
            while i + run < n and run < 127 and p[i + run] == p[i]:
                run += 1
            if run >= 3:
                out += struct.pack("b", run - 1) + p[i:i + 1]; i += run
            else:
                j = i
                while j < n and j - i < 127 and not (j + 2 < n and p[j] == p[j + 1] == p[j + 2]):
                    j += 1
                out += struct.pack("b", -(j - i)) + p[i:j]; i = j
        out = bytes(out)
    return out if len(out) < len(raw) else raw                                 # the format stores the raw block when compression does not help


def write_exr(channels: dict, compression: int = 3, line_order: int = 0) -> bytes:
    """channels: name -> [h, w] array of dtype float16, float32 or uint32.  Scan-line, single part."""
    names = sorted(channels)
    h, w = channels[names[0]].shape
    types = {np.dtype(np.uint32): 0, np.dtype(np.float16): 1, np.dtype(np.float32): 2}
    chlist = b"".join(n.encode() + b"\0" + struct.pack("<iBBBBii", types[channels[n].dtype], 0, 0, 0, 0, 1, 1) for n in names) + b"\0"
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    header = struct.pack("<II", 20000630, 2)
    header += _exr_attr("channels", "chlist", chlist) + _exr_attr("compression", "compression", bytes([compression]))
    header += _exr_attr("dataWindow", "box2i", box) + _exr_attr("displayWindow", "box2i", box)
    header += _exr_attr("lineOrder", "lineOrder", bytes([line_order])) + _exr_attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
    header += _exr_attr("screenWindowCenter", "v2f", struct.pack("<ff", 0, 0)) + _exr_attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0"
    lines = {0: 1, 1: 1, 2: 1, 3: 16}[compression]
    blocks = []
    for y0 in range(0, h, lines):
        raw = b"".join(np.ascontiguousarray(channels[n][y]).tobytes() for y in range(y0, min(h, y0 + lines)) for n in names)
        data = _exr_pack_block(raw, compression)
        blocks.append(struct.pack("<ii", y0, len(data)) + data)
    order = range(len(blocks)) if line_order == 0 else reversed(range(len(blocks)))       # DECREASING_Y: blocks stored bottom-up, table still by y
    offsets = [0] * len(blocks)
    pos = len(header) + 8 * len(blocks)
    body = b""
    for b in order:
        offsets[b] = pos; body += blocks[b]; pos += len(blocks[b])
    return header + b"".join(struct.pack("<Q", o) for o in offsets) + body
