/* rt2_ingest.h — asset readers of the host side: what src/gltf/mod.rs::load_file (through the `gltf`
 * crate, Cargo.lock gltf 1.4.1) and image::open("skybox2.exr") (`image` 0.25 / `exr`) hand to Model::from_gltf and
 * Image::new_from_data.  C++ because the reference's host is compiled code and this image has no Rust toolchain.
 * The reference's own assets (src/models/box.glb, skybox2.exr) are git-ignored upstream, so these readers are tested on
 * files written by tests/ (round trips against the procedural scenes) and by independent encoders in this image — PIL
 * (libjpeg-turbo, zlib-ng) for PNG / JPEG, OpenCV (the OpenEXR library) for EXR — not against the reference's files.
 */
#ifndef RT2_INGEST_H
#define RT2_INGEST_H
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/rt2_abi.h"

namespace rt2ingest {

/* src/gltf/mod.rs:16-24 `Model` flattened the way src/model.rs:185-476 consumes it: one geometry per (node, primitive). */
struct GltfModel {
    std::vector<rt2_vertex> vertices;
    std::vector<uint32_t> indices;
    std::vector<rt2_geometry_info> geometries;     /* transform = the node's LOCAL matrix (mod.rs:129: no hierarchy walk) */
    std::vector<uint32_t> index_counts;
    std::vector<std::vector<uint8_t>> image_rgba8; /* src/gltf/image.rs:57-109: expanded to RGBA8 */
    std::vector<uint32_t> image_w, image_h;
    std::vector<rt2_sampler_desc> samplers;        /* [0] = default sampler (mod.rs:145-152), glTF sampler i at i + 1 */
    std::vector<rt2_texture_desc> textures;
};

/* .glb (binary container) or .gltf (JSON + external / data-URI buffers and images).  false + err on failure. */
bool load_gltf(const std::string& path, GltfModel& out, std::string& err);

/* 8-bit PNG (grey, grey+alpha, RGB, RGBA, palette; non-interlaced) -> the channel layout the `image` crate reports
 * (1, 2, 3 or 4 channels).  16-bit and interlaced files are rejected like src/gltf/image.rs:38 rejects wide formats. */
bool decode_png(const uint8_t* data, size_t size, std::vector<uint8_t>& pixels, uint32_t& w, uint32_t& h, uint32_t& channels, std::string& err);

/* Baseline / extended-sequential Huffman JPEG (ITU-T T.81 SOF0 / SOF1, 8 bits; grey or YCbCr; any 1-2 x 1-2 sampling; restart
 * intervals) -> 1 or 3 channels, as the `image` crate reports L8 / Rgb8.  Arithmetic pinned to libjpeg-turbo's defaults (integer
 * "islow" IDCT, fancy chroma upsampling, 16-bit fixed-point YCbCr -> RGB).  Progressive (SOF2), lossless, arithmetic-coded, CMYK
 * and 12-bit files are reported as unsupported. */
bool decode_jpeg(const uint8_t* data, size_t size, std::vector<uint8_t>& pixels, uint32_t& w, uint32_t& h, uint32_t& channels, std::string& err);

/* one PIZ-compressed block -> the bytes of the same block stored uncompressed (rt2_ingest_codecs.cpp) */
bool exr_piz_decode(const uint8_t* src, size_t n, const std::vector<int>& chan_words, uint32_t width, uint32_t rows, std::vector<uint8_t>& out, std::string& err);

/* OpenEXR scan-line image (NONE / RLE / ZIPS / ZIP / PIZ compression; HALF, FLOAT or UINT channels R,G,B[,A] or Y) ->
 * RGBA32F rows top to bottom, alpha 1 when absent: DynamicImage::to_rgba32f (src/context.rs:1476). */
bool load_exr(const std::string& path, std::vector<float>& rgba, uint32_t& w, uint32_t& h, std::string& err);

}  // namespace rt2ingest
#endif
