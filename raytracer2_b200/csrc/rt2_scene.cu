// rt2_scene.cu — scene upload (Model::from_gltf, src/model.rs:185-453; sky src/main.rs:145) and
// acceleration-structure build (src/model.rs:454-459, src/main.rs:153-191).
#include "rt2_internal.cuh"
#include "rt2_build.cuh"

#include <cmath>

namespace rt2 {

// shading records in slot order: same fp32 vertex attributes GetGeometryFromHit gathers (Hit.glsl:12-33)
__global__ void k_build_tri_shade(BuildInput in, const uint32_t* __restrict__ slot_to_prim, uint32_t n, TriShade* __restrict__ out) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t i = slot_to_prim[s];
    const uint32_t g = find_geometry(in.geom_tri_offset, in.num_geoms, i);
    const uint32_t p = i - in.geom_tri_offset[g];
    const rt2_geometry_info& gi = in.geoms[g];
    const rt2_vertex* v[3];
    for (int k = 0; k < 3; k++) v[k] = &in.vertices[gi.vertex_offset + in.indices[gi.index_offset + 3u * p + (uint32_t)k]];
    TriShade t;
    for (int c = 0; c < 3; c++) { t.n0[c] = v[0]->normal[c]; t.n1[c] = v[1]->normal[c]; t.n2[c] = v[2]->normal[c]; }
    for (int c = 0; c < 2; c++) { t.uv0[c] = v[0]->uv[c]; t.uv1[c] = v[1]->uv[c]; t.uv2[c] = v[2]->uv[c]; }
    for (int c = 0; c < 3; c++) { t.c0[c] = v[0]->color[c]; t.c1[c] = v[1]->color[c]; t.c2[c] = v[2]->color[c]; }
    out[s] = t;
}

// instance world boxes from the BLAS root box (8 corners through the 3x4), padded: the TLAS box is
// only a culling aid and must never be tighter than what object-space traversal can hit
__global__ void k_instance_boxes(const InstanceRec* __restrict__ inst, uint32_t n, V3 blo, V3 bhi, float4* lo, float4* hi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* m = inst[i].fwd;
    V3 l = v3(INFINITY, INFINITY, INFINITY), h = v3(-INFINITY, -INFINITY, -INFINITY);
    for (int c = 0; c < 8; c++) {
        V3 p = xform_point(m, v3((c & 1) ? bhi.x : blo.x, (c & 2) ? bhi.y : blo.y, (c & 4) ? bhi.z : blo.z));
        l = v3(fminf(l.x, p.x), fminf(l.y, p.y), fminf(l.z, p.z)); h = v3(fmaxf(h.x, p.x), fmaxf(h.y, p.y), fmaxf(h.z, p.z));
    }
    V3 e = (h - l) * 1e-4f + v3(1e-6f, 1e-6f, 1e-6f);
    // also cover |coordinate| * 2^-20 so fp32 rounding of the object-space ray cannot matter
    e = e + v3(fmaxf(fabsf(l.x), fabsf(h.x)), fmaxf(fabsf(l.y), fabsf(h.y)), fmaxf(fabsf(l.z), fabsf(h.z))) * 9.5367431640625e-07f;
    lo[i] = make_float4(l.x - e.x, l.y - e.y, l.z - e.z, 0.f);
    hi[i] = make_float4(h.x + e.x, h.y + e.y, h.z + e.z, 0.f);
}

// Flattened instancing: world-space box of (instance i / T, BLAS leaf slot i % T), padded like the instance boxes — the box is a
// culling aid only (the triangle is tested in object space) and must never be tighter than what that test can accept.
__global__ void k_flat_boxes(const InstanceRec* __restrict__ inst, const TriPos* __restrict__ tris, uint32_t T, uint64_t n, float4* lo, float4* hi) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t in = (uint32_t)(i / T), s = (uint32_t)(i % T);
    const float* m = inst[in].fwd;
    const TriPos t = tris[s];
    const V3 a = xform_point(m, v3(t.v0.x, t.v0.y, t.v0.z)), b = xform_point(m, v3(t.v1.x, t.v1.y, t.v1.z)), c = xform_point(m, v3(t.v2.x, t.v2.y, t.v2.z));
    const V3 l = v3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
    const V3 h = v3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
    V3 e = (h - l) * 1e-4f + v3(1e-6f, 1e-6f, 1e-6f);
    e = e + v3(fmaxf(fabsf(l.x), fabsf(h.x)), fmaxf(fabsf(l.y), fabsf(h.y)), fmaxf(fabsf(l.z), fabsf(h.z))) * 1.9073486328125e-06f;   // 2^-19
    lo[i] = make_float4(l.x - e.x, l.y - e.y, l.z - e.z, 0.f);
    hi[i] = make_float4(h.x + e.x, h.y + e.y, h.z + e.z, 0.f);
}
__global__ void k_flat_leaves(const uint32_t* __restrict__ out_prim, uint32_t T, uint64_t n, FlatLeaf* leaves) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = out_prim[i];
    leaves[i] = FlatLeaf{p / T, p % T};
}

// inverse of an affine 3x4 in double (adjugate / determinant), rounded to float — the same formula,
// in the same order, as the oracle's invert_affine (host code; compiled with -ffp-contract=off).
static void invert_affine(const float* m, float* out) {
    double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    double tx = m[3], ty = m[7], tz = m[11];
    double A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    double det = a * A + b * B + c * C;
    double id = 1.0 / det;
    double r00 = A * id, r01 = -(b * i - c * h) * id, r02 = (b * f - c * e) * id;
    double r10 = B * id, r11 = (a * i - c * g) * id, r12 = -(a * f - c * d) * id;
    double r20 = C * id, r21 = -(a * h - b * g) * id, r22 = (a * e - b * d) * id;
    out[0] = (float)r00; out[1] = (float)r01; out[2] = (float)r02; out[3] = (float)(-(r00 * tx + r01 * ty + r02 * tz));
    out[4] = (float)r10; out[5] = (float)r11; out[6] = (float)r12; out[7] = (float)(-(r10 * tx + r11 * ty + r12 * tz));
    out[8] = (float)r20; out[9] = (float)r21; out[10] = (float)r22; out[11] = (float)(-(r20 * tx + r21 * ty + r22 * tz));
}

template <class T>
static int upload(rt2_ctx* ctx, T** dst, const T* src, size_t n) {
    *dst = nullptr;
    RT2_CUDA(ctx, cudaMalloc((void**)dst, std::max<size_t>(n, 1) * sizeof(T)));
    ctx->device_bytes += std::max<size_t>(n, 1) * sizeof(T);
    if (n) RT2_CUDA(ctx, cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return RT2_OK;
}

void scene_release(rt2_ctx* ctx) {
    DeviceScene& s = ctx->scene;
    void* ptrs[] = {s.vertices, s.indices, s.geoms, s.geom_tri_offset, s.geom_shade, s.texels, s.textures, s.srgb_lut, s.sky, s.instances, s.tri_shade};
    for (void* p : ptrs) if (p) cudaFree(p);
    s.blas.release(); s.tlas.release(); s.flat.release();
    if (s.flat_leaf) cudaFree(s.flat_leaf);
    s = DeviceScene();
}

int scene_upload(rt2_ctx* ctx, const rt2_scene_desc* d) {
    if (!d) return fail(ctx, RT2_E_INVALID, "scene is NULL");
    if ((d->num_vertices && !d->vertices) || (d->num_indices && !d->indices) || (d->num_geometries && (!d->geometries || !d->index_counts)))
        return fail(ctx, RT2_E_INVALID, "scene has counts without pointers");
    if (d->num_geometries == 0xFFFFFFFFu) return fail(ctx, RT2_E_INVALID, "too many geometries");
    // validate ranges up front: the kernels index without bounds checks (like the reference's shaders)
    uint64_t total_tris = 0;
    for (uint32_t g = 0; g < d->num_geometries; g++) {
        const rt2_geometry_info& gi = d->geometries[g];
        uint32_t ic = d->index_counts[g];
        if ((uint64_t)gi.index_offset + ic > d->num_indices) return fail(ctx, RT2_E_INVALID, "geometry index range out of bounds");
        for (uint32_t k = 0; k < ic; k++)
            if ((uint64_t)gi.vertex_offset + d->indices[gi.index_offset + k] >= d->num_vertices) return fail(ctx, RT2_E_INVALID, "vertex index out of bounds");
        if (gi.base_color_texture_index >= (int32_t)d->num_textures && gi.base_color_texture_index > -1) return fail(ctx, RT2_E_INVALID, "texture index out of bounds");
        total_tris += ic / 3;
    }
    if (total_tris >= 0x7FFFFFFFull) return fail(ctx, RT2_E_INVALID, "too many triangles");
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    scene_release(ctx);
    DeviceScene& s = ctx->scene;
    int rc;
    s.num_vertices = d->num_vertices; s.num_indices = d->num_indices; s.num_geoms = d->num_geometries; s.num_tris = (uint32_t)total_tris;
    if ((rc = upload(ctx, &s.vertices, d->vertices, d->num_vertices))) return rc;
    if ((rc = upload(ctx, &s.indices, d->indices, d->num_indices))) return rc;
    if ((rc = upload(ctx, &s.geoms, d->geometries, d->num_geometries))) return rc;
    s.h_geoms.assign(d->geometries, d->geometries + d->num_geometries);
    s.h_index_counts.assign(d->index_counts, d->index_counts + d->num_geometries);
    std::vector<uint32_t> offs(d->num_geometries + 1, 0);
    std::vector<GeomShade> gs(d->num_geometries);
    s.num_lights = 0;
    for (uint32_t g = 0; g < d->num_geometries; g++) {
        const rt2_geometry_info& gi = d->geometries[g];
        offs[g + 1] = offs[g] + d->index_counts[g] / 3;
        GeomShade& o = gs[g];
        memcpy(o.transform, gi.transform, 64);
        for (int c = 0; c < 3; c++) { o.base_color[c] = gi.base_color[c]; o.emission12[c] = gi.emission[c] * 12.0f; o.emission[c] = gi.emission[c]; }
        o.texture = gi.base_color_texture_index; o.metallic = gi.metallic_factor; o.roughness = gi.roughness;
        if (gi.emission[0] != 0.0f || gi.emission[1] != 0.0f || gi.emission[2] != 0.0f) s.num_lights += d->index_counts[g] / 3;   // model.rs:411-413
    }
    if ((rc = upload(ctx, &s.geom_tri_offset, offs.data(), offs.size()))) return rc;
    if ((rc = upload(ctx, &s.geom_shade, gs.data(), gs.size()))) return rc;
    // textures: resolved (image, sampler) pairs over one RGBA8 texel pool
    {
        std::vector<TextureRec> recs(d->num_textures);
        std::vector<uint32_t> img_off(d->num_images, 0);
        size_t pool = 0;
        for (uint32_t i = 0; i < d->num_images; i++) {
            if (!d->images[i].rgba8 || !d->images[i].width || !d->images[i].height) return fail(ctx, RT2_E_INVALID, "empty image");
            img_off[i] = (uint32_t)pool; pool += (size_t)d->images[i].width * d->images[i].height;
            if (pool >= 0xFFFFFFFFull) return fail(ctx, RT2_E_INVALID, "texel pool too large");
        }
        std::vector<uint32_t> white(1, 0xFFFFFFFFu);
        RT2_CUDA(ctx, cudaMalloc((void**)&s.texels, (pool + 1) * 4)); ctx->device_bytes += (pool + 1) * 4;
        for (uint32_t i = 0; i < d->num_images; i++)
            RT2_CUDA(ctx, cudaMemcpyAsync(s.texels + img_off[i], d->images[i].rgba8, (size_t)d->images[i].width * d->images[i].height * 4, cudaMemcpyHostToDevice, ctx->stream));
        RT2_CUDA(ctx, cudaMemcpyAsync(s.texels + pool, white.data(), 4, cudaMemcpyHostToDevice, ctx->stream));
        for (uint32_t t = 0; t < d->num_textures; t++) {
            const rt2_texture_desc& td = d->textures[t];
            TextureRec r{(uint32_t)pool, 1, 1, RT2_FILTER_LINEAR};
            if (td.image_index < d->num_images) { r.offset = img_off[td.image_index]; r.width = d->images[td.image_index].width; r.height = d->images[td.image_index].height; }
            if (td.sampler_index < d->num_samplers) r.filter = d->samplers[td.sampler_index].mag_filter;   // implicit LOD 0 => magnification filter
            recs[t] = r;
        }
        s.num_textures = d->num_textures;
        if ((rc = upload(ctx, &s.textures, recs.data(), recs.size()))) return rc;
        float lut[256];
        for (int i = 0; i < 256; i++) { double c = i / 255.0; lut[i] = (float)(c <= 0.04045 ? c / 12.92 : std::pow((c + 0.055) / 1.055, 2.4)); }
        if ((rc = upload(ctx, &s.srgb_lut, lut, 256))) return rc;
    }
    if (d->sky_rgba32f && d->sky_width && d->sky_height) {
        s.sky_w = d->sky_width; s.sky_h = d->sky_height;
        if ((rc = upload(ctx, &s.sky, reinterpret_cast<const float4*>(d->sky_rgba32f), (size_t)s.sky_w * s.sky_h))) return rc;
    }
    if (d->num_instances && d->instance_transforms) {
        std::vector<InstanceRec> ir(d->num_instances);
        for (uint32_t i = 0; i < d->num_instances; i++) {
            memcpy(ir[i].fwd, d->instance_transforms + (size_t)i * 12, 48);
            invert_affine(ir[i].fwd, ir[i].inv);
        }
        s.num_instances = d->num_instances;
        if ((rc = upload(ctx, &s.instances, ir.data(), ir.size()))) return rc;
    }
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // host pointers are the caller's again
    return RT2_OK;
}

int scene_build_accel(rt2_ctx* ctx) {
    DeviceScene& s = ctx->scene;
    if (!s.geom_tri_offset) return fail(ctx, RT2_E_STATE, "rt2_build_accel before rt2_upload_scene");
    s.blas.release(); s.tlas.release(); s.flat.release();
    if (s.flat_leaf) { cudaFree(s.flat_leaf); s.flat_leaf = nullptr; }
    if (s.tri_shade) { cudaFree(s.tri_shade); s.tri_shade = nullptr; }
    s.accel_built = false;
    LaunchCounter lc;
    std::string err;
    BuildInput in{s.vertices, s.indices, s.geoms, s.geom_tri_offset, s.num_geoms, s.num_tris};
    int rc = build_blas(ctx->stream, in, &s.blas, &lc, err);
    if (rc != RT2_OK) return fail(ctx, rc, "BLAS build: " + err);
    RT2_CUDA(ctx, cudaMalloc((void**)&s.tri_shade, std::max<size_t>(s.num_tris, 1) * sizeof(TriShade)));
    if (s.num_tris) { k_build_tri_shade<<<(s.num_tris + 255) / 256, 256, 0, ctx->stream>>>(in, s.blas.slot_to_prim, s.num_tris, s.tri_shade); lc.add(1); }
    if (s.num_instances) {
        float4 *lo = nullptr, *hi = nullptr;
        RT2_CUDA(ctx, cudaMalloc((void**)&lo, (size_t)s.num_instances * 16));
        RT2_CUDA(ctx, cudaMalloc((void**)&hi, (size_t)s.num_instances * 16));
        k_instance_boxes<<<(s.num_instances + 255) / 256, 256, 0, ctx->stream>>>(s.instances, s.num_instances,
            v3(s.blas.box_lo[0], s.blas.box_lo[1], s.blas.box_lo[2]), v3(s.blas.box_hi[0], s.blas.box_hi[1], s.blas.box_hi[2]), lo, hi);
        lc.add(1);
        rc = build_cwbvh_from_boxes(ctx->stream, lo, hi, s.num_instances, &s.tlas.nodes, &s.tlas.num_nodes, &s.tlas.slot_to_prim, &s.tlas.depth, &lc, err);
        cudaFree(lo); cudaFree(hi);
        if (rc != RT2_OK) return fail(ctx, rc, "TLAS build: " + err);
        s.tlas.num_prims = s.num_instances;
        // Flattened instancing when it fits: n = instances x triangles primitives cost ~40 B each when built (80-byte node per ~3.7
        // primitives + 8-byte leaf record) and ~450 B each of build scratch (sort keys, BVH2, the reinsertion rounds' and the collapse's per-node state).  RT2_FLATTEN=0 keeps the two-level walk only,
        // RT2_FLATTEN=1 forces the attempt; the default takes it when the scratch fits in half of the free device memory.
        const uint64_t n_flat = (uint64_t)s.num_instances * s.num_tris;
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        int want = n_flat > 0 && n_flat < 0x7FFFFFF0ull && n_flat * 480ull < free_b / 2 ? 1 : 0;
        if (const char* e = getenv("RT2_FLATTEN")) want = atoi(e) != 0 && n_flat > 0 && n_flat < 0x7FFFFFF0ull;
        if (want) {
            RT2_CUDA(ctx, cudaMalloc((void**)&lo, n_flat * 16));
            RT2_CUDA(ctx, cudaMalloc((void**)&hi, n_flat * 16));
            k_flat_boxes<<<(uint32_t)((n_flat + 255) / 256), 256, 0, ctx->stream>>>(s.instances, s.blas.tris, s.num_tris, n_flat, lo, hi);
            lc.add(1);
            Node8* nodes = nullptr; uint32_t* out_prim = nullptr; uint32_t num_nodes = 0, depth = 0;
            rc = build_cwbvh_from_boxes(ctx->stream, lo, hi, (uint32_t)n_flat, &nodes, &num_nodes, &out_prim, &depth, &lc, err);
            cudaFree(lo); cudaFree(hi);
            if (rc != RT2_OK) return fail(ctx, rc, "flattened-instance BVH build: " + err);
            // the builder sizes the node array for the worst case (n + 1): keep only what was used
            RT2_CUDA(ctx, cudaMalloc((void**)&s.flat.nodes, (size_t)num_nodes * sizeof(Node8)));
            RT2_CUDA(ctx, cudaMemcpyAsync(s.flat.nodes, nodes, (size_t)num_nodes * sizeof(Node8), cudaMemcpyDeviceToDevice, ctx->stream));
            RT2_CUDA(ctx, cudaMalloc((void**)&s.flat_leaf, n_flat * sizeof(FlatLeaf)));
            k_flat_leaves<<<(uint32_t)((n_flat + 255) / 256), 256, 0, ctx->stream>>>(out_prim, s.num_tris, n_flat, s.flat_leaf);
            lc.add(1);
            RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            cudaFree(nodes); cudaFree(out_prim);
            s.flat.num_nodes = num_nodes; s.flat.num_prims = (uint32_t)n_flat; s.flat.depth = depth;
            ctx->device_bytes += (size_t)num_nodes * sizeof(Node8) + n_flat * sizeof(FlatLeaf);
        }
    }
    RT2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    RT2_CUDA(ctx, cudaGetLastError());
    s.accel_built = true;
    ctx->stats.kernel_launches = lc.n;
    ctx->stats.bvh_nodes = (uint64_t)s.blas.num_nodes + s.tlas.num_nodes + s.flat.num_nodes;
    ctx->stats.bvh_triangles = s.num_tris;
    return RT2_OK;
}

}  // namespace rt2
