/* rt2_host.h — host side above the C ABI, in C++ because the reference's host is compiled code
 * (Rust; no Rust toolchain in this image).  Mirrors the reference's host-side arithmetic for the
 * path: camera / view constants (src/camera.rs:99-157), the default GConst (src/main.rs:237-400)
 * and its per-frame mutation (src/main.rs:641-651), the three pure helpers of
 * src/light_passes.rs:671-731, and deterministic procedural generators that write scenes straight
 * into the reference's upload layout (what src/gltf/mod.rs::load_file produces).
 * Exposed with C linkage so tests and bench.py can drive it through ctypes.
 */
#ifndef RT2_HOST_H
#define RT2_HOST_H
#include "../../include/rt2_abi.h"
#include "../../include/rt2_gconst.h"
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

/* src/camera.rs:16-24 `Camera` */
typedef struct rt2h_camera {
    float position[3];
    float direction[3];   /* normalised by rt2h_camera_new like Camera::new (camera.rs:27-44) */
    float fov;            /* degrees */
    float aspect_ratio;
    float z_near, z_far;
} rt2h_camera;

void rt2h_camera_new(rt2h_camera* c, const float position[3], const float direction[3], float fov,
                     float aspect_ratio, float z_near, float z_far);
/* Camera::view_matrix (camera.rs:99-101, glam look_at_rh with UP=(0,-1,0), camera.rs:14) */
void rt2h_camera_view_matrix(const rt2h_camera* c, float out_col_major[16]);
/* perspective() (camera.rs:146-157) */
void rt2h_perspective(float fovy_rad, float aspect, float z_near, float z_far, float out_col_major[16]);
/* Camera::planar_view_constants (camera.rs:111-142); window size is a parameter here
 * (the reference hard-codes WINDOW_SIZE = 1920x1080, src/main.rs:58). */
void rt2h_planar_view_constants(const rt2h_camera* c, uint32_t width, uint32_t height, rt2_planar_view_constants* out);

/* src/light_passes.rs:671-698; returns bytes written (2*count) */
uint32_t rt2h_fill_neighbor_offset_buffer(uint32_t neighbor_offset_count, uint8_t* out);
/* src/light_passes.rs:700-716 */
void rt2h_compute_pdf_texture_size(uint32_t max_items, uint32_t* w, uint32_t* h, uint32_t* mips);
/* src/light_passes.rs:718-731 */
void rt2h_calculate_reservoir_buffer_parameters(uint32_t w, uint32_t h, rt2_reservoir_buffer_params* out);

/* Default GConst: src/main.rs:237-400 (incl. NEIGHBOR_OFFSET_COUNT = 8192, main.rs:57), with the
 * size-dependent fields derived from width/height/num_lights/env size exactly as main.rs does. */
void rt2h_default_gconst(rt2_gconst* g, const rt2h_camera* cam, uint32_t width, uint32_t height,
                         uint32_t num_lights, uint32_t env_w, uint32_t env_h);
/* per-frame mutation, src/main.rs:641-651: prev_view <- view, view <- camera, frame++, fresh
 * uniform_random_numbers (here: a counter-based hash of frame so runs are repeatable). */
void rt2h_advance_frame(rt2_gconst* g, const rt2h_camera* cam, uint32_t width, uint32_t height, uint32_t frame);

/* ---- procedural scenes ------------------------------------------------- */
typedef struct rt2h_scene rt2h_scene;
enum rt2h_scene_kind {
    RT2H_SCENE_CORNELL   = 1,  /* C1: 32 tris, 8 geometries, 1 area light                       */
    RT2H_SCENE_HALL      = 2,  /* C2/C4: 64 geometries x 4096 tris, 64 textured materials        */
    RT2H_SCENE_MANYLIGHT = 3,  /* C3: 1M tris, 200 geometries, 10k emissive triangles            */
    RT2H_SCENE_INSTANCED = 5,  /* C5: 2048-tri BLAS x 10k instances + procedural sky             */
    RT2H_SCENE_FURNACE   = 6,  /* closed white box, uniform emission (analytic test)             */
    RT2H_SCENE_RANDOM    = 7   /* random triangle soup (property tests)                          */
};
/* `scale` shrinks the big scenes for tests: triangle counts scale ~ scale^2 (1.0 = BASELINE size);
 * texture_size 0 => default (512 for HALL). */
rt2h_scene* rt2h_scene_generate(int kind, float scale, uint32_t texture_size, uint64_t seed);
/* SURVEY §8f N4 — the reference's asset path: gltf::load_file("./src/models/box.glb") (src/main.rs:62, src/gltf/mod.rs:50-168)
 * and image::open("./src/models/skybox2.exr") -> to_rgba32f (src/main.rs:63, src/context.rs:1476), flattened into the
 * upload layout Model::from_gltf consumes.  gltf_path: .glb or .gltf; sky_exr_path may be NULL.  The camera is the
 * reference's start-up camera (src/main.rs:114-121).  Returns NULL and writes a message into err (if given) on failure. */
rt2h_scene* rt2h_scene_load_gltf(const char* gltf_path, const char* sky_exr_path, char* err, size_t err_len);
/* the EXR reader alone: RGBA32F rows, top to bottom; caller frees with rt2h_free */
float*      rt2h_load_exr(const char* path, uint32_t* width, uint32_t* height, char* err, size_t err_len);
/* the image decoders alone (PNG or baseline JPEG, by magic bytes): `channels` x 8-bit samples as the `image` crate reports them
 * (src/gltf/image.rs:57-109 expands them to RGBA8); caller frees with rt2h_free */
uint8_t*    rt2h_decode_image(const uint8_t* data, size_t size, uint32_t* width, uint32_t* height, uint32_t* channels, char* err, size_t err_len);
void        rt2h_free(void* p);
void        rt2h_scene_destroy(rt2h_scene* s);
const rt2_scene_desc* rt2h_scene_desc(const rt2h_scene* s);
/* recommended camera for the scene */
void        rt2h_scene_camera(const rt2h_scene* s, float aspect_ratio, rt2h_camera* out);
/* number of emissive triangles (Model.lights, src/model.rs:411-413) */
uint32_t    rt2h_scene_num_lights(const rt2h_scene* s);
/* Gives any scene the instanced scene's procedural sky (equirect gradient + sun, width x width/2): the reference always has a
 * sky box (src/main.rs:63), the closed test scenes have none.  Invalidates nothing: the descriptor is refreshed in place. */
void        rt2h_scene_add_procedural_sky(rt2h_scene* s, uint32_t width);
uint64_t    rt2h_scene_num_triangles(const rt2h_scene* s);   /* instanced: BLAS tris x instances */

#ifdef __cplusplus
}
#endif
#endif
