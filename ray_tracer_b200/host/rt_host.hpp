// rt_host.hpp — C++ mirror of the reference's host side of the path, above the C ABI.
//
// The reference is a Rust crate; no Rust toolchain exists in this image, so the host
// layer a Rust user would keep (scene/camera/material/mesh types, Context::update/render
// frame logic, model loaders) is mirrored here in C++ with the same names, argument
// meaning and arithmetic, and drives librt_b200.so exactly as the Rust FFI stub in
// INTEGRATION.md would.  Citations are to /root/reference/src/core/.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/rt_b200.h"

namespace rtb {

struct Vec3 { float x, y, z; };
struct Vec4 { float x, y, z, w; };

// glam 0.23 scalar Vec3 operations (operation order matters for bit-equal uniforms)
Vec3 operator+(Vec3 a, Vec3 b);
Vec3 operator-(Vec3 a, Vec3 b);
Vec3 operator*(float s, Vec3 a);
Vec3 operator/(Vec3 a, float s);
float dot(Vec3 a, Vec3 b);
Vec3 cross(Vec3 a, Vec3 b);
float length(Vec3 a);
Vec3 normalize(Vec3 a);

// scene.rs:46-58 — Sphere::new (smoothness = specular clamped to <= 1, -1 passes = glass)
rt_sphere Sphere(Vec3 position, float radius, Vec4 color, Vec4 emission_color, float emission_strength, float specular);
// scene.rs:61-77 — Mesh::new
rt_mesh Mesh(Vec3 pos, uint32_t first, uint32_t triangles, uint32_t offset, Vec4 color, Vec4 emission_color,
             float emission_strength, float specular);
// scene.rs:88-92 — Vertex::new
rt_vertex Vertex(Vec3 pos, Vec3 normal);

// camera.rs:171-247 — key / mouse / scroll state of the fly-cam (winit key codes reduced to
// the ones process_keyboard matches)
enum class Key { W, S, A, D, Up, Down, Left, Right, Space, LShift, Other };
struct CameraController {
    float amount_left = 0, amount_right = 0, amount_forward = 0, amount_backward = 0, amount_up = 0, amount_down = 0;
    float rotate_horizontal = 0, rotate_vertical = 0, scroll = 0;
    float speed = 3.0f, sensitivity = 0.35f;                       // CameraController::new(3.0, 0.35), camera.rs:57
    bool process_keyboard(Key key, bool pressed);                  // camera.rs:201-232: 5.0 while pressed
    void process_mouse(double mouse_dx, double mouse_dy);          // :233-236: x3
    bool process_scroll_lines(float lines);                        // :238-247 LineDelta: -(lines * 10000)
    bool process_scroll_pixels(double pixels);                     //          PixelDelta: -pixels
};

// camera.rs:31-47 (state)
struct Camera {
    Vec3 origin{0, 0, 0}, look_at{0, 0, -1}, view_up{0, 1, 0};
    float fov = 45.0f, aspect = 1.0f, near_ = 0.1f, far_ = 100.0f, aperture = 0.0f, focus_dist = 0.1f;
    CameraController controller;
    // camera.rs:122-165 — one fly-cam step of dt seconds: yaw/pitch re-derived from look_at every
    // call, x/z swapped in the scroll direction and the pitch clamp compared in degrees, as upstream
    void update_camera(float dt);
    Camera() = default;
    // camera.rs:49-78 — Camera::new(origin, look_at, view_up, fov, aspect, near, far, aperture, focus_dist)
    Camera(Vec3 origin, Vec3 look_at, Vec3 view_up, float fov, float aspect, float near_, float far_, float aperture,
           float focus_dist);
    // camera.rs:79-121
    rt_camera_uniform to_uniform() const;
};
// camera.rs:248-250
float radians(float deg);

// scene.rs:94-100
struct Scene {
    Camera camera;
    std::vector<rt_sphere> spheres;
    std::vector<rt_vertex> vertices;
    std::vector<uint32_t> indices;
    std::vector<rt_mesh> meshes;

    static Scene empty(float aspect);                          // scene.rs:103-119
    static Scene random_balls(float aspect, uint32_t seed);    // scene.rs:121-197 (thread_rng -> fixed stream)
    static Scene room(float aspect);                           // scene.rs:198-310
    static Scene metal(float aspect);                          // scene.rs:311-378
    static Scene balls(float aspect);                          // scene.rs:379-476
    static Scene by_id(int id, float aspect);                  // context.rs:261-278: 0 balls, 1 random_balls, 2 room, 3 metal
    void add_dummy_mesh();                                     // the 0-triangle mesh of scene.rs:432-445
};

// resource.rs:27-45 — dispatch on extension; appends to the three arrays. Throws std::runtime_error.
void load_model(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices,
                std::vector<rt_mesh>& meshes);
void load_model_obj(const std::string& path, std::vector<rt_vertex>& v, std::vector<uint32_t>& i, std::vector<rt_mesh>& m);   // :47-107
void load_model_gltf(const std::string& path, std::vector<rt_vertex>& v, std::vector<uint32_t>& i, std::vector<rt_mesh>& m);  // :109-192
void load_model_glb(const std::string& path, std::vector<rt_vertex>& v, std::vector<uint32_t>& i, std::vector<rt_mesh>& m);   // :194-278
// Texture-extension loaders (the reference drops `vt` / TEXCOORD_0 and never opens an image):
// keep_uv = true stores the texture coordinates in the Vertex padding words (OBJ v flipped so
// that image row 0 is the top row).
void load_model_uv(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices,
                   std::vector<rt_mesh>& meshes);
// Scene-graph import of a .glb or .gltf (extension): walks the default scene, applies every node's transform
// (matrix or T*R*S, parents first) to positions and normals, one Mesh record per primitive INSTANCE at
// pos = 0, colour / smoothness / emission from the material (baseColorFactor, 1 - roughnessFactor,
// emissiveFactor), texture handles = image index + 1 (the order of load_glb_images), UVs kept.
// An .obj path imports the file's .mtl instead: one Mesh per `usemtl` run (Kd, Ke, Ns -> colour, emission, smoothness;
// map_Kd / map_Bump -> texture handles in the order load_glb_images returns the images).
void load_scene_glb(const std::string& path, std::vector<rt_vertex>& vertices, std::vector<uint32_t>& indices,
                    std::vector<rt_mesh>& meshes);
struct Image { uint32_t width = 0, height = 0; std::vector<uint8_t> rgba; };
Image load_png(const std::string& path);                            // grey (1-16 bit), grey+alpha, RGB, RGBA, palette (1-8 bit); Adam7 interlace
Image decode_png(const uint8_t* data, size_t size);
Image decode_jpeg(const uint8_t* data, size_t size);                // sequential and progressive DCT JPEG (8 bit, Huffman); arithmetic / lossless files throw
Image decode_image(const uint8_t* data, size_t size);               // PNG or JPEG by signature
Image load_image(const std::string& path);
// images[] of a .glb (bufferViews of the binary chunk) or a .gltf (bufferViews, sibling files, base64 data: URIs), in order;
// entries this decoder does not read (other formats, arithmetic-coded JPEG, missing files) come back empty
std::vector<Image> load_glb_images(const std::string& path);
void write_png(const std::string& path, const uint8_t* rgba, uint32_t width, uint32_t height);
// Image comparison (rt_imgdiff.cpp; SURVEY §8(c) tiers T2/T3, §8(f3)): RMSE over n_values floats; mean LDR-FLIP
// (Andersson et al. 2020) of two sRGB-encoded [0,1] RGB images, optional per-pixel error map (width*height floats).
double image_rmse(const float* a, const float* b, size_t n_values);
double image_flip(const float* ref_srgb, const float* test_srgb, uint32_t width, uint32_t height, double pixels_per_degree,
                  float* error_map);
// the hard-coded record each loaded primitive gets (resource.rs:73-85)
rt_mesh loader_mesh(uint32_t mesh_count, uint32_t n_indices, uint32_t first, uint32_t offset);

// context.rs:23-37 — the part of Context on the path: params, scene, the compute pipeline,
// frames/accumulate logic. Surface, swapchain, blit and imgui are out of scope.
class Context {
public:
    rt_params params{};
    Scene scene;
    int selected_scene = 0;

    // context.rs:40-124 — defaults bounces=3, rpp=1, skybox=0, frames=0, accumulate=1, Scene::balls
    Context(int cuda_device, uint32_t width, uint32_t height);
    ~Context();
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;

    void resize(uint32_t width, uint32_t height);   // context.rs:126-142
    void clear_accululation();                      // context.rs:143-146 (spelling as upstream)
    void update();                                  // context.rs:176-187 without the fly-cam step (dt = 0 leaves the camera's look_at re-derivation out)
    void update(float dt);                          // context.rs:176-187: camera.update_camera(dt), to_uniform, frames += 1
    // context.rs:148-175 — every window event resets the accumulation, then keyboard / wheel /
    // left button go to the controller; returns what upstream's `input` returns
    bool input_key(Key key, bool pressed);
    bool input_scroll_lines(float lines);
    bool input_mouse_button(bool left_pressed);
    bool input_other();
    void mouse_motion(double dx, double dy);        // lib.rs:25-27: only while the left button is held
    bool mouse_pressed = false;
    // What the blit pass + sRGB swapchain show (render.wgsl:34-39, context.rs:75-80), written as a PNG:
    // clamp, sRGB OETF, 8 bit; top row first; mirror_x reproduces the on-screen flip of renderer.rs:35-38
    void save_png(const std::string& path, bool mirror_x = false);
    void render();                                  // context.rs:196-208 — the compute pass only
    void set_scene(int id);                         // context.rs:259-283
    void set_scene(Scene s);                        // RayTracer::new on a caller-built scene
    // N redraws (update + render each), fused on the device when nothing changes between them
    void run_frames(uint32_t n);
    void sync();
    // Several GPUs from one process (no reference counterpart): every shard context gets the same scene,
    // inputs and update()/run_frames() calls and renders only its share of the tiles / frames (rt_set_shard);
    // combine() folds the shards' images into shards[0]'s combined image with one kernel over NVLink peer memory
    // (rt_combine_peers); until shards[0] renders again, its read_texture() / save_png() show that image.  The shards'
    // own accumulation images are untouched, so the same shards can go on rendering and be combined again.
    void set_shard(uint32_t tile_rank, uint32_t tile_world, uint32_t frame_rank, uint32_t frame_world);
    static void combine(const std::vector<Context*>& shards);
    std::vector<float> read_texture();              // RGBA32F, row 0 = bottom
    rt_ctx* handle() { return ctx_; }

private:
    rt_ctx* ctx_ = nullptr;
    std::vector<rt_ctx*> combined_from_;            // non-empty: the last combine's shard set, still current
    void check(int rc, const char* what);
    void upload();
};

} // namespace rtb
