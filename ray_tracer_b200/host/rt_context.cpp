// rt_context.cpp — Context (frames/accumulate logic + the compute pass) over the C ABI,
// and flat C wrappers (rth_*) so Python tests and bench.py can drive this C++ host layer.
#include <cstring>
#include <stdexcept>
#include <string>
#include "rt_host.hpp"

namespace rtb {

void Context::check(int rc, const char* what)
{
    if (rc != RT_OK) throw std::runtime_error(std::string(what) + ": " + rt_last_error(ctx_));
}

Context::Context(int cuda_device, uint32_t width, uint32_t height)
{
    int rc = rt_create(&ctx_, cuda_device, width, height);
    if (rc != RT_OK) throw std::runtime_error(std::string("rt_create: ") + rt_last_error(nullptr));
    // context.rs:86-94
    params.width = width; params.height = height;
    params.number_of_bounces = 3; params.rays_per_pixel = 1; params.skybox = 0; params.frames = 0; params.accumulate = 1;
    scene = Scene::balls((float)width / (float)height);        // context.rs:105
    try { upload(); } catch (...) { rt_destroy(ctx_); ctx_ = nullptr; throw; }
}

Context::~Context() { rt_destroy(ctx_); }

void Context::upload()
{
    // RayTracer::new: the four scene buffers (ray_tracer.rs:97-130)
    check(rt_upload_scene(ctx_, scene.spheres.data(), (uint32_t)scene.spheres.size(), scene.vertices.data(),
                          (uint32_t)scene.vertices.size(), scene.indices.data(), (uint32_t)scene.indices.size(),
                          scene.meshes.data(), (uint32_t)scene.meshes.size()), "rt_upload_scene");
}

void Context::resize(uint32_t width, uint32_t height)
{
    if (width == 0 || height == 0) return;                     // context.rs:127
    scene.camera.aspect = (float)width / (float)height;        // :130
    check(rt_resize(ctx_, width, height), "rt_resize");        // :132 new texture
    combined_from_.clear();
    params.width = width; params.height = height;
    params.frames = -1;                                        // :136
    check(rt_set_params(ctx_, &params), "rt_set_params");      // :138
}

void Context::clear_accululation()
{
    params.frames = -1;                                        // context.rs:144
    check(rt_set_params(ctx_, &params), "rt_set_params");
}

void Context::update()
{
    rt_camera_uniform u = scene.camera.to_uniform();           // context.rs:179
    if (params.accumulate != 0) params.frames += 1;            // :180-184
    else params.frames = -1;
    check(rt_set_camera(ctx_, &u), "rt_set_camera");           // :185
    check(rt_set_params(ctx_, &params), "rt_set_params");      // :186
}

void Context::update(float dt)
{
    scene.camera.update_camera(dt);                            // context.rs:178
    update();
}

bool Context::input_key(Key key, bool pressed) { clear_accululation(); return scene.camera.controller.process_keyboard(key, pressed); }
bool Context::input_scroll_lines(float lines) { clear_accululation(); return scene.camera.controller.process_scroll_lines(lines); }
bool Context::input_mouse_button(bool left_pressed) { clear_accululation(); mouse_pressed = left_pressed; return true; }
bool Context::input_other() { clear_accululation(); return false; }                       // context.rs:149: any window event resets
void Context::mouse_motion(double dx, double dy) { if (mouse_pressed) scene.camera.controller.process_mouse(dx, dy); }

void Context::save_png(const std::string& path, bool mirror_x)
{
    const uint32_t W = params.width, H = params.height;
    std::vector<uint8_t> px((size_t)W * H * 4), out((size_t)W * H * 4);
    if (!combined_from_.empty()) check(rt_combine_peers_rgba8(combined_from_.data(), (uint32_t)combined_from_.size(), nullptr, px.data(), px.size()), "rt_combine_peers_rgba8");
    else check(rt_read_rgba8(ctx_, px.data(), px.size()), "rt_read_rgba8");
    for (uint32_t y = 0; y < H; y++)                           // texture row 0 is the bottom image row
        for (uint32_t x = 0; x < W; x++)
            memcpy(&out[((size_t)(H - 1 - y) * W + (mirror_x ? W - 1 - x : x)) * 4], &px[((size_t)y * W + x) * 4], 4);
    write_png(path, out.data(), W, H);
}

void Context::render()
{
    combined_from_.clear();
    check(rt_dispatch(ctx_), "rt_dispatch");                   // context.rs:196-208
}

void Context::run_frames(uint32_t n)
{
    if (n == 0) return;
    if (params.accumulate == 0) {                              // frames pinned to -1: every redraw identical
        for (uint32_t i = 0; i < n; i++) { update(); render(); }
        return;
    }
    // n x (update; render) with frames = f+1 .. f+n, fused on the device
    rt_camera_uniform u = scene.camera.to_uniform();
    int32_t first = params.frames + 1;
    params.frames += (int32_t)n;
    check(rt_set_camera(ctx_, &u), "rt_set_camera");
    check(rt_set_params(ctx_, &params), "rt_set_params");
    combined_from_.clear();
    check(rt_render_frames(ctx_, first, n), "rt_render_frames");
}

void Context::set_shard(uint32_t tile_rank, uint32_t tile_world, uint32_t frame_rank, uint32_t frame_world)
{
    check(rt_set_shard(ctx_, tile_rank, tile_world, frame_rank, frame_world), "rt_set_shard");
}

void Context::combine(const std::vector<Context*>& shards)
{
    if (shards.empty() || !shards[0]) throw std::runtime_error("Context::combine: no shards");
    std::vector<rt_ctx*> h;
    for (Context* c : shards) { if (!c) throw std::runtime_error("Context::combine: null shard"); h.push_back(c->ctx_); }
    shards[0]->check(rt_combine_peers(h.data(), (uint32_t)h.size(), nullptr), "rt_combine_peers");
    shards[0]->combined_from_ = h;
}

void Context::set_scene(int id)
{
    set_scene(Scene::by_id(id, (float)params.width / (float)params.height));   // context.rs:261-278
    selected_scene = id;
}

void Context::set_scene(Scene s)
{
    scene = std::move(s);
    upload();                                                  // context.rs:279 RayTracer::new
    params.frames = -1;                                        // :280
    check(rt_set_params(ctx_, &params), "rt_set_params");
}

void Context::sync() { check(rt_sync(ctx_), "rt_sync"); }

std::vector<float> Context::read_texture()
{
    std::vector<float> out((size_t)params.width * params.height * 4);
    if (!combined_from_.empty()) check(rt_read_combined(ctx_, out.data(), out.size()), "rt_read_combined");
    else check(rt_read_accum(ctx_, out.data(), out.size()), "rt_read_accum");
    return out;
}

} // namespace rtb

// ---- flat C wrappers -------------------------------------------------------------------
static thread_local std::string g_err;
#define RTH_TRY(body) try { body; return 0; } catch (const std::exception& e) { g_err = e.what(); return -1; }

extern "C" {

const char* rth_last_error(void) { return g_err.c_str(); }

int rth_camera_to_uniform(const float* origin, const float* look_at, const float* up, float fov, float aspect, float near_,
                          float far_, float aperture, float focus_dist, rt_camera_uniform* out)
{
    RTH_TRY(*out = rtb::Camera({origin[0], origin[1], origin[2]}, {look_at[0], look_at[1], look_at[2]}, {up[0], up[1], up[2]},
                               fov, aspect, near_, far_, aperture, focus_dist).to_uniform());
}

// scene handles ---------------------------------------------------------------------------
void* rth_scene_builtin(int id, float aspect, uint32_t seed)
{
    try {
        if (id == 1) return new rtb::Scene(rtb::Scene::random_balls(aspect, seed));
        if (id == -1) return new rtb::Scene(rtb::Scene::empty(aspect));
        return new rtb::Scene(rtb::Scene::by_id(id, aspect));
    } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void* rth_scene_from_model(const char* path)
{
    try {
        rtb::Scene* s = new rtb::Scene();
        try { rtb::load_model(path, s->vertices, s->indices, s->meshes); } catch (...) { delete s; throw; }
        return s;
    } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void* rth_scene_from_model_uv(const char* path)               // texture extension: keeps vt / TEXCOORD_0
{
    try {
        rtb::Scene* s = new rtb::Scene();
        try { rtb::load_model_uv(path, s->vertices, s->indices, s->meshes); } catch (...) { delete s; throw; }
        return s;
    } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void* rth_scene_from_glb_graph(const char* path)             // scene-graph import: node transforms + materials (extension)
{
    try {
        rtb::Scene* s = new rtb::Scene();
        try { rtb::load_scene_glb(path, s->vertices, s->indices, s->meshes); } catch (...) { delete s; throw; }
        return s;
    } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
// PNG in / out. rth_png_load returns a handle (NULL on error); *w, *h receive the size.
void* rth_png_load(const char* path, uint32_t* w, uint32_t* h)
{
    try { rtb::Image* im = new rtb::Image(rtb::load_image(path)); *w = im->width; *h = im->height; return im; }
    catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void rth_image_copy(const void* im, uint8_t* rgba) { const rtb::Image* i = (const rtb::Image*)im; if (!i->rgba.empty()) memcpy(rgba, i->rgba.data(), i->rgba.size()); }
void rth_image_free(void* im) { delete (rtb::Image*)im; }
int rth_png_write(const char* path, const uint8_t* rgba, uint32_t w, uint32_t h) { RTH_TRY(rtb::write_png(path, rgba, w, h)); }
// images embedded in a .glb: returns a handle to the list, *n = count; rth_glb_image gives entry k (may be 0x0 for non-PNG)
void* rth_glb_images(const char* path, uint32_t* n)
{
    try { auto* v = new std::vector<rtb::Image>(rtb::load_glb_images(path)); *n = (uint32_t)v->size(); return v; }
    catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void rth_glb_image(const void* list, uint32_t k, uint32_t* w, uint32_t* h, uint8_t* rgba)
{
    const rtb::Image& i = ((const std::vector<rtb::Image>*)list)->at(k);
    *w = i.width; *h = i.height;
    if (rgba && !i.rgba.empty()) memcpy(rgba, i.rgba.data(), i.rgba.size());
}
double rth_image_rmse(const float* a, const float* b, size_t n_values) { return rtb::image_rmse(a, b, n_values); }
int rth_image_flip(const float* ref_srgb, const float* test_srgb, uint32_t w, uint32_t h, double ppd, float* error_map, double* mean)
{
    RTH_TRY(*mean = rtb::image_flip(ref_srgb, test_srgb, w, h, ppd, error_map));
}
void rth_glb_images_free(void* list) { delete (std::vector<rtb::Image>*)list; }
void rth_scene_free(void* s) { delete (rtb::Scene*)s; }
void rth_scene_counts(const void* h, uint32_t* ns, uint32_t* nv, uint32_t* ni, uint32_t* nm)
{
    const rtb::Scene* s = (const rtb::Scene*)h;
    *ns = (uint32_t)s->spheres.size(); *nv = (uint32_t)s->vertices.size(); *ni = (uint32_t)s->indices.size(); *nm = (uint32_t)s->meshes.size();
}
void rth_scene_copy(const void* h, rt_camera_uniform* cam, rt_sphere* sp, rt_vertex* vs, uint32_t* ix, rt_mesh* ms)
{
    const rtb::Scene* s = (const rtb::Scene*)h;
    if (cam) *cam = s->camera.to_uniform();
    if (sp && !s->spheres.empty()) memcpy(sp, s->spheres.data(), s->spheres.size() * sizeof(rt_sphere));
    if (vs && !s->vertices.empty()) memcpy(vs, s->vertices.data(), s->vertices.size() * sizeof(rt_vertex));
    if (ix && !s->indices.empty()) memcpy(ix, s->indices.data(), s->indices.size() * sizeof(uint32_t));
    if (ms && !s->meshes.empty()) memcpy(ms, s->meshes.data(), s->meshes.size() * sizeof(rt_mesh));
}
void rth_loader_mesh(uint32_t mesh_count, uint32_t n_indices, uint32_t first, uint32_t offset, rt_mesh* out)
{
    *out = rtb::loader_mesh(mesh_count, n_indices, first, offset);
}

// Context ----------------------------------------------------------------------------------
void* rth_context_new(int device, uint32_t w, uint32_t h)
{
    try { return new rtb::Context(device, w, h); } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void rth_context_free(void* c) { delete (rtb::Context*)c; }
int rth_context_update(void* c) { RTH_TRY(((rtb::Context*)c)->update()); }
int rth_context_update_dt(void* c, float dt) { RTH_TRY(((rtb::Context*)c)->update(dt)); }
int rth_context_input_key(void* c, int key, int pressed) { RTH_TRY(return ((rtb::Context*)c)->input_key((rtb::Key)key, pressed != 0) ? 1 : 0); }
int rth_context_input_scroll_lines(void* c, float lines) { RTH_TRY(return ((rtb::Context*)c)->input_scroll_lines(lines) ? 1 : 0); }
int rth_context_input_mouse_button(void* c, int pressed) { RTH_TRY(return ((rtb::Context*)c)->input_mouse_button(pressed != 0) ? 1 : 0); }
int rth_context_input_other(void* c) { RTH_TRY(return ((rtb::Context*)c)->input_other() ? 1 : 0); }
int rth_context_mouse_motion(void* c, double dx, double dy) { RTH_TRY(((rtb::Context*)c)->mouse_motion(dx, dy)); }
void rth_context_get_camera(void* c, float* origin3_look_at3) {
    const rtb::Camera& k = ((rtb::Context*)c)->scene.camera;
    origin3_look_at3[0] = k.origin.x; origin3_look_at3[1] = k.origin.y; origin3_look_at3[2] = k.origin.z;
    origin3_look_at3[3] = k.look_at.x; origin3_look_at3[4] = k.look_at.y; origin3_look_at3[5] = k.look_at.z;
}
int rth_context_save_png(void* c, const char* path, int mirror_x) { RTH_TRY(((rtb::Context*)c)->save_png(path, mirror_x != 0)); }
// stand-alone fly-cam step for CPU tests: state = origin3, look_at3, then the controller's 11 floats
int rth_camera_update(float* state17, float dt)
{
    try {
        rtb::Camera k; float* q = state17;
        k.origin.x = q[0]; k.origin.y = q[1]; k.origin.z = q[2]; k.look_at.x = q[3]; k.look_at.y = q[4]; k.look_at.z = q[5];
        rtb::CameraController& t = k.controller;
        t.amount_left = q[6]; t.amount_right = q[7]; t.amount_forward = q[8]; t.amount_backward = q[9]; t.amount_up = q[10]; t.amount_down = q[11];
        t.rotate_horizontal = q[12]; t.rotate_vertical = q[13]; t.scroll = q[14]; t.speed = q[15]; t.sensitivity = q[16];
        k.update_camera(dt);
        q[0] = k.origin.x; q[1] = k.origin.y; q[2] = k.origin.z; q[3] = k.look_at.x; q[4] = k.look_at.y; q[5] = k.look_at.z;
        q[12] = t.rotate_horizontal; q[13] = t.rotate_vertical; q[14] = t.scroll;
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
int rth_controller_key(float* state17, int key, int pressed)
{
    rtb::CameraController t; float* q = state17;
    t.amount_left = q[6]; t.amount_right = q[7]; t.amount_forward = q[8]; t.amount_backward = q[9]; t.amount_up = q[10]; t.amount_down = q[11];
    bool r = t.process_keyboard((rtb::Key)key, pressed != 0);
    q[6] = t.amount_left; q[7] = t.amount_right; q[8] = t.amount_forward; q[9] = t.amount_backward; q[10] = t.amount_up; q[11] = t.amount_down;
    return r ? 1 : 0;
}
int rth_context_render(void* c) { RTH_TRY(((rtb::Context*)c)->render()); }
int rth_context_run_frames(void* c, uint32_t n) { RTH_TRY(((rtb::Context*)c)->run_frames(n)); }
int rth_context_resize(void* c, uint32_t w, uint32_t h) { RTH_TRY(((rtb::Context*)c)->resize(w, h)); }
int rth_context_clear_accululation(void* c) { RTH_TRY(((rtb::Context*)c)->clear_accululation()); }
int rth_context_set_scene(void* c, int id) { RTH_TRY(((rtb::Context*)c)->set_scene(id)); }
int rth_context_sync(void* c) { RTH_TRY(((rtb::Context*)c)->sync()); }
int rth_context_set_shard(void* c, uint32_t tr, uint32_t tw, uint32_t fr, uint32_t fw) { RTH_TRY(((rtb::Context*)c)->set_shard(tr, tw, fr, fw)); }
int rth_context_combine(void* const* ctxs, uint32_t n)
{
    RTH_TRY({ std::vector<rtb::Context*> v; for (uint32_t k = 0; k < n; k++) v.push_back((rtb::Context*)ctxs[k]); rtb::Context::combine(v); });
}
int rth_context_set_scene_arrays(void* c, const float* cam9_fov_aspect_etc /* origin3, look_at3, up3, fov, aspect, near, far, aperture, focus */,
                                 const rt_sphere* sp, uint32_t ns, const rt_vertex* vs, uint32_t nv, const uint32_t* ix,
                                 uint32_t ni, const rt_mesh* ms, uint32_t nm)
{
    RTH_TRY({
        const float* q = cam9_fov_aspect_etc;
        rtb::Scene s;
        s.camera = rtb::Camera({q[0], q[1], q[2]}, {q[3], q[4], q[5]}, {q[6], q[7], q[8]}, q[9], q[10], q[11], q[12], q[13], q[14]);
        s.spheres.assign(sp, sp + ns); s.vertices.assign(vs, vs + nv); s.indices.assign(ix, ix + ni); s.meshes.assign(ms, ms + nm);
        ((rtb::Context*)c)->set_scene(std::move(s));
    });
}
void rth_context_get_params(void* c, rt_params* out) { *out = ((rtb::Context*)c)->params; }
void rth_context_set_params(void* c, const rt_params* p) { ((rtb::Context*)c)->params = *p; }
int rth_context_read_texture(void* c, float* out, size_t n)
{
    RTH_TRY({
        rtb::Context* x = (rtb::Context*)c;
        std::vector<float> v = x->read_texture();
        if (v.size() != n) throw std::runtime_error("rth_context_read_texture: size must be W*H*4");
        memcpy(out, v.data(), n * sizeof(float));
    });
}
void* rth_context_handle(void* c) { return ((rtb::Context*)c)->handle(); }

} // extern "C"
