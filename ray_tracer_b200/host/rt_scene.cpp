// rt_scene.cpp — Camera::to_uniform, record constructors and the built-in scenes.
// Compiled with -ffp-contract=off and without -mfma: every float operation below is one
// IEEE binary32 operation in the order glam 0.23 / the reference performs it.
#include <cmath>
#include <stdexcept>
#include "rt_host.hpp"

namespace rtb {

Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
Vec3 operator*(float s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
float dot(Vec3 a, Vec3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }
Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
float length(Vec3 a) { return std::sqrt(dot(a, a)); }
Vec3 normalize(Vec3 a) { float r = 1.0f / length(a); return {a.x * r, a.y * r, a.z * r}; }

float radians(float deg) { return deg * (3.14159274101257324f / 180.0f); }

static void put3(float* d, Vec3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
static void put4(float* d, Vec4 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w; }

rt_sphere Sphere(Vec3 position, float radius, Vec4 color, Vec4 emission_color, float emission_strength, float specular)
{
    rt_sphere s{};
    put3(s.position, position); s.radius = radius; put4(s.color, color); put4(s.emission_color, emission_color);
    s.emission_strength = emission_strength;
    s.smoothness = specular < 1.0f ? specular : 1.0f;
    return s;
}

rt_mesh Mesh(Vec3 pos, uint32_t first, uint32_t triangles, uint32_t offset, Vec4 color, Vec4 emission_color,
             float emission_strength, float specular)
{
    rt_mesh m{};
    m.first = first; m.triangles = triangles; m.offset = offset;
    put3(m.pos, pos); put4(m.color, color); put4(m.emission_color, emission_color);
    m.emission_strength = emission_strength;
    m.specular = specular < 1.0f ? specular : 1.0f;
    return m;
}

rt_vertex Vertex(Vec3 pos, Vec3 normal)
{
    rt_vertex v{};
    put3(v.pos, pos); put3(v.normal, normal);
    return v;
}

Camera::Camera(Vec3 o, Vec3 la, Vec3 up, float fov_, float aspect_, float n, float f, float ap, float fd)
    : origin(o), look_at(la), view_up(up), fov(fov_), aspect(aspect_), near_(n), far_(f), aperture(ap), focus_dist(fd) {}

rt_camera_uniform Camera::to_uniform() const
{
    float theta = radians(fov);
    float height = 2.0f * std::tan(theta / 2.0f);
    float width = aspect * height;
    Vec3 w = normalize(origin - look_at);
    Vec3 u = normalize(cross(view_up, w));
    Vec3 v = cross(w, u);
    Vec3 horizontal = (focus_dist * width) * u;
    Vec3 vertical = (focus_dist * height) * v;
    Vec3 llc = origin - horizontal / 2.0f - vertical / 2.0f - focus_dist * w;
    rt_camera_uniform cu{};
    put3(cu.origin, origin); put3(cu.lower_left_corner, llc); put3(cu.horizontal, horizontal); put3(cu.vertical, vertical);
    cu.near_ = near_; cu.far_ = far_;
    put3(cu.w, w); put3(cu.u, u); put3(cu.v, v);
    cu.lens_radius = aperture / 2.0f;
    return cu;
}

// ---- fly-cam (camera.rs:122-247) ---------------------------------------------------------
bool CameraController::process_keyboard(Key key, bool pressed)
{
    float amount = pressed ? 5.0f : 0.0f;
    switch (key) {
    case Key::W: case Key::Up: amount_forward = amount; return true;
    case Key::S: case Key::Down: amount_backward = amount; return true;
    case Key::A: case Key::Left: amount_left = amount; return true;
    case Key::D: case Key::Right: amount_right = amount; return true;
    case Key::Space: amount_up = amount; return true;
    case Key::LShift: amount_down = amount; return true;
    default: return false;
    }
}
void CameraController::process_mouse(double dx, double dy) { rotate_horizontal = (float)dx * 3.0f; rotate_vertical = (float)dy * 3.0f; }
bool CameraController::process_scroll_lines(float lines) { scroll = -(lines * 10000.0f); return true; }
bool CameraController::process_scroll_pixels(double pixels) { scroll = -(float)pixels; return true; }

static Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }

void Camera::update_camera(float dt)
{
    // SAFE_FRAC_PI_2_DEG: (FRAC_PI_2 - 0.0001) * (180/PI), compared against radians (camera.rs:6-7,159-163)
    const float safe_deg = (1.57079637050628662f - 0.0001f) * (180.0f / 3.14159274101257324f);
    CameraController& c = controller;
    Vec3 direction = normalize(look_at - origin);
    float pitch = std::asin(direction.y);
    float yaw = std::atan2(direction.x, direction.z);
    float yaw_sin = std::sin(yaw), yaw_cos = std::cos(yaw);
    Vec3 forward = normalize(Vec3{yaw_sin, 0.0f, yaw_cos});
    Vec3 right = normalize(Vec3{yaw_cos, 0.0f, -yaw_sin});
    origin = origin + ((forward * (c.amount_forward - c.amount_backward)) * c.speed) * dt;
    origin = origin + ((right * (c.amount_right - c.amount_left)) * c.speed) * dt;
    float pitch_sin = std::sin(pitch), pitch_cos = std::cos(pitch);
    Vec3 scrollward = normalize(Vec3{pitch_cos * yaw_cos, pitch_sin, pitch_cos * yaw_sin});
    origin = origin - (((scrollward * c.scroll) * c.speed) * c.sensitivity) * dt;
    c.scroll = 0.0f;
    origin.y += (c.amount_up - c.amount_down) * c.speed * dt;
    yaw += c.rotate_horizontal * c.sensitivity * dt;
    pitch += -c.rotate_vertical * c.sensitivity * dt;
    c.rotate_horizontal = 0.0f; c.rotate_vertical = 0.0f;
    if (pitch < -safe_deg) pitch = -safe_deg;
    else if (pitch > safe_deg) pitch = safe_deg;
    look_at = origin + Vec3{std::cos(pitch) * std::sin(yaw), std::sin(pitch), std::cos(pitch) * std::cos(yaw)};
}

void Scene::add_dummy_mesh()
{
    vertices.push_back(Vertex({0.0f, 0.0f, 1.0f}, {2.0f, -3.0f, -1.0f}));
    indices.push_back(1u);
    meshes.push_back(Mesh({0, 0, 0}, 0, 0, 0, {0.0f, 0.6f, 0.0f, 1.0f}, {1, 1, 1, 1}, 0.0f, 0.5f));
}

Scene Scene::empty(float aspect)
{
    Scene s;
    s.camera = Camera({-2.764473f, 5.8210998f, 3.839141f}, {-2.0999293f, 5.1703076f, 3.4719195f}, {0, 1, 0}, 45.0f,
                      aspect, 0.1f, 100.0f, 1.0f, 2.0f);
    return s;
}

Scene Scene::balls(float aspect)
{
    Scene s;
    s.camera = Camera({3.089f, 1.53f, -3.0f}, {-2.0f, -1.0f, 2.0f}, {0, 1, 0}, 45.0f, aspect, 0.1f, 100.0f, 0.0f, 0.1f);
    const Vec4 k{0, 0, 0, 1}, one{1, 1, 1, 1};
    s.spheres = {
        Sphere({-3.64f, -0.42f, 0.8028f}, 0.75f, one, k, 0.0f, 0.7f),
        Sphere({-2.54f, -0.72f, 0.5f}, 0.6f, {1, 0, 0, 1}, k, 0.0f, 0.5f),
        Sphere({-1.27f, -0.72f, 1.0f}, 0.5f, {0, 1, 0, 1}, one, 0.0f, 0.2f),
        Sphere({-0.5f, -0.9f, 1.55f}, 0.35f, {0, 0, 1, 1}, one, 0.0f, 0.0f),
        Sphere({-3.46f, -15.88f, 2.76f}, 15.0f, {0.5f, 0.0f, 0.8f, 1.0f}, one, 0.0f, 0.0f),       // floor
        Sphere({-7.44f, -0.72f, 20.0f}, 15.0f, {0.1f, 0.1f, 0.1f, 0.0f}, one, 2.0f, 0.0f),        // light
    };
    s.add_dummy_mesh();
    return s;
}

Scene Scene::metal(float aspect)
{
    Scene s;
    s.camera = Camera({0, 0, 3}, {0, 0, -1}, {0, 1, 0}, 45.0f, aspect, 0.1f, 100.0f, 0.0f, 0.1f);
    const Vec4 k{0, 0, 0, 1};
    s.spheres = {
        Sphere({0.0f, -100.5f, -1.0f}, 100.0f, {0.8f, 0.8f, 0.0f, 1.0f}, k, 0.0f, 0.0f),
        Sphere({0.0f, 0.0f, -1.0f}, 0.5f, {0.7f, 0.3f, 0.3f, 1.0f}, k, 0.0f, 0.0f),
        Sphere({-1.0f, 0.0f, -1.0f}, 0.5f, {0.8f, 0.8f, 0.8f, 1.0f}, k, 0.0f, -1.0f),
        Sphere({1.0f, 0.0f, -1.0f}, 0.5f, {0.8f, 0.6f, 0.2f, 1.0f}, k, 0.0f, 0.15f),
    };
    s.add_dummy_mesh();
    return s;
}

Scene Scene::room(float aspect)
{
    Scene s;
    s.camera = Camera({-7, 0, 0}, {1, 0, 0}, {0, 1, 0}, 45.0f, aspect, 0.1f, 100.0f, 0.0f, 0.1f);
    const Vec4 k{0, 0, 0, 1}, one{1, 1, 1, 1};
    s.spheres = {
        Sphere({4.0f, 0.0f, 1.7f}, 1.2f, one, k, 0.0f, 1.0f),
        Sphere({4.0f, 0.0f, -1.7f}, 1.2f, one, k, 0.0f, 0.5f),
    };
    const Vec3 n{3.0f, -4.0f, 2.0f};
    s.vertices = {
        Vertex({3, -3, -3}, {2, -3, -3}), Vertex({3, -3, 3}, {4, -3, 0}), Vertex({-3, -3, 3}, n), Vertex({-3, -3, -3}, n),
        Vertex({3, 3, -3}, n), Vertex({3, 3, 3}, n), Vertex({-3, 3, 3}, n), Vertex({-3, 3, -3}, n),
        Vertex({1, 1, -1}, n), Vertex({1, 1, 1}, n), Vertex({-1, 1, 1}, n), Vertex({-1, 1, -1}, n),
    };
    s.indices = {3, 2, 1, 3, 1, 0, 7, 0, 4, 7, 3, 0, 7, 6, 2, 7, 2, 3, 2, 6, 5, 2, 5, 1,
                 1, 5, 4, 1, 4, 0, 5, 6, 7, 5, 7, 4, 9, 10, 11, 9, 11, 8};
    s.meshes = {
        Mesh({3, 0, 0}, 0, 2, 0, {1, 0, 0, 1}, one, 0.0f, 0.5f),
        Mesh({3, 0, 0}, 6, 2, 0, {0, 1, 0, 1}, one, 0.0f, 0.5f),
        Mesh({3, 0, 0}, 12, 2, 0, {0, 0, 1, 1}, one, 0.0f, 0.5f),
        Mesh({3, 0, 0}, 18, 2, 0, {0.5f, 0.5f, 0.0f, 1.0f}, one, 0.0f, 0.5f),
        Mesh({3, 0, 0}, 24, 2, 0, {0.0f, 0.5f, 0.5f, 1.0f}, one, 0.0f, 0.5f),
        Mesh({3, 0, 0}, 30, 2, 0, one, one, 0.0f, 0.5f),
        Mesh({3.0f, 1.9f, 0.0f}, 36, 2, 0, one, one, 10.5f, 0.0f),
    };
    return s;
}

namespace {
// Stand-in for rand::thread_rng (scene.rs:138, unseeded upstream): the shader's own PCG
// hash stream (wgsl:205-210); gen() = top 24 bits * 2^-24 in [0,1).
struct FixedRng {
    uint32_t s;
    float gen()
    {
        s = s * 747796405u + 2891336453u;
        uint32_t r = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
        r = (r >> 22u) ^ r;
        return (float)(r >> 8) * (1.0f / 16777216.0f);
    }
    float gen_range(float lo, float hi) { return lo + (hi - lo) * gen(); }
};
} // namespace

Scene Scene::random_balls(float aspect, uint32_t seed)
{
    Scene s;
    s.camera = Camera({10.5f, 2.0f, 3.0f}, {0, 0, 0}, {0, 1, 0}, 45.0f, aspect, 0.1f, 100.0f, 0.1f, 10.0f);
    FixedRng rng{seed};
    const Vec4 zero{0, 0, 0, 0}, one{1, 1, 1, 1};
    s.spheres.push_back(Sphere({0.0f, -1000.0f, 0.0f}, 1000.0f, {0.5f, 0.5f, 0.5f, 1.0f}, zero, 0.0f, 0.0f));
    for (int a = -11; a < 11; a++) {
        for (int b = -11; b < 11; b++) {
            float mat = rng.gen();
            float cx = (float)a + 0.9f * rng.gen();
            float cz = (float)b + 0.9f * rng.gen();
            Vec3 center{cx, 0.2f, cz};
            if (length(center - Vec3{4.0f, 0.2f, 0.0f}) > 0.9f) {
                if (mat < 0.8f) {
                    float r = rng.gen(), g = rng.gen(), bl = rng.gen();
                    s.spheres.push_back(Sphere(center, 0.2f, {r, g, bl, 1.0f}, zero, 0.0f, 0.0f));
                } else if (mat < 0.95f) {
                    float r = rng.gen_range(0.5f, 1.0f), g = rng.gen_range(0.5f, 1.0f), bl = rng.gen_range(0.5f, 1.0f);
                    float fuzz = rng.gen_range(0.0f, 0.5f);
                    s.spheres.push_back(Sphere(center, 0.2f, {r, g, bl, 1.0f}, zero, 0.0f, fuzz));
                } else {
                    s.spheres.push_back(Sphere(center, 0.2f, one, zero, 0.0f, -1.0f));
                }
            }
        }
    }
    s.spheres.push_back(Sphere({0, 1, 0}, 1.0f, one, zero, 0.0f, -1.0f));
    s.spheres.push_back(Sphere({-4, 1, 0}, 1.0f, {0.4f, 0.2f, 0.1f, 1.0f}, zero, 0.0f, 0.0f));
    s.spheres.push_back(Sphere({4, 1, 0}, 1.0f, {0.7f, 0.6f, 0.5f, 1.0f}, zero, 0.0f, 0.9f));
    s.add_dummy_mesh();
    return s;
}

Scene Scene::by_id(int id, float aspect)
{
    switch (id) {
    case 0: return balls(aspect);
    case 1: return random_balls(aspect, 12345u);
    case 2: return room(aspect);
    case 3: return metal(aspect);
    default: throw std::runtime_error("Scene::by_id: scene id must be 0..3");
    }
}

} // namespace rtb
