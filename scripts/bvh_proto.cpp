// bvh_proto.cpp — CPU study tool (not product, not oracle): how many node visits / triangle
// tests per ray segment do different Morton-order builders cost on a config's geometry?
// Used to decide the GPU builder: LBVH (Karras radix tree, what rt_bvh.cu
// built first) vs PLOC (nearest-neighbour merging inside a Morton window) vs a full-sweep
// SAH top-down build as the quality ceiling.
//
//   g++ -O2 -o /tmp/bvh_proto scripts/bvh_proto.cpp && python scripts/bvh_study.py
//
// Input: binary file of n triangles, 9 floats each (A, B, C), then camera: origin, llc,
// horizontal, vertical (12 floats).  Rays: primary rays on a pixel grid + uniform-hemisphere
// bounces (up to 8), which is the ray mix of the benchmark.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

struct V3 { float x, y, z; };
static V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static V3 vmin(V3 a, V3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; }
static V3 vmax(V3 a, V3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }
struct Box { V3 lo, hi; };
static Box merge(Box a, Box b) { return {vmin(a.lo, b.lo), vmax(a.hi, b.hi)}; }
static float area(Box b) { V3 d = b.hi - b.lo; return 2.f * (d.x * d.y + d.y * d.z + d.z * d.x); }

struct Tri { V3 a, b, c; };
// binary tree over primitives: child >= 0 inner, child < 0 -> ~prim
struct Node { int l, r; Box box; int count; };
struct Tree { std::vector<Node> nodes; int root; };

static uint64_t spread21(uint32_t v)
{
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull; x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full; x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

static std::vector<Tri> tris;
static std::vector<Box> pbox;
static std::vector<int> order;       // Morton order of primitives
static std::vector<uint64_t> keys;   // sorted keys

static void morton_sort()
{
    int n = (int)tris.size();
    Box sb = pbox[0];
    for (auto& b : pbox) sb = merge(sb, b);
    V3 ext = sb.hi - sb.lo;
    float e = std::max(ext.x, std::max(ext.y, ext.z));
    std::vector<std::pair<uint64_t, int>> kv(n);
    for (int i = 0; i < n; i++) {
        V3 c = (pbox[i].lo + pbox[i].hi) * 0.5f;
        auto q = [&](float v, float lo) { float t = std::min(std::max((v - lo) / e, 0.f), 1.f); return (uint32_t)std::min(t * 2097152.0f, 2097151.0f); };
        kv[i] = {(spread21(q(c.x, sb.lo.x)) << 2) | (spread21(q(c.y, sb.lo.y)) << 1) | spread21(q(c.z, sb.lo.z)), i};
    }
    std::stable_sort(kv.begin(), kv.end());
    order.resize(n); keys.resize(n);
    for (int i = 0; i < n; i++) { keys[i] = kv[i].first; order[i] = kv[i].second; }
}

// ---- LBVH: split at the highest differing key bit (equivalent to the Karras radix tree) ----
static int lbvh_rec(Tree& t, int first, int last)
{
    if (first == last) return ~order[first];
    auto delta = [&](int i, int j) { uint64_t a = keys[i], b = keys[j]; return a == b ? 64 + __builtin_clz((uint32_t)i ^ (uint32_t)j) : __builtin_clzll(a ^ b); };
    int d = delta(first, last);
    int lo = first, hi = last;          // largest s in [first,last) with delta(first,s) > d
    while (hi - lo > 1) { int mid = (lo + hi) / 2; if (delta(first, mid) > d) lo = mid; else hi = mid; }
    int s = lo;
    int id = (int)t.nodes.size(); t.nodes.push_back({});
    int l = lbvh_rec(t, first, s), r = lbvh_rec(t, s + 1, last);
    t.nodes[id].l = l; t.nodes[id].r = r;
    return id;
}

// ---- PLOC (Meister & Bittner 2018): merge mutual nearest neighbours inside a window of
// `radius` clusters on each side in Morton order, until one cluster is left ----
static Tree build_ploc(int radius)
{
    Tree t; int n = (int)tris.size();
    std::vector<int> c(n), c2; std::vector<Box> b(n), b2;
    for (int i = 0; i < n; i++) { c[i] = ~order[i]; b[i] = pbox[order[i]]; }
    int iters = 0;
    while (c.size() > 1) {
        int m = (int)c.size();
        std::vector<int> nn(m);
        for (int i = 0; i < m; i++) {
            float best = INFINITY; int bj = -1;
            for (int j = std::max(0, i - radius); j <= std::min(m - 1, i + radius); j++) {
                if (j == i) continue;
                float a = area(merge(b[i], b[j]));
                if (a < best) { best = a; bj = j; }      // ties: smallest j
            }
            nn[i] = bj;
        }
        c2.clear(); b2.clear();
        for (int i = 0; i < m; i++) {
            int j = nn[i];
            if (nn[j] == i) {
                if (i < j) {
                    int id = (int)t.nodes.size();
                    Node nd; nd.l = c[i]; nd.r = c[j]; nd.box = merge(b[i], b[j]); nd.count = 0;
                    t.nodes.push_back(nd);
                    c2.push_back(id); b2.push_back(nd.box);
                }
            } else { c2.push_back(c[i]); b2.push_back(b[i]); }
        }
        c.swap(c2); b.swap(b2); iters++;
    }
    t.root = c[0];
    fprintf(stderr, "ploc r=%d: %d iterations\n", radius, iters);
    return t;
}

// ---- full-sweep SAH, top-down (quality ceiling) ----
static int sah_rec(Tree& t, std::vector<int>& ids, int first, int last)
{
    if (first == last) return ~ids[first];
    int n = last - first + 1;
    float bestc = INFINITY; int bax = 0, bsplit = first;
    std::vector<float> ra(n);
    for (int ax = 0; ax < 3; ax++) {
        std::sort(ids.begin() + first, ids.begin() + last + 1, [&](int a, int b) {
            float ca = ax == 0 ? pbox[a].lo.x + pbox[a].hi.x : ax == 1 ? pbox[a].lo.y + pbox[a].hi.y : pbox[a].lo.z + pbox[a].hi.z;
            float cb = ax == 0 ? pbox[b].lo.x + pbox[b].hi.x : ax == 1 ? pbox[b].lo.y + pbox[b].hi.y : pbox[b].lo.z + pbox[b].hi.z;
            return ca < cb || (ca == cb && a < b); });
        Box rb = pbox[ids[last]];
        for (int i = last; i > first; i--) { rb = merge(rb, pbox[ids[i]]); ra[i - first] = area(rb); }
        Box lb = pbox[ids[first]];
        for (int i = first; i < last; i++) {
            lb = merge(lb, pbox[ids[i]]);
            float c = area(lb) * (i - first + 1) + ra[i + 1 - first] * (last - i);
            if (c < bestc) { bestc = c; bax = ax; bsplit = i; }
        }
    }
    if (bax != 2)
        std::sort(ids.begin() + first, ids.begin() + last + 1, [&](int a, int b) {
            float ca = bax == 0 ? pbox[a].lo.x + pbox[a].hi.x : pbox[a].lo.y + pbox[a].hi.y;
            float cb = bax == 0 ? pbox[b].lo.x + pbox[b].hi.x : pbox[b].lo.y + pbox[b].hi.y;
            return ca < cb || (ca == cb && a < b); });
    int id = (int)t.nodes.size(); t.nodes.push_back({});
    int l = sah_rec(t, ids, first, bsplit), r = sah_rec(t, ids, bsplit + 1, last);
    t.nodes[id].l = l; t.nodes[id].r = r;
    return id;
}

static void fit(Tree& t, int n, Box& box, int& count)
{
    if (n < 0) { box = pbox[~n]; count = 1; return; }
    Box a, b; int ca, cb;
    fit(t, t.nodes[n].l, a, ca); fit(t, t.nodes[n].r, b, cb);
    t.nodes[n].box = box = merge(a, b); t.nodes[n].count = count = ca + cb;
}

static double sah_cost(const Tree& t, int leaf_size)
{
    double c = 0; float ra = area(t.nodes[t.root].box);
    std::function<void(int)> rec = [&](int n) {
        if (n < 0) { c += area(pbox[~n]) / ra; return; }
        const Node& nd = t.nodes[n];
        if (nd.count <= leaf_size) { c += nd.count * (double)area(nd.box) / ra; return; }
        c += 1.2 * area(nd.box) / ra;
        rec(nd.l); rec(nd.r);
    };
    rec(t.root);
    return c;
}

// ---- traversal statistics: same policy as the CUDA kernel (near child first, prune on tn <= best) ----
struct Stats { uint64_t rays = 0, nodes = 0, tris = 0, hits = 0; };
static int g_leaf = 4;
static bool g_popcull = false;   // stack entries carry their entry distance; popped nodes beyond the closest hit are skipped unfetched

static bool slab(const Box& b, V3 o, V3 inv, float best, float& tn)
{
    float t0 = (b.lo.x - o.x) * inv.x, t1 = (b.hi.x - o.x) * inv.x; if (t0 > t1) std::swap(t0, t1);
    float u0 = (b.lo.y - o.y) * inv.y, u1 = (b.hi.y - o.y) * inv.y; if (u0 > u1) std::swap(u0, u1);
    float v0 = (b.lo.z - o.z) * inv.z, v1 = (b.hi.z - o.z) * inv.z; if (v0 > v1) std::swap(v0, v1);
    tn = std::max(std::max(t0, u0), std::max(v0, 0.f));
    float tf = std::min(std::min(t1, u1), std::min(v1, best));
    return tn <= tf;
}

static void leaf_prims(const Tree& t, int n, std::vector<int>& out)
{
    if (n < 0) { out.push_back(~n); return; }
    leaf_prims(t, t.nodes[n].l, out); leaf_prims(t, t.nodes[n].r, out);
}

static int trace(const Tree& t, V3 o, V3 d, float& best, Stats& st)
{
    V3 inv = {1.f / d.x, 1.f / d.y, 1.f / d.z};
    int stack[128], sp = 0, hit = -1; float stn[128];
    int node = t.root;
    auto is_leaf = [&](int n) { return n < 0 || t.nodes[n].count <= g_leaf; };
    auto test_leaf = [&](int n) {
        static thread_local std::vector<int> ps; ps.clear(); leaf_prims(t, n, ps);
        for (int p : ps) {
            st.tris++;
            const Tri& tr = tris[p];
            V3 e1 = tr.b - tr.a, e2 = tr.c - tr.a, nrm = cross(e1, e2);
            float det = -dot(d, nrm);
            if (!(det >= 1e-6f)) continue;
            V3 ao = o - tr.a, dao = cross(ao, d);
            float invd = 1.f / det, dst = dot(ao, nrm) * invd, u = dot(e2, dao) * invd, v = -dot(e1, dao) * invd, w = 1.f - u - v;
            if (dst >= 0 && u >= 0 && v >= 0 && w >= 0 && dst < best) { best = dst; hit = p; }
        }
    };
    if (is_leaf(node)) { test_leaf(node); return hit; }
    for (;;) {
        const Node& nd = t.nodes[node];
        st.nodes++;
        auto cbox = [&](int c) { return c < 0 ? pbox[~c] : t.nodes[c].box; };
        float tl, tr_; bool hl = slab(cbox(nd.l), o, inv, best, tl), hr = slab(cbox(nd.r), o, inv, best, tr_);
        int next = INT32_MAX;
        if (hl && hr) { bool lf = tl <= tr_; stn[sp] = lf ? tr_ : tl; stack[sp++] = lf ? nd.r : nd.l; next = lf ? nd.l : nd.r; }
        else if (hl) next = nd.l; else if (hr) next = nd.r;
        for (;;) {
            if (next == INT32_MAX) { if (sp == 0) return hit; next = stack[--sp]; if (g_popcull && stn[sp] > best) { next = INT32_MAX; continue; } }
            if (is_leaf(next)) { test_leaf(next); next = INT32_MAX; continue; }
            break;
        }
        node = next;
    }
}

static uint32_t rng_state = 12345;
static float rnd() { rng_state = rng_state * 747796405u + 2891336453u; uint32_t r = ((rng_state >> ((rng_state >> 28) + 4)) ^ rng_state) * 277803737u; r = (r >> 22) ^ r; return (r >> 8) * (1.f / 16777216.f); }

static Stats run(const Tree& t, const float* cam, int W, int H, int step, int bounces)
{
    Stats st; rng_state = 12345;
    V3 org = {cam[0], cam[1], cam[2]}, llc = {cam[3], cam[4], cam[5]}, hor = {cam[6], cam[7], cam[8]}, ver = {cam[9], cam[10], cam[11]};
    for (int y = 0; y < H; y += step) for (int x = 0; x < W; x += step) {
        V3 o = org, d = llc + hor * ((x + rnd()) / W) + ver * ((y + rnd()) / H) - org;
        for (int s = 0; s <= bounces; s++) {
            float best = 3.4e38f; st.rays++;
            int h = trace(t, o, d, best, st);
            if (h < 0) break;
            st.hits++;
            const Tri& tr = tris[h];
            V3 n = cross(tr.b - tr.a, tr.c - tr.a); n = n * (1.f / std::sqrt(dot(n, n)));
            o = o + d * best;
            V3 r;
            do { r = {rnd() * 2 - 1, rnd() * 2 - 1, rnd() * 2 - 1}; } while (dot(r, r) > 1.f || dot(r, r) < 1e-4f);
            if (dot(r, n) < 0) r = r * -1.f;
            d = r * (1.f / std::sqrt(dot(r, r)));
        }
    }
    return st;
}

int main(int argc, char** argv)
{
    if (argc < 2) { fprintf(stderr, "usage: bvh_proto scene.bin [W H step]\n"); return 1; }
    FILE* f = fopen(argv[1], "rb"); if (!f) return 1;
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    int n = (int)((sz - 48) / 36);
    tris.resize(n); float cam[12];
    if (fread(tris.data(), 36, n, f) != (size_t)n || fread(cam, 4, 12, f) != 12) return 1;
    fclose(f);
    int W = argc > 2 ? atoi(argv[2]) : 1920, H = argc > 3 ? atoi(argv[3]) : 1080, step = argc > 4 ? atoi(argv[4]) : 8;
    pbox.resize(n);
    for (int i = 0; i < n; i++) pbox[i] = {vmin(tris[i].a, vmin(tris[i].b, tris[i].c)), vmax(tris[i].a, vmax(tris[i].b, tris[i].c))};
    morton_sort();
    auto report = [&](const char* name, Tree& t) {
        Box b; int c; fit(t, t.root, b, c);
        for (int leaf : {1, 2, 4}) {
            g_leaf = leaf;
            Stats s = run(t, cam, W, H, step, 8);
            printf("%-12s leaf<=%d  SAH %.2f  rays %llu  nodes/ray %.2f  tris/ray %.2f  hit %.3f\n", name, leaf, sah_cost(t, leaf),
                   (unsigned long long)s.rays, (double)s.nodes / s.rays, (double)s.tris / s.rays, (double)s.hits / s.rays);
        }
    };
    { Tree t; t.root = lbvh_rec(t, 0, n - 1); report("lbvh", t); }
    { Tree t = build_ploc(3); report("ploc r=3", t); g_popcull = true; report("ploc3+popcull", t); g_popcull = false; }
    for (int r : {4, 10, 25, 50}) { Tree t = build_ploc(r); char nm[32]; snprintf(nm, sizeof nm, "ploc r=%d", r); report(nm, t); }
    if (n <= 200000) { Tree t; std::vector<int> ids(n); for (int i = 0; i < n; i++) ids[i] = i; t.root = sah_rec(t, ids, 0, n - 1); report("sah sweep", t); }
    return 0;
}
