/* wgsl_ref_hooks.h — TEST INFRASTRUCTURE (oracle). Call hooks of the generated oracle/_ref code.
 *
 * make_ref.py wraps every function of the translated shader as
 *     R f(args) { WGSL_HOOK_ENTER(FN_f, args); R r = f_body(args); WGSL_HOOK_RETURN(FN_f, r); return r; }
 * The hooks only observe: they count calls per function (=> path segments, primitive tests, RNG draws,
 * camera paths per dispatch) and capture the primary-hit AOV the hand oracle and the CUDA path export,
 * neither of which the shader itself writes anywhere.  Included by the generated file right after
 * `enum fn_id` / `fn_names[]`, inside no namespace.
 *
 * The AOV part is the one place that names things of the reference shader (it has to say WHICH calls
 * it watches): the closest-hit function `calculate_ray_collions` (wgsl:149), the two primitive tests it
 * calls once per primitive in scan order, `ray_sphere` (wgsl:106) and `ray_triangle` (wgsl:125), and the
 * `hit` / `dst` members of their result (wgsl:93-99).  The primitive's id is the ordinal of the test
 * within the closest-hit call — the global scan order, spheres then triangles mesh by mesh — and the
 * winner is tracked with the shader's own rule (strict `<`, wgsl:154,177); when the closest-hit call
 * returns, its `dst` must equal the tracked winner's or the mismatch counter is raised (tests assert 0).
 */
#ifndef WGSL_REF_HOOKS_H
#define WGSL_REF_HOOKS_H
#include <stdint.h>
#include <string.h>

namespace wgsl { namespace ref {

struct hook_tls {
    uint64_t calls[FN_COUNT];
    /* primary-hit capture for the current invocation of the entry point */
    bool armed, capturing, captured;
    uint32_t ordinal, best_ordinal;
    float best_dst;
    bool prim_hit; float prim_dst; uint32_t prim_id;
    uint64_t mismatches;
};
static thread_local hook_tls g_hook;

static int fn_index(const char* name)
{
    for (int k = 0; k < FN_COUNT; k++) if (strcmp(fn_names[k], name) == 0) return k;
    return -1;
}
static const int FN_CLOSEST = fn_index("calculate_ray_collions");
static const int FN_SPHERE_TEST = fn_index("ray_sphere");
static const int FN_TRIANGLE_TEST = fn_index("ray_triangle");

/* members `hit` and `dst`, when the type has them */
template <class R> static auto hit_of(const R& r, int) -> decltype((bool)r.hit) { return (bool)r.hit; }
template <class R> static bool hit_of(const R&, long) { return false; }
template <class R> static auto dst_of(const R& r, int) -> decltype((float)r.dst) { return (float)r.dst; }
template <class R> static float dst_of(const R&, long) { return 0.0f; }

static inline void hook_enter(int fn)
{
    hook_tls& h = g_hook;
    h.calls[fn]++;
    if (fn == FN_CLOSEST && h.armed && !h.captured && !h.capturing) {
        h.capturing = true; h.ordinal = 0; h.best_ordinal = 0xFFFFFFFFu; h.best_dst = 3.40282346638528859812e+38f;
    }
}
template <class R> static inline void hook_return(int fn, const R& r)
{
    hook_tls& h = g_hook;
    if (!h.capturing) return;
    if (fn == FN_SPHERE_TEST || fn == FN_TRIANGLE_TEST) {
        if (hit_of(r, 0) && dst_of(r, 0) < h.best_dst) { h.best_dst = dst_of(r, 0); h.best_ordinal = h.ordinal; }
        h.ordinal++;
    } else if (fn == FN_CLOSEST) {
        h.capturing = false; h.captured = true;
        h.prim_hit = hit_of(r, 0); h.prim_dst = dst_of(r, 0);
        h.prim_id = h.prim_hit ? h.best_ordinal : 0xFFFFFFFFu;
        if (h.prim_hit && !(h.prim_dst == h.best_dst)) h.mismatches++;
        if (!h.prim_hit && h.best_ordinal != 0xFFFFFFFFu) h.mismatches++;
    }
}

} }

#define WGSL_HOOK_ENTER(id, ...) ::wgsl::ref::hook_enter(id)
#define WGSL_HOOK_RETURN(id, r) ::wgsl::ref::hook_return(id, r)
#define WGSL_HOOK_LEAVE(id) ((void)0)
#endif
