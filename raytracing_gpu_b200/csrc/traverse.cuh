// traverse.cuh — closest-hit query over the LBVH + media: what color_f asks of the world (render.h:63).
// Replaces hittable_list::hit / bvh_node::hit (hittable_list.h:23, bvh.h:348): no device new/delete, no powf,
// ordered descent with a shrinking t_max, one 64-byte node fetch per two child boxes.
#pragma once

#include "intersect.cuh"
#include "rng.cuh"
#include <climits>

namespace rtb {

#define RT_STACK RT_STACK_DEPTH   // Karras depth bound: 63 key bits + index bits for duplicate keys; PLOC depth is checked at build

struct Closest {
    float t;     // +inf on miss
    int prim;    // >= 0: surface primitive (device order); -1: miss; <= -2: medium index = -2 - prim
};

struct TraversalCounters { uint32_t nodes, prim_tests; };

// ties: the reference's hittable_list lets the LATER object win (t == closest_so_far is accepted,
// sphere.h:49, aarect.h:66) — with an unordered traversal that becomes "higher leaf id wins".
__device__ __forceinline__ void consider(const SceneView& sv, int prim, float t, Closest& best) {
    if (t < best.t || best.prim < 0 || sv.prim_leaf_id[prim] > sv.prim_leaf_id[best.prim]) { best.t = t; best.prim = prim; }
}

// shared-memory accesses the compiler neither hoists nor keeps in registers (see closest_surface_stream)
__device__ __forceinline__ float4 lds4(const float4* p) {   // not hoisted, not kept in registers across the node phase
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
    return v;
}
__device__ __forceinline__ void sts4(float4* p, float4 v) {
    asm volatile("st.shared.v4.f32 [%4], {%0,%1,%2,%3};" :: "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"((uint32_t)__cvta_generic_to_shared(p)) : "memory");
}

// Two loop shapes (profiles/r01_v4_experiments.md):
//   PARKED = false  fused node-or-leaf step per iteration.  ncu: 35 % of its issue slots go to the leaf code at
//                   3.2 active lanes of 32, because almost every iteration SOME lane is at a leaf.
//   PARKED = true   a lane that reaches a leaf parks it (one register) and keeps walking; the warp enters the leaf
//                   code only when some lane reaches a second leaf or nobody has nodes left, then every parked leaf
//                   is tested together.  Parking only delays the shrinking of t_max and consider() is
//                   order-independent, so the closest hit is the same; measured 2.7 % faster on the final scene.
//                   WARP-COLLECTIVE (votes): every lane of the warp must call, idle lanes with active = false.
#define RT_NODE_DONE INT_MIN

template <bool COUNT, bool PARKED = false>
__device__ __forceinline__ void closest_surface(const SceneView& sv, const Ray& ray, float t_min, Closest& best, TraversalCounters* cnt, bool active) {
    // slab set-up: t = plane * inv - o * inv as one FMA; zero direction components are nudged so inv stays finite
    const float tiny = 1e-20f;
    const float ix = 1.0f / (fabsf(ray.d.x) < tiny ? copysignf(tiny, ray.d.x) : ray.d.x);
    const float iy = 1.0f / (fabsf(ray.d.y) < tiny ? copysignf(tiny, ray.d.y) : ray.d.y);
    const float iz = 1.0f / (fabsf(ray.d.z) < tiny ? copysignf(tiny, ray.d.z) : ray.d.z);
    const float ox = ray.o.x * ix, oy = ray.o.y * iy, oz = ray.o.z * iz;

    // object-space ray cache: consecutive leaves usually share a wrapper chain (e.g. 1000 spheres under one)
    uint32_t cached_chain = 0;
    Ray cached = ray;

    int stack[RT_STACK];
    int sp = 0;
    int node = (active && sv.n_prims) ? 0 : RT_NODE_DONE;
    if constexpr (PARKED) {
        int parked = RT_NODE_DONE;
        auto pop = [&]() { return sp ? stack[--sp] : RT_NODE_DONE; };
        for (;;) {
            if (node >= 0) {
                const BNode* np = sv.nodes + node;
                const float4 n0 = __ldg(&np->n0), n1 = __ldg(&np->n1), n2 = __ldg(&np->n2);
                const int2 ch = __ldg(reinterpret_cast<const int2*>(&np->child0));
                if (COUNT) cnt->nodes++;
                const float c0lox = fmaf(n0.x, ix, -ox), c0hix = fmaf(n0.y, ix, -ox), c0loy = fmaf(n0.z, iy, -oy), c0hiy = fmaf(n0.w, iy, -oy);
                const float c1lox = fmaf(n1.x, ix, -ox), c1hix = fmaf(n1.y, ix, -ox), c1loy = fmaf(n1.z, iy, -oy), c1hiy = fmaf(n1.w, iy, -oy);
                const float c0loz = fmaf(n2.x, iz, -oz), c0hiz = fmaf(n2.y, iz, -oz), c1loz = fmaf(n2.z, iz, -oz), c1hiz = fmaf(n2.w, iz, -oz);
                const float t0n = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
                const float t0f = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), best.t));
                const float t1n = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
                const float t1f = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), best.t));
                const bool h0 = t0n <= t0f, h1 = t1n <= t1f;
                if (h0 && h1) {
                    const bool swap = t1n < t0n;
                    stack[sp++] = swap ? ch.x : ch.y;
                    node = swap ? ch.y : ch.x;
                } else if (h0) node = ch.x;
                else if (h1) node = ch.y;
                else node = pop();
                if (node < 0 && node != RT_NODE_DONE && parked == RT_NODE_DONE) { parked = node; node = pop(); }
            }
            const bool blocked = node < 0 && node != RT_NODE_DONE;   // a second leaf while one is parked
            // two votes, not one __reduce_or_sync of a 2-bit state: VOTE.ANY writes a predicate the branch consumes directly,
            // REDUX goes through a uniform register (+2 % on the final scene and the 1M-triangle mesh, same-box A/B)
            const bool any_blocked = __any_sync(0xffffffffu, blocked), any_walking = __any_sync(0xffffffffu, node >= 0);
            if (any_blocked || !any_walking) {
                if (parked != RT_NODE_DONE) {
                    const int prim = ~parked;
                    parked = RT_NODE_DONE;
                    const DPrim p = sv.prims[prim];
                    const uint32_t chain = prim_chain(p);
                    if (chain != cached_chain) { cached = to_object_space(sv, chain, ray); cached_chain = chain; }
                    if (COUNT) cnt->prim_tests++;
                    float t;
                    if (prim_t_local(p, chain ? cached : ray, t_min, best.t, sv.semantics, t)) consider(sv, prim, t, best);
                }
                if (blocked) { parked = node; node = pop(); }
                if (!__any_sync(0xffffffffu, node != RT_NODE_DONE || parked != RT_NODE_DONE)) break;
            }
        }
        return;
    }
    while (node != RT_NODE_DONE) {
        if (node >= 0) {
            const BNode* np = sv.nodes + node;
            const float4 n0 = __ldg(&np->n0), n1 = __ldg(&np->n1), n2 = __ldg(&np->n2);
            const int2 ch = __ldg(reinterpret_cast<const int2*>(&np->child0));
            if (COUNT) cnt->nodes++;
            const float c0lox = fmaf(n0.x, ix, -ox), c0hix = fmaf(n0.y, ix, -ox), c0loy = fmaf(n0.z, iy, -oy), c0hiy = fmaf(n0.w, iy, -oy);
            const float c1lox = fmaf(n1.x, ix, -ox), c1hix = fmaf(n1.y, ix, -ox), c1loy = fmaf(n1.z, iy, -oy), c1hiy = fmaf(n1.w, iy, -oy);
            const float c0loz = fmaf(n2.x, iz, -oz), c0hiz = fmaf(n2.y, iz, -oz), c1loz = fmaf(n2.z, iz, -oz), c1hiz = fmaf(n2.w, iz, -oz);
            const float t0n = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
            const float t0f = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), best.t));
            const float t1n = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
            const float t1f = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), best.t));
            const bool h0 = t0n <= t0f, h1 = t1n <= t1f;
            if (h0 && h1) {
                const bool swap = t1n < t0n;
                stack[sp++] = swap ? ch.x : ch.y;
                node = swap ? ch.y : ch.x;
                continue;
            }
            if (h0) { node = ch.x; continue; }
            if (h1) { node = ch.y; continue; }
        } else {
            const int prim = ~node;
            const DPrim p = sv.prims[prim];
            const uint32_t chain = prim_chain(p);
            if (chain != cached_chain) { cached = to_object_space(sv, chain, ray); cached_chain = chain; }
            if (COUNT) cnt->prim_tests++;
            float t;
            if (prim_t_local(p, chain ? cached : ray, t_min, best.t, sv.semantics, t)) consider(sv, prim, t, best);
        }
        node = sp ? stack[--sp] : RT_NODE_DONE;
    }
}

// Suspendable walk (k_extend_pass): closest_surface<PARKED = true> whose traversal state is the caller's, so a walk can
// be cut short and taken up again by another warp.  Rays of one 32-ray batch finish at very different times (only ~60 %
// of the node steps a warp pays for carry a live lane on the final scene); instead of letting the last few lanes run the
// warp alone, the walk stops once no more than `suspend_below` lanes are still live (checked where the warp leaves the
// node loop anyway: the leaf phase), the caller stores those lanes' state, and a later pass packs 32 such stragglers
// into one warp.  A lane is live while it has a node to visit or a leaf parked.  The closest hit of a ray depends
// neither on where its walk is cut nor on its neighbours: the state is the complete traversal state.
// WARP-COLLECTIVE; lanes without a ray enter with node = parked = RT_NODE_DONE.  A warp does not suspend while some lane's
// stack is deeper than RT_SUSPEND_STACK (what a continuation record holds).
#define RT_SUSPEND_STACK 24
struct Walk { int node, parked, sp; };

template <bool COUNT>
__device__ __forceinline__ void walk_suspendable(const SceneView& sv, const Ray& ray, float t_min, Closest& best, TraversalCounters* cnt,
                                                 Walk& w, int* stack, uint32_t suspend_below, uint32_t* steps) {
    const float tiny = 1e-20f;
    const float ix = 1.0f / (fabsf(ray.d.x) < tiny ? copysignf(tiny, ray.d.x) : ray.d.x);
    const float iy = 1.0f / (fabsf(ray.d.y) < tiny ? copysignf(tiny, ray.d.y) : ray.d.y);
    const float iz = 1.0f / (fabsf(ray.d.z) < tiny ? copysignf(tiny, ray.d.z) : ray.d.z);
    const float ox = ray.o.x * ix, oy = ray.o.y * iy, oz = ray.o.z * iz;
    uint32_t cached_chain = 0;
    Ray cached = ray;
    int node = w.node, parked = w.parked, sp = w.sp;
    auto pop = [&]() { return sp ? stack[--sp] : RT_NODE_DONE; };
    for (;;) {
        if (node >= 0) {
            const BNode* np = sv.nodes + node;
            const float4 n0 = __ldg(&np->n0), n1 = __ldg(&np->n1), n2 = __ldg(&np->n2);
            const int2 ch = __ldg(reinterpret_cast<const int2*>(&np->child0));
            if (COUNT) cnt->nodes++;
            const float c0lox = fmaf(n0.x, ix, -ox), c0hix = fmaf(n0.y, ix, -ox), c0loy = fmaf(n0.z, iy, -oy), c0hiy = fmaf(n0.w, iy, -oy);
            const float c1lox = fmaf(n1.x, ix, -ox), c1hix = fmaf(n1.y, ix, -ox), c1loy = fmaf(n1.z, iy, -oy), c1hiy = fmaf(n1.w, iy, -oy);
            const float c0loz = fmaf(n2.x, iz, -oz), c0hiz = fmaf(n2.y, iz, -oz), c1loz = fmaf(n2.z, iz, -oz), c1hiz = fmaf(n2.w, iz, -oz);
            const float t0n = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
            const float t0f = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), best.t));
            const float t1n = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
            const float t1f = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), best.t));
            const bool h0 = t0n <= t0f, h1 = t1n <= t1f;
            if (h0 && h1) {
                const bool swap = t1n < t0n;
                stack[sp++] = swap ? ch.x : ch.y;
                node = swap ? ch.y : ch.x;
            } else if (h0) node = ch.x;
            else if (h1) node = ch.y;
            else node = pop();
            if (node < 0 && node != RT_NODE_DONE && parked == RT_NODE_DONE) { parked = node; node = pop(); }
        }
        const bool blocked = node < 0 && node != RT_NODE_DONE;   // a second leaf while one is parked
        const bool any_blocked = __any_sync(0xffffffffu, blocked), any_walking = __any_sync(0xffffffffu, node >= 0);
        if (COUNT) (*steps)++;
        if (any_blocked || !any_walking) {
            if (parked != RT_NODE_DONE) {
                const int prim = ~parked;
                parked = RT_NODE_DONE;
                const DPrim p = sv.prims[prim];
                const uint32_t chain = prim_chain(p);
                if (chain != cached_chain) { cached = to_object_space(sv, chain, ray); cached_chain = chain; }
                if (COUNT) cnt->prim_tests++;
                float t;
                if (prim_t_local(p, chain ? cached : ray, t_min, best.t, sv.semantics, t)) consider(sv, prim, t, best);
            }
            if (blocked) { parked = node; node = pop(); }
            const uint32_t live = __ballot_sync(0xffffffffu, node != RT_NODE_DONE || parked != RT_NODE_DONE);
            if (!live) break;
            if ((uint32_t)__popc(live) <= suspend_below && !__any_sync(0xffffffffu, sp > RT_SUSPEND_STACK)) break;
        }
    }
    w.node = node; w.parked = parked; w.sp = sp;
}

// Streaming closest-hit for a persistent warp (k_extend): rays of one warp finish at very different times — measured on
// the final scene, only 59 % of the node steps a warp pays for carry a live lane (rt_render_stats::node_slots) — so a
// lane that has finished does not wait for the slowest ray of its batch: once fewer than REFILL_BELOW lanes still walk
// the tree, every finished lane stores its result and takes the next ray of the queue.  The node step, leaf parking and
// flush are those of closest_surface<PARKED = true>; one warp reduction per iteration returns the number of walking
// lanes (low half) and of blocked lanes (high half).
// Registers: the node phase needs the inverse direction, the t range, node / stack / parked and the node's boxes.
// Everything else a lane owns — the ray itself, its object-space copy, its pool slot — lives in shared memory
// (LaneSlots) and is read where a leaf is tested; left in registers it pushes the node phase's operands into local
// memory (first version: 9 LDL per node step, 45 % slower than the batch kernel).
//   exchange(fin, fin_mask, valid, best) -> dry     WARP-COLLECTIVE.  Lanes with fin && valid store the result of the ray
//       they hold (media pass included) and clear valid; finished lanes may receive a new ray (valid = true, ray written
//       to the lane's slots).  Returns warp-uniform true once the queue can hand out nothing more.
// The closest hit of a ray does not depend on which lane or warp traces it, nor on what its neighbours do.
struct LaneSlots {      // this lane's shared-memory slots
    float4* ray_o;      // origin, time
    float4* ray_d;      // direction, caller's word
    float4* obj_o;      // object-space origin, wrapper chain the copy belongs to (0 = none)
    float4* obj_d;
    float4* inv_d;      // slab set-up of the node phase: 1 / d and o / d.  Constant per ray, yet RE-READ after every leaf
    float4* inv_o;      //   phase: that ends their register live range there, so the leaf code's temporaries cannot
};                      //   push them into local memory (ptxas spills whole live ranges, hot loop included)
__device__ __forceinline__ Ray lane_ray(const LaneSlots& ls) {
    const float4 o = lds4(ls.ray_o), d = lds4(ls.ray_d);
    Ray r; r.o = make_float3(o.x, o.y, o.z); r.d = make_float3(d.x, d.y, d.z); r.tm = o.w;
    return r;
}

template <bool COUNT, int REFILL_BELOW, class Exchange>
__device__ __forceinline__ void closest_surface_stream(const SceneView& sv, float t_min, const LaneSlots ls, Exchange&& exchange,
                                                       TraversalCounters* cnt, unsigned long long* slots) {
    Closest best; best.t = INFINITY; best.prim = -1;
    int stack[RT_STACK];
    int sp = 0, node = RT_NODE_DONE, parked = RT_NODE_DONE;
    bool valid = false;                     // this lane holds a ray whose result is not stored yet
    uint32_t refill_below = REFILL_BELOW;   // 1 once the queue is dry: then only "nobody walks" ends the node phase
    auto pop = [&]() { return sp ? stack[--sp] : RT_NODE_DONE; };
    sts4(ls.inv_d, make_float4(0.f, 0.f, 0.f, 0.f));
    sts4(ls.inv_o, make_float4(0.f, 0.f, 0.f, 0.f));
    for (;;) {
        // node phase
        uint32_t state;
        bool blocked;
        const float4 id4 = lds4(ls.inv_d), io4 = lds4(ls.inv_o);
        const float ix = id4.x, iy = id4.y, iz = id4.z, ox = io4.x, oy = io4.y, oz = io4.z;
        do {
            if (node >= 0) {
                const BNode* np = sv.nodes + node;
                const float4 n0 = __ldg(&np->n0), n1 = __ldg(&np->n1), n2 = __ldg(&np->n2);
                const int2 ch = __ldg(reinterpret_cast<const int2*>(&np->child0));
                if (COUNT) cnt->nodes++;
                const float c0lox = fmaf(n0.x, ix, -ox), c0hix = fmaf(n0.y, ix, -ox), c0loy = fmaf(n0.z, iy, -oy), c0hiy = fmaf(n0.w, iy, -oy);
                const float c1lox = fmaf(n1.x, ix, -ox), c1hix = fmaf(n1.y, ix, -ox), c1loy = fmaf(n1.z, iy, -oy), c1hiy = fmaf(n1.w, iy, -oy);
                const float c0loz = fmaf(n2.x, iz, -oz), c0hiz = fmaf(n2.y, iz, -oz), c1loz = fmaf(n2.z, iz, -oz), c1hiz = fmaf(n2.w, iz, -oz);
                const float t0n = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
                const float t0f = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), best.t));
                const float t1n = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
                const float t1f = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), best.t));
                const bool h0 = t0n <= t0f, h1 = t1n <= t1f;
                if (h0 && h1) {
                    const bool swap = t1n < t0n;
                    stack[sp++] = swap ? ch.x : ch.y;
                    node = swap ? ch.y : ch.x;
                } else if (h0) node = ch.x;
                else if (h1) node = ch.y;
                else node = pop();
                if (node < 0 && node != RT_NODE_DONE && parked == RT_NODE_DONE) { parked = node; node = pop(); }
            }
            blocked = node < 0 && node != RT_NODE_DONE;   // a second leaf while one is parked
            state = __reduce_add_sync(0xffffffffu, (blocked ? 0x10000u : 0u) | (node >= 0 ? 1u : 0u));
            if (COUNT && (state & 0xffffu)) *slots += 32;
        } while (state - refill_below < 0x10000u - refill_below);   // refill_below <= walking lanes, nobody blocked

        // leaf phase: every parked leaf is tested, blocked lanes park theirs
        if (parked != RT_NODE_DONE) {
            const int prim = ~parked;
            parked = RT_NODE_DONE;
            const DPrim p = sv.prims[prim];
            const uint32_t chain = prim_chain(p);
            Ray r = lane_ray(ls);
            if (chain) {   // object-space copy: consecutive leaves usually share a wrapper chain
                const float4 co = lds4(ls.obj_o);
                if (__float_as_uint(co.w) == chain) {
                    const float4 cd = lds4(ls.obj_d);
                    r.o = make_float3(co.x, co.y, co.z); r.d = make_float3(cd.x, cd.y, cd.z);
                } else {
                    r = to_object_space(sv, chain, r);
                    sts4(ls.obj_o, make_float4(r.o.x, r.o.y, r.o.z, __uint_as_float(chain)));
                    sts4(ls.obj_d, make_float4(r.d.x, r.d.y, r.d.z, 0.f));
                }
            }
            if (COUNT) cnt->prim_tests++;
            float t;
            if (prim_t_local(p, r, t_min, best.t, sv.semantics, t)) consider(sv, prim, t, best);
        }
        if (blocked) { parked = node; node = pop(); }
        const bool fin = node == RT_NODE_DONE && parked == RT_NODE_DONE;
        const uint32_t fm = __ballot_sync(0xffffffffu, fin);
        if (refill_below > 1u ? __popc(fm) > 32 - REFILL_BELOW : fm != 0u) {
            const bool dry = exchange(fin, fm, valid, best);
            if (fin && valid) {
                const Ray r = lane_ray(ls);
                const float tiny = 1e-20f;
                const float ix = 1.0f / (fabsf(r.d.x) < tiny ? copysignf(tiny, r.d.x) : r.d.x);
                const float iy = 1.0f / (fabsf(r.d.y) < tiny ? copysignf(tiny, r.d.y) : r.d.y);
                const float iz = 1.0f / (fabsf(r.d.z) < tiny ? copysignf(tiny, r.d.z) : r.d.z);
                sts4(ls.inv_d, make_float4(ix, iy, iz, 0.f));
                sts4(ls.inv_o, make_float4(r.o.x * ix, r.o.y * iy, r.o.z * iz, 0.f));
                best.t = INFINITY; best.prim = -1;
                sts4(ls.obj_o, make_float4(0.f, 0.f, 0.f, 0.f));
                sp = 0;
                node = sv.n_prims ? 0 : RT_NODE_DONE;
            }
            if (dry) {
                refill_below = 1u;
                if (!__any_sync(0xffffffffu, valid)) break;
            }
        }
    }
}

// The media pass of closest_hit for one lane (not collective).
template <class XiOf>
__device__ __forceinline__ void closest_medium(const SceneView& sv, const Ray& ray, float t_min, Closest& best, XiOf&& xi_of) {
    const float tiny = 1e-20f;
    const float ix = 1.0f / (fabsf(ray.d.x) < tiny ? copysignf(tiny, ray.d.x) : ray.d.x);
    const float iy = 1.0f / (fabsf(ray.d.y) < tiny ? copysignf(tiny, ray.d.y) : ray.d.y);
    const float iz = 1.0f / (fabsf(ray.d.z) < tiny ? copysignf(tiny, ray.d.z) : ray.d.z);
    const float ox = ray.o.x * ix, oy = ray.o.y * iy, oz = ray.o.z * iz;
    for (uint32_t a = 0; a < sv.n_active_media; a++) {
        const uint32_t m = sv.active_media[a];
        const DMedium med = sv.media[m];
        {   // the ray's LINE against the medium's (padded) world box: a miss means no boundary crossing at any t.  And the
            // SEGMENT [t_min, closest surface]: constant_medium::hit clips entry and exit to it and gives up when nothing is
            // left (constant_medium.h:42-46), which is certain when the box is entered behind the closest surface or left
            // before t_min — no need to intersect the boundary to find that out
            const float x0 = fmaf(med.lo.x, ix, -ox), x1 = fmaf(med.hi.x, ix, -ox), y0 = fmaf(med.lo.y, iy, -oy), y1 = fmaf(med.hi.y, iy, -oy);
            const float z0 = fmaf(med.lo.z, iz, -oz), z1 = fmaf(med.hi.z, iz, -oz);
            const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fminf(z0, z1));
            const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
            if (!(tn <= tf) || tn > best.t || tf < t_min) continue;
        }
        float t;
        if (medium_t(sv, med, ray, t_min, best.t, [&]() { return xi_of(m); }, t)) { best.t = t; best.prim = -2 - (int)m; }
    }
}

// Media are evaluated after the surfaces with t_max = closest surface.  In the reference a medium sees only
// the objects listed BEFORE it (hittable_list.h:29-35); because constant_medium clips its exit to t_max before
// sampling (constant_medium.h:44) the result is the same except where the sampled distance lands within one
// rounding step of a later surface.  xi_of(m): the uniform draw of medium m (scripted for first-hit queries,
// Philox in the renderer); only evaluated when the ray really crosses the boundary.
// Lanes without a ray pass active = false (callers keep whole warps converged around the query).
template <bool COUNT, bool PARKED = false, class XiOf>
__device__ __forceinline__ Closest closest_hit(const SceneView& sv, const Ray& ray, float t_min, bool with_media, XiOf&& xi_of,
                                               TraversalCounters* cnt, bool active = true) {
    Closest best; best.t = INFINITY; best.prim = -1;
    closest_surface<COUNT, PARKED>(sv, ray, t_min, best, cnt, active);
    if (with_media && active && sv.n_active_media) closest_medium(sv, ray, t_min, best, xi_of);
    return best;
}

} // namespace rtb
