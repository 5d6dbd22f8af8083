// traverse.cuh — closest-hit query over the LBVH + media: what color_f asks of the world (render.h:63).
// Replaces hittable_list::hit / bvh_node::hit (hittable_list.h:23, bvh.h:348): no device new/delete, no powf,
// ordered descent with a shrinking t_max, one 64-byte node fetch per two child boxes.
#pragma once

#include "intersect.cuh"
#include "rng.cuh"
#include <climits>

namespace rtb {

#define RT_STACK (RT_STACK_DEPTH + 1)   // + the sentinel entry; Karras depth bound: 63 key bits + index bits for duplicate keys; PLOC depth is checked at build

struct Closest {
    float t;     // +inf on miss
    int prim;    // >= 0: surface primitive (device order); -1: miss; <= -2: medium index = -2 - prim
};

struct TraversalCounters { uint32_t nodes, prim_tests; };

#define RT_NODE_DONE INT_MIN   // "no node": a finished lane, an empty stack, an empty slot of a wide node

// ties: the reference's hittable_list lets the LATER object win (t == closest_so_far is accepted,
// sphere.h:49, aarect.h:66) — with an unordered traversal that becomes "higher leaf id wins".
__device__ __forceinline__ void consider(const SceneView& sv, int prim, float t, Closest& best) {
    if (t < best.t || best.prim < 0 || sv.prim_leaf_id[prim] > sv.prim_leaf_id[best.prim]) { best.t = t; best.prim = prim; }
}

// ---- the node step --------------------------------------------------------------------------------------------------
// One 64-byte node = both children's boxes.  Slab distances are t = plane * (1/d) - o * (1/d), one FMA per plane.
// Which plane of a pair is the near one depends only on the sign of the ray direction, so the min / max of each pair can
// be chosen ahead of time: with a = max(1/d, 0) and b = min(1/d, 0) (one of them is zero),
//     near = lo * a + (hi * b - o/d)        far = lo * b + (hi * a - o/d)
// are the SAME floats as min / max of the two FMAs (0 * plane adds an exact zero), computed on the FMA pipe — ncu showed
// k_extend limited by the ALU pipe (FMNMX, ISETP, SEL: 67 % busy) with the FMA pipe at 15 % (profiles/r01_v8_sah.md), and
// this moves 12 of the 20 min / max of a node step over: 24 FFMA + 8 FMNMX(3) instead of 12 FFMA + 20 FMNMX(3).
struct SlabRay {
    float ax, bx, ay, by, az, bz;   // positive / negative part of 1 / d per axis
    float ox, oy, oz;               // o / d
};
__device__ __forceinline__ SlabRay slab_ray(const Ray& ray) {
    // zero direction components are nudged so 1 / d stays finite
    const float tiny = 1e-20f;
    const float ix = 1.0f / (fabsf(ray.d.x) < tiny ? copysignf(tiny, ray.d.x) : ray.d.x);
    const float iy = 1.0f / (fabsf(ray.d.y) < tiny ? copysignf(tiny, ray.d.y) : ray.d.y);
    const float iz = 1.0f / (fabsf(ray.d.z) < tiny ? copysignf(tiny, ray.d.z) : ray.d.z);
    SlabRay s;
    s.ax = fmaxf(ix, 0.f); s.bx = fminf(ix, 0.f); s.ay = fmaxf(iy, 0.f); s.by = fminf(iy, 0.f); s.az = fmaxf(iz, 0.f); s.bz = fminf(iz, 0.f);
    s.ox = ray.o.x * ix; s.oy = ray.o.y * iy; s.oz = ray.o.z * iz;
    return s;
}
struct NodeHit { bool h0, h1; float t0n, t1n; int c0, c1; };
template <int V>
__device__ __forceinline__ NodeHit node_step(const BNode* np, const SlabRay& s, float t_min, float t_max) {
    const float4 n0 = __ldg(&np->n0), n1 = __ldg(&np->n1), n2 = __ldg(&np->n2);
    const int2 ch = __ldg(reinterpret_cast<const int2*>(&np->child0));
    NodeHit r;
    float t0f, t1f;
    if constexpr ((V & 4) == 0) {
        const float c0nx = fmaf(n0.x, s.ax, fmaf(n0.y, s.bx, -s.ox)), c0fx = fmaf(n0.x, s.bx, fmaf(n0.y, s.ax, -s.ox));
        const float c0ny = fmaf(n0.z, s.ay, fmaf(n0.w, s.by, -s.oy)), c0fy = fmaf(n0.z, s.by, fmaf(n0.w, s.ay, -s.oy));
        const float c1nx = fmaf(n1.x, s.ax, fmaf(n1.y, s.bx, -s.ox)), c1fx = fmaf(n1.x, s.bx, fmaf(n1.y, s.ax, -s.ox));
        const float c1ny = fmaf(n1.z, s.ay, fmaf(n1.w, s.by, -s.oy)), c1fy = fmaf(n1.z, s.by, fmaf(n1.w, s.ay, -s.oy));
        const float c0nz = fmaf(n2.x, s.az, fmaf(n2.y, s.bz, -s.oz)), c0fz = fmaf(n2.x, s.bz, fmaf(n2.y, s.az, -s.oz));
        const float c1nz = fmaf(n2.z, s.az, fmaf(n2.w, s.bz, -s.oz)), c1fz = fmaf(n2.z, s.bz, fmaf(n2.w, s.az, -s.oz));
        r.t0n = fmaxf(fmaxf(c0nx, c0ny), fmaxf(c0nz, t_min));
        t0f = fminf(fminf(c0fx, c0fy), fminf(c0fz, t_max));
        r.t1n = fmaxf(fmaxf(c1nx, c1ny), fmaxf(c1nz, t_min));
        t1f = fminf(fminf(c1fx, c1fy), fminf(c1fz, t_max));
    } else {   // TV_MINMAX: round 1's arithmetic (A/B)
        const float ix = s.ax + s.bx, iy = s.ay + s.by, iz = s.az + s.bz;
        const float c0lox = fmaf(n0.x, ix, -s.ox), c0hix = fmaf(n0.y, ix, -s.ox), c0loy = fmaf(n0.z, iy, -s.oy), c0hiy = fmaf(n0.w, iy, -s.oy);
        const float c1lox = fmaf(n1.x, ix, -s.ox), c1hix = fmaf(n1.y, ix, -s.ox), c1loy = fmaf(n1.z, iy, -s.oy), c1hiy = fmaf(n1.w, iy, -s.oy);
        const float c0loz = fmaf(n2.x, iz, -s.oz), c0hiz = fmaf(n2.y, iz, -s.oz), c1loz = fmaf(n2.z, iz, -s.oz), c1hiz = fmaf(n2.w, iz, -s.oz);
        r.t0n = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), t_min));
        t0f = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), t_max));
        r.t1n = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), t_min));
        t1f = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), t_max));
    }
    r.h0 = r.t0n <= t0f; r.h1 = r.t1n <= t1f;
    r.c0 = ch.x; r.c1 = ch.y;
    return r;
}

// ---- the wide node step (WNode: four children, 8-bit boxes) ---------------------------------------------------------------
// Per node: three scales 2^e and the three offsets (origin - ray origin) / d, then per child and axis one byte -> float
// (PRMT into the mantissa of 2^23, minus 2^23: exact) and one FMA per plane.  Near / far planes are chosen for all four
// children at once by swapping the packed lo / hi words of an axis when the direction is negative.  The four candidates
// are sorted by entry distance (misses last) with a five-comparator network.  RT_WIDE_FOLD (default): the byte goes into the
// mantissa of 1.0 instead (1 + q 2^-15) and the subtraction is folded into the plane offset — 21 instructions fewer per step, an
// extra error of <= 1/512 step that a 1/128-step guard band in the builder covers: +0.4 % node visits, extend -2.4 %, mesh +2 %.
// Measured and not adopted (profiles/r02_wide_nodes.md): the same fold through 2^23 (error of half a step: the builder must pad
// every box by a whole step, +7 % node visits, extend +1.8 %), and sorting only the nearest child to the front (+1 % visits, +0.8 %).
struct WideHits { float t[4]; int c[4]; };   // ascending entry distance; a miss is (+inf, RT_NODE_DONE)
__device__ __forceinline__ float byte_to_float(uint32_t w, uint32_t sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)) - 8388608.0f; }
__device__ __forceinline__ WideHits wide_step(const WNode* np, const SlabRay& s, float t_min, float t_max) {
    const uint4 w0 = __ldg(reinterpret_cast<const uint4*>(np)), w1 = __ldg(reinterpret_cast<const uint4*>(np) + 1);
    const uint4 w2 = __ldg(reinterpret_cast<const uint4*>(np) + 2), w3 = __ldg(reinterpret_cast<const uint4*>(np) + 3);
    const float ix = s.ax + s.bx, iy = s.ay + s.by, iz = s.az + s.bz;                 // one of each pair is zero: exact
    const float sx = __uint_as_float((w0.w & 0xffu) << 23) * ix, sy = __uint_as_float(((w0.w >> 8) & 0xffu) << 23) * iy,
                sz = __uint_as_float(((w0.w >> 16) & 0xffu) << 23) * iz;
    const float bx = fmaf(__uint_as_float(w0.x), ix, -s.ox), by = fmaf(__uint_as_float(w0.y), iy, -s.oy), bz = fmaf(__uint_as_float(w0.z), iz, -s.oz);
    const bool px = ix >= 0.f, py = iy >= 0.f, pz = iz >= 0.f;
    const uint32_t nx = px ? w2.x : w2.w, fx = px ? w2.w : w2.x;                      // qlo_x / qhi_x
    const uint32_t ny = py ? w2.y : w3.x, fy = py ? w3.x : w2.y;                      // qlo_y / qhi_y
    const uint32_t nz = pz ? w2.z : w3.y, fz = pz ? w3.y : w2.z;                      // qlo_z / qhi_z
    const int code[4] = {(int)w1.x, (int)w1.y, (int)w1.z, (int)w1.w};
    WideHits h;
#if RT_WIDE_FOLD
    // plane byte q as the float 1 + q * 2^-15 (one PRMT into the mantissa of 1.0): t = (1 + q 2^-15) * (2^15 s) + (b - 2^15 s), one
    // FMA and no subtraction per plane.  The offset b - 2^15 s is rounded: an error of up to 1/512 of a quantisation step, which the
    // builder's 1/128-step guard band around every plane covers (bvh_build.cu, RT_WIDE_FOLD)
    const float kx = sx * 32768.0f, ky = sy * 32768.0f, kz = sz * 32768.0f;
    const float cx = bx - kx, cy = by - ky, cz = bz - kz;
    #define RT_WPLANE(w, sel, k, c) fmaf(__uint_as_float(__byte_perm(w, 0x3F800000u, sel)), k, c)
    #pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint32_t sel = 0x7604u | ((uint32_t)c << 4);
        const float tn = fmaxf(fmaxf(RT_WPLANE(nx, sel, kx, cx), RT_WPLANE(ny, sel, ky, cy)), fmaxf(RT_WPLANE(nz, sel, kz, cz), t_min));
        const float tf = fminf(fminf(RT_WPLANE(fx, sel, kx, cx), RT_WPLANE(fy, sel, ky, cy)), fminf(RT_WPLANE(fz, sel, kz, cz), t_max));
#else
    #pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint32_t sel = 0x7540u | (uint32_t)c;
        const float tn = fmaxf(fmaxf(fmaf(byte_to_float(nx, sel), sx, bx), fmaf(byte_to_float(ny, sel), sy, by)), fmaxf(fmaf(byte_to_float(nz, sel), sz, bz), t_min));
        const float tf = fminf(fminf(fmaf(byte_to_float(fx, sel), sx, bx), fmaf(byte_to_float(fy, sel), sy, by)), fminf(fmaf(byte_to_float(fz, sel), sz, bz), t_max));
#endif
        const bool hit = tn <= tf && code[c] != RT_WIDE_EMPTY;
        h.t[c] = hit ? tn : INFINITY;
        h.c[c] = hit ? code[c] : RT_NODE_DONE;
    }
    auto cswap = [&](int a, int b) {
        const bool sw = h.t[b] < h.t[a];
        const float ta = sw ? h.t[b] : h.t[a], tb = sw ? h.t[a] : h.t[b];
        const int ca = sw ? h.c[b] : h.c[a], cb = sw ? h.c[a] : h.c[b];
        h.t[a] = ta; h.t[b] = tb; h.c[a] = ca; h.c[b] = cb;
    };
    cswap(0, 1); cswap(2, 3); cswap(0, 2); cswap(1, 3); cswap(1, 2);
    return h;
}

// ---- the traversal stack --------------------------------------------------------------------------------------------
// Variants (template bit mask V), each measured against the plain one on the same box (profiles/r02_*):
//   TV_TCULL  an entry carries the distance at which the ray enters the pushed box; a pop discards entries that the
//             closest hit found meanwhile has made unreachable WITHOUT fetching their node (the entry test equals the
//             box test the parent made, against the smaller t_max: same closest hit, fewer node visits)
//   TV_SMEM   the stack lives in shared memory, [entry][thread]: every access is conflict-free and costs one wavefront
//             however the lanes' depths differ — in local memory a warp's access costs one wavefront per distinct depth, on
//             the L1 data pipe the node fetches already keep 66 % busy.  RT_SHORT_STACK entries: chosen by the host for
//             trees no deeper than that (a stack never holds more entries than the tree has levels); deeper trees walk
//             with the local-memory stack.
//   TV_MINMAX min / max plane selection on the ALU pipe, as in round 1
//   TV_WIDE   the 4-wide compressed tree: a step sorts up to four entered children, descends into the nearest and pushes
//             the rest far-to-near (RT_NODE_DONE is INT_MIN = RT_WIDE_EMPTY: "no node" in both encodings)
#define RT_SHORT_STACK 16
#ifndef RT_STACK_SENTINEL
#define RT_STACK_SENTINEL 1   // final scene +1.4 %, big_scene1 +1.9 % (profiles/r02_traversal_ab.md)
#endif
#ifndef RT_STEP_UNROLL_DEFER
#define RT_STEP_UNROLL_DEFER 2   // the same for walk_suspendable (1 M-triangle mesh: 1 / 2 / 3 / 4 = 114.5 / 111.3 / 116.1 / 121.1 ms)
#endif
#ifndef RT_STEP_UNROLL
#define RT_STEP_UNROLL 2   // node steps between two looks at the leaf-phase condition (two votes + a branch per look): 1 / 2 / 3 / 4 =
                           // final scene 74.99 / 72.83 / 73.09 / 72.58 ms, mesh 114.5 / 111.3 / 116.1 / 121.1 (profiles/r02_traversal_ab.md)
#endif
enum : int { TV_PLAIN = 0, TV_TCULL = 1, TV_SMEM = 2, TV_MINMAX = 4, TV_WIDE = 8 /* walks SceneView::wide (4-wide, 8-bit boxes); plain stack only */,
             TV_PACK = 16 /* two parked leaves per lane, the warp's leaf tests packed into its lowest lanes (closest_surface_packed) */ };

// Closest surface hit.  A lane that reaches a leaf parks it (one register) and keeps walking; the warp enters the leaf code
// only when some lane reaches a second leaf or nobody has nodes left, then every parked leaf is tested together (round 1:
// fused node-or-leaf iterations spent 35 % of their issue slots in leaf code at 3.2 active lanes).  Parking only delays the
// shrinking of t_max and consider() is order-independent, so the closest hit is the same.
// WARP-COLLECTIVE (votes): every lane of the warp must call, idle lanes with active = false.
template <bool COUNT, int V, int STRIDE>
__device__ __forceinline__ void closest_surface(const SceneView& sv, const Ray& ray, float t_min, Closest& best, TraversalCounters* cnt, bool active,
                                                int* stack_column) {
    const SlabRay s = slab_ray(ray);
    // object-space ray cache: consecutive leaves usually share a wrapper chain (e.g. 1000 spheres under one)
    uint32_t cached_chain = 0;
    Ray cached = ray;
    constexpr bool TC = (V & TV_TCULL) != 0, SM = (V & TV_SMEM) != 0;
    int node_loc[SM ? 1 : RT_STACK];                    // local-memory stack (only the arrays: sp stays a register)
    float tn_loc[(SM || !TC) ? 1 : RT_STACK];
    int sp = 0;
#if RT_STACK_SENTINEL
    // entry 0 is a sentinel "no node": a pop never has to ask whether the stack is empty (a lane that pops the sentinel is done and
    // never pops again)
    if constexpr (!TC && !SM) { node_loc[0] = RT_NODE_DONE; sp = 1; }
#endif
    // shared-memory column of this thread: node words, then RT_SHORT_STACK rows further the entry distances (TC)
    auto push = [&](int n, float tn) {
        RT_DEV_ASSERT(sp < (SM ? RT_SHORT_STACK : RT_STACK));
        RT_DEV_ASSERT(n < 0 ? (uint32_t)~n < sv.n_prims : (uint32_t)n < ((V & TV_WIDE) ? sv.n_wide : sv.n_nodes));
        if constexpr (SM) { stack_column[sp * STRIDE] = n; if constexpr (TC) stack_column[(RT_SHORT_STACK + sp) * STRIDE] = __float_as_int(tn); }
        else { node_loc[sp] = n; if constexpr (TC) tn_loc[sp] = tn; }
        sp++;
    };
    auto pop = [&](float t_max) {
        if constexpr (!TC) {
#if RT_STACK_SENTINEL
            if constexpr (!SM) { RT_DEV_ASSERT(sp > 0); sp--; return node_loc[sp]; }
#endif
            if (!sp) return RT_NODE_DONE;
            sp--;
            if constexpr (SM) return stack_column[sp * STRIDE]; else return node_loc[sp];
        } else {
            while (sp) {
                sp--;
                int n; float tn;
                if constexpr (SM) { n = stack_column[sp * STRIDE]; tn = __int_as_float(stack_column[(RT_SHORT_STACK + sp) * STRIDE]); }
                else { n = node_loc[sp]; tn = tn_loc[sp]; }
                if (!(tn > t_max)) return n;
            }
            return RT_NODE_DONE;
        }
    };
    int node = (active && sv.n_prims) ? 0 : RT_NODE_DONE;
    int parked = RT_NODE_DONE;
    for (;;) {
        #pragma unroll
        for (int unrolled = 0; unrolled < RT_STEP_UNROLL; unrolled++)
        if (node >= 0) {
            if constexpr ((V & TV_WIDE) != 0) {
                const WideHits h = wide_step(sv.wide + node, s, t_min, best.t);
                if (COUNT) cnt->nodes++;
                if (h.c[3] != RT_NODE_DONE) push(h.c[3], h.t[3]);
                if (h.c[2] != RT_NODE_DONE) push(h.c[2], h.t[2]);
                if (h.c[1] != RT_NODE_DONE) push(h.c[1], h.t[1]);
                node = h.c[0] != RT_NODE_DONE ? h.c[0] : pop(best.t);
            } else {
                const NodeHit h = node_step<V>(sv.nodes + node, s, t_min, best.t);
                if (COUNT) cnt->nodes++;
                if (h.h0 && h.h1) {
                    const bool swap = h.t1n < h.t0n;
                    push(swap ? h.c0 : h.c1, swap ? h.t0n : h.t1n);
                    node = swap ? h.c1 : h.c0;
                } else if (h.h0) node = h.c0;
                else if (h.h1) node = h.c1;
                else node = pop(best.t);
            }
            if (node < 0 && node != RT_NODE_DONE && parked == RT_NODE_DONE) { parked = node; node = pop(best.t); }
        }
        const bool blocked = node < 0 && node != RT_NODE_DONE;   // a second leaf while one is parked
        // two votes, not one __reduce_or_sync of a 2-bit state: VOTE.ANY writes a predicate the branch consumes directly,
        // REDUX goes through a uniform register (+2 % on the final scene and the 1M-triangle mesh, same-box A/B)
        const bool any_blocked = __any_sync(0xffffffffu, blocked), any_walking = __any_sync(0xffffffffu, node >= 0);
        if (any_blocked || !any_walking) {
            if (parked != RT_NODE_DONE) {
                const int prim = ~parked;
                parked = RT_NODE_DONE;
                RT_DEV_ASSERT((uint32_t)prim < sv.n_prims);
                const DPrim p = sv.prims[prim];
                const uint32_t chain = prim_chain(p);
                if (chain != cached_chain) { cached = to_object_space(sv, chain, ray); cached_chain = chain; }
                if (COUNT) cnt->prim_tests++;
                float t;
                if (prim_t_local(p, chain ? cached : ray, t_min, best.t, sv.semantics, t)) consider(sv, prim, t, best);
            }
            if (blocked) { parked = node; node = pop(best.t); }
            if (!__any_sync(0xffffffffu, node != RT_NODE_DONE || parked != RT_NODE_DONE)) break;
        }
    }
}

// TV_PACK: the leaf phase of closest_surface runs at 4-7 of 32 lanes (a lane parks ONE leaf, the phase starts when any lane meets
// its second) and is a fifth of k_extend's instructions.  Here a lane parks up to TWO leaves and the phase hands the warp's
// pending tests to its lowest lanes — test i runs on lane i with the owner's ray fetched by shuffles, the distance goes back the
// same way — so a phase that starts later (a lane's THIRD leaf) tests about twice as many leaves in one pass.  Costs: t_max
// shrinks later (more node visits), no object-space ray cache, ~30 shuffles per phase.  Same closest hit: consider() is
// order-independent once a test made against an older t_max is re-checked against the current one.
template <bool COUNT>
__device__ __forceinline__ void closest_surface_packed(const SceneView& sv, const Ray& ray, float t_min, Closest& best, TraversalCounters* cnt, bool active) {
    const SlabRay s = slab_ray(ray);
    int node_loc[RT_STACK];
    int sp = 0;
    auto pop = [&]() { return sp ? node_loc[--sp] : RT_NODE_DONE; };
    auto is_leaf = [](int n) { return n < 0 && n != RT_NODE_DONE; };
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    int node = (active && sv.n_prims) ? 0 : RT_NODE_DONE;
    int park0 = RT_NODE_DONE, park1 = RT_NODE_DONE;
    for (;;) {
        if (node >= 0) {
            const NodeHit h = node_step<0>(sv.nodes + node, s, t_min, best.t);
            if (COUNT) cnt->nodes++;
            if (h.h0 && h.h1) {
                const bool swap = h.t1n < h.t0n;
                RT_DEV_ASSERT(sp < RT_STACK);
                node_loc[sp++] = swap ? h.c0 : h.c1;
                node = swap ? h.c1 : h.c0;
            } else if (h.h0) node = h.c0;
            else if (h.h1) node = h.c1;
            else node = pop();
            if (is_leaf(node) && park0 == RT_NODE_DONE) { park0 = node; node = pop(); }
            if (is_leaf(node) && park1 == RT_NODE_DONE) { park1 = node; node = pop(); }
        }
        const bool blocked = is_leaf(node);                        // a third leaf while two are parked
        const bool any_blocked = __any_sync(0xffffffffu, blocked), any_walking = __any_sync(0xffffffffu, node >= 0);
        if (any_blocked || !any_walking) {
            const uint32_t m0 = __ballot_sync(0xffffffffu, park0 != RT_NODE_DONE), m1 = __ballot_sync(0xffffffffu, park1 != RT_NODE_DONE);
            const uint32_t n0 = __popc(m0), total = n0 + __popc(m1);
            const uint32_t i0 = __popc(m0 & lt), i1 = n0 + __popc(m1 & lt);   // where this lane's tests sit in the packed order
            for (uint32_t base = 0; base < total; base += 32u) {              // one pass unless more than 32 leaves are pending
                const uint32_t i = base + lane;
                const bool second = i >= n0, mine = i < total;
                const uint32_t owner = mine ? __fns(second ? m1 : m0, 0u, (int)(second ? i - n0 : i) + 1) : lane;
                const int l0 = __shfl_sync(0xffffffffu, park0, owner), l1 = __shfl_sync(0xffffffffu, park1, owner);
                Ray r;
                r.o.x = __shfl_sync(0xffffffffu, ray.o.x, owner); r.o.y = __shfl_sync(0xffffffffu, ray.o.y, owner); r.o.z = __shfl_sync(0xffffffffu, ray.o.z, owner);
                r.d.x = __shfl_sync(0xffffffffu, ray.d.x, owner); r.d.y = __shfl_sync(0xffffffffu, ray.d.y, owner); r.d.z = __shfl_sync(0xffffffffu, ray.d.z, owner);
                r.tm = __shfl_sync(0xffffffffu, ray.tm, owner);
                const float t_max = __shfl_sync(0xffffffffu, best.t, owner);
                float t_hit = 0.f;
                bool hit = false;
                if (mine) {
                    const int prim = ~(second ? l1 : l0);
                    RT_DEV_ASSERT((uint32_t)prim < sv.n_prims);
                    const DPrim p = sv.prims[prim];
                    const uint32_t chain = prim_chain(p);
                    if (chain) r = to_object_space(sv, chain, r);
                    if (COUNT) cnt->prim_tests++;
                    float t;
                    if (prim_t_local(p, r, t_min, t_max, sv.semantics, t)) { t_hit = t; hit = true; }
                }
                const uint32_t hits = __ballot_sync(0xffffffffu, hit);
                const float r0 = __shfl_sync(0xffffffffu, t_hit, (i0 - base) & 31u), r1 = __shfl_sync(0xffffffffu, t_hit, (i1 - base) & 31u);
                // a test made against the t_max of the phase's start counts only if it still is within the current one (written as the
                // reference writes its range test, so that a NaN distance behaves as it does there)
                if (park0 != RT_NODE_DONE && i0 - base < 32u && ((hits >> (i0 - base)) & 1u) && !(r0 > best.t)) consider(sv, ~park0, r0, best);
                if (park1 != RT_NODE_DONE && i1 - base < 32u && ((hits >> (i1 - base)) & 1u) && !(r1 > best.t)) consider(sv, ~park1, r1, best);
            }
            park0 = park1 = RT_NODE_DONE;
            if (is_leaf(node)) { park0 = node; node = pop(); }
            if (is_leaf(node)) { park1 = node; node = pop(); }
            if (!__any_sync(0xffffffffu, node != RT_NODE_DONE || park0 != RT_NODE_DONE)) break;
        }
    }
}

// Suspendable walk (k_extend_defer): closest_surface whose traversal state is the caller's, so a walk can
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

template <bool COUNT, bool WIDE>
__device__ __forceinline__ void walk_suspendable(const SceneView& sv, const Ray& ray, float t_min, Closest& best, TraversalCounters* cnt,
                                                 Walk& w, int* stack, uint32_t suspend_below, uint32_t* steps) {
    const SlabRay s = slab_ray(ray);
    uint32_t cached_chain = 0;
    Ray cached = ray;
    int node = w.node, parked = w.parked, sp = w.sp;
#if RT_STACK_SENTINEL
    auto pop = [&]() { RT_DEV_ASSERT(sp > 0); return stack[--sp]; };   // stack[0] is the caller's sentinel RT_NODE_DONE (see closest_surface)
#else
    auto pop = [&]() { return sp ? stack[--sp] : RT_NODE_DONE; };
#endif
    for (;;) {
        #pragma unroll
        for (int unrolled = 0; unrolled < RT_STEP_UNROLL_DEFER; unrolled++)
        if (node >= 0) {
            if constexpr (WIDE) {
                const WideHits h = wide_step(sv.wide + node, s, t_min, best.t);
                if (COUNT) cnt->nodes++;
                RT_DEV_ASSERT(sp + 3 <= RT_STACK);
                if (h.c[3] != RT_NODE_DONE) stack[sp++] = h.c[3];
                if (h.c[2] != RT_NODE_DONE) stack[sp++] = h.c[2];
                if (h.c[1] != RT_NODE_DONE) stack[sp++] = h.c[1];
                node = h.c[0] != RT_NODE_DONE ? h.c[0] : pop();
            } else {
                const NodeHit h = node_step<0>(sv.nodes + node, s, t_min, best.t);
                if (COUNT) cnt->nodes++;
                if (h.h0 && h.h1) {
                    const bool swap = h.t1n < h.t0n;
                    RT_DEV_ASSERT(sp < RT_STACK);
                    stack[sp++] = swap ? h.c0 : h.c1;
                    node = swap ? h.c1 : h.c0;
                } else if (h.h0) node = h.c0;
                else if (h.h1) node = h.c1;
                else node = pop();
            }
            if (node < 0 && node != RT_NODE_DONE && parked == RT_NODE_DONE) { parked = node; node = pop(); }
        }
        const bool blocked = node < 0 && node != RT_NODE_DONE;   // a second leaf while one is parked
        const bool any_blocked = __any_sync(0xffffffffu, blocked), any_walking = __any_sync(0xffffffffu, node >= 0);
        if (COUNT) (*steps) += RT_STEP_UNROLL_DEFER;
        if (any_blocked || !any_walking) {
            if (parked != RT_NODE_DONE) {
                const int prim = ~parked;
                parked = RT_NODE_DONE;
                RT_DEV_ASSERT((uint32_t)prim < sv.n_prims);
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
template <bool COUNT, int V, int STRIDE, class XiOf>
__device__ __forceinline__ Closest closest_hit(const SceneView& sv, const Ray& ray, float t_min, bool with_media, XiOf&& xi_of,
                                               TraversalCounters* cnt, bool active, int* stack_column) {
    Closest best; best.t = INFINITY; best.prim = -1;
    if constexpr ((V & TV_PACK) != 0) closest_surface_packed<COUNT>(sv, ray, t_min, best, cnt, active);
    else closest_surface<COUNT, V, STRIDE>(sv, ray, t_min, best, cnt, active, stack_column);
    if (with_media && active && sv.n_active_media) closest_medium(sv, ray, t_min, best, xi_of);
    return best;
}

} // namespace rtb
