// Traversal of the compressed 8-wide BVH (Bvh8Node, 80 bytes = five 16-byte loads per
// node) with a short stack of (child group base, pending-hit mask) pairs; children are
// visited in octant order, so no distance sort is needed.  The traversal is written as a
// resumable state machine with two kinds of step -- node_step (open one inner child, slab-test
// its 8 children) and prim_step (test one pending primitive) -- so the wavefront trace kernels
// (lj_wavefront.cu) can (a) hand a finished lane a new ray and (b) let the warp VOTE on which
// phase to run next, postponing primitive tests until enough lanes have one pending: the two
// phases then run with mostly full warps instead of serialising behind each other.  The
// whole-path engines simply run a ray to completion (trace_closest / trace_any below, the
// rtcIntersect1 / rtcOccluded1 of this build).  Postponing changes only the ORDER of primitive
// tests; the closest hit is the lexicographic minimum of (t, geomID, primID) and therefore
// independent of that order.
// Primitive tests and the closest-hit rule are those of lj_traverse.cuh; the running best
// hit is kept as (t, leaf slot) only and the full record (u, v, Ng) is recomputed once at
// the end with the same arithmetic.
#pragma once
#include "lj_traverse.cuh"

#ifndef LJ_BVH8_STACK
#define LJ_BVH8_STACK 32
#endif
#ifndef LJ_STACK_TOP_CACHE
#define LJ_STACK_TOP_CACHE 0
#endif

LJ_D unsigned byte_of(unsigned w, int i) { return (w >> (8 * i)) & 0xffu; }

// byte j of w spliced into the mantissa of 2^15: returns the float 32768 + byte (exact)
// (`m15` = 0x47000000, handed in as a runtime value -- DScene::m15_bits -- so that the selector is PRMT's immediate)
LJ_D float byte_to_m15(unsigned w, int j, unsigned m15) {
    return __uint_as_float(__byte_perm(w, m15, 0x7404u | ((unsigned)j << 4)));
}

struct Trav8 {
    RayF r;
    float idx, idy, idz;
    unsigned oct_inv;
    float best_t; int best_slot;
    uint2 ng;        // node group: x = child base, y = (pending inner-child hits << 24) | imask
    uint2 tg;        // primitive group waiting to be tested: x = first slot, y = 24-bit pending mask
    int sp;
    uint2 top;                    // the stack's top entry lives in registers (LJ_STACK_TOP_CACHE): logical stack = stack[0 .. sp-2] + top
    uint2 stack[LJ_BVH8_STACK];   // node groups (y & 0xff000000 != 0) and postponed primitive groups (top byte 0)
    unsigned n_nodes, n_prims;

    LJ_D void init(const RayF &ray) {
        r = ray;
        // inverse direction with zero components replaced by a huge finite value of the right sign
        const float sx = fabsf(r.dx) > 1e-20f ? r.dx : copysignf(1e-20f, r.dx);
        const float sy = fabsf(r.dy) > 1e-20f ? r.dy : copysignf(1e-20f, r.dy);
        const float sz = fabsf(r.dz) > 1e-20f ? r.dz : copysignf(1e-20f, r.dz);
        idx = 1.0f / sx; idy = 1.0f / sy; idz = 1.0f / sz;
        oct_inv = (sx >= 0.f ? 1u : 0u) | (sy >= 0.f ? 2u : 0u) | (sz >= 0.f ? 4u : 0u);
        best_t = r.tfar; best_slot = -1;
        ng = make_uint2(0u, 0x80000000u);      // root: one pending inner child, empty imask
        tg = make_uint2(0u, 0u);
        sp = 0; top = make_uint2(0u, 0u); n_nodes = 0; n_prims = 0;
    }

    // LJ_STACK_TOP_CACHE=1 (experiment, OFF): push / pop with the top entry cached in `top` -- a pop hands out the register
    // and starts the load of the entry below, which is not needed before the NEXT pop, so the local-memory latency leaves
    // the refill's critical path (the refill loop is 19 % of the extend kernel's stall samples at 4-5 active lanes).  The
    // two extra registers do not fit the 48 the trace kernels are compiled for: k_wf_extend's spills went from 84 to 404
    // bytes and sponza's extend stage from 74.4 to 84.4 ms (profiles/r02_experiments.md).  Same entries, same order.
#if LJ_STACK_TOP_CACHE
    LJ_D void push(uint2 e) { if (sp > 0) stack[sp - 1] = top; top = e; sp++; }
    LJ_D uint2 peek() const { return top; }
    LJ_D void pop() { sp--; if (sp > 0) top = stack[sp - 1]; }
#else
    LJ_D void push(uint2 e) { stack[sp++] = e; }
    LJ_D uint2 peek() const { return stack[sp - 1]; }
    LJ_D void pop() { sp--; }
#endif
    LJ_D bool wants_node() const { return (ng.y & 0xff000000u) != 0u; }
    LJ_D bool wants_prim() const { return tg.y != 0u; }
    LJ_D bool active() const { return wants_node() || wants_prim(); }

    // closest-hit rule on ties: smaller (geomID, primID) wins
    LJ_D bool tie_less(const DScene &sc, int slot_a, int slot_b) const {
        int ga = __float_as_int(__ldg(&sc.tri_verts[3 * slot_a]).w), gb = __float_as_int(__ldg(&sc.tri_verts[3 * slot_b]).w);
        if (ga != gb) return ga < gb;
        return __float_as_int(__ldg(&sc.tri_verts[3 * slot_a + 1]).w) < __float_as_int(__ldg(&sc.tri_verts[3 * slot_b + 1]).w);
    }

    // test one primitive slot; returns true if it is an accepted hit (ANY: stop)
    template <bool ANY>
    LJ_D bool prim(const DScene &sc, int slot) {
        const float4 a = __ldg(&sc.tri_verts[3 * slot + 0]);
        const float4 b = __ldg(&sc.tri_verts[3 * slot + 1]);
        const float4 c = __ldg(&sc.tri_verts[3 * slot + 2]);
        float t;
        if (__float_as_int(c.w) == 0) {
            TriHit h;
            const f3 O{r.ox, r.oy, r.oz}, D{r.dx, r.dy, r.dz};
            if (!tri_test(O, D, r.tnear, r.tfar, f3{a.x, a.y, a.z}, f3{b.x, b.y, b.z}, f3{c.x, c.y, c.z}, h)) return false;
            t = h.t;
        } else {
            double ts = sphere_test(r, sc.spheres[__float_as_int(a.x)], r.tfar);
            if (ts < 0) return false;
            t = (float)ts;
        }
        if (ANY) return true;
        if (best_slot < 0 || t < best_t || (t == best_t && tie_less(sc, slot, best_slot))) { best_t = t; best_slot = slot; }
        return false;
    }

    // After a phase: if the node group is exhausted, pull work from the stack.  Node groups go to
    // `ng`; postponed primitive groups go to `tg` if it is free, otherwise the lane stays blocked
    // on its primitive work (wants_node() false, wants_prim() true).
    LJ_D void refill_from_stack() {
        while (!wants_node() && sp > 0) {
            const uint2 e = peek();
            if (e.y & 0xff000000u) { ng = e; pop(); }
            else if (tg.y == 0u) { tg = e; pop(); }
            else break;
        }
    }

    // Node phase: open one pending inner child and slab-test its 8 children.  Inner hits become
    // the new node group; leaf hits become a primitive group, held in `tg` or postponed on the stack.
    // Requires wants_node().  Returns true only if an occluder was found (ANY, stack-overflow path).
    // REFILL = false: the caller pulls from the stack itself (the wavefront trace loops do it once per iteration with
    // all lanes converged, instead of inside the divergent tail of each phase).
    template <bool ANY, bool COUNT, bool REFILL = true>
    LJ_D bool node_step(const DScene &sc) {
        const int bit = 31 - __clz(ng.y);
        ng.y &= ~(1u << bit);
        if ((ng.y & 0xff000000u) && sp < LJ_BVH8_STACK) push(ng);
        const unsigned slot = (unsigned)(bit - 24) ^ oct_inv;
        const unsigned rel = __popc(ng.y & 0xffu & ((1u << slot) - 1u));
        const float4 *np = reinterpret_cast<const float4 *>(&sc.bvh8[ng.x + rel]);
        const float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
        if (COUNT) n_nodes++;
        const unsigned ew = __float_as_uint(n0.w);
        // Child plane along one axis: t = q * A + B with A = 2^e / d, B = (p - o) / d, q the quantised
        // byte.  The byte is spliced into the mantissa of 2^15 with one PRMT (M = 32768 + q, exact), so
        // t = fma(M, A, C) with C = B - 32768 A: one PRMT + one FFMA per plane, no int->float conversion.
        // Rounding: A is off by 2^-24 relative (the reciprocal), B by 3 roundings, C by one more and by the
        // cancellation against 32768 A, t by its own rounding: |t~ - t| <= 2^-24 (|t| + 4 |C| + 5 * 32768 |A|)
        // <= 2^-24 * 5 |C| + 0.012 |A|.  The near (far) planes are therefore moved out by |A|/64 + |C| * 2^-21 once
        // per node and axis; boxes grow by < 2 % of a quantisation step and the test stays conservative without
        // any per-child padding.
        const float ax = __uint_as_float(byte_of(ew, 0) << 23) * idx;
        const float ay = __uint_as_float(byte_of(ew, 1) << 23) * idy;
        const float az = __uint_as_float(byte_of(ew, 2) << 23) * idz;
        const float cx = fmaf(-32768.0f, ax, (n0.x - r.ox) * idx);
        const float cy = fmaf(-32768.0f, ay, (n0.y - r.oy) * idy);
        const float cz = fmaf(-32768.0f, az, (n0.z - r.oz) * idz);
        const float px =fmaf(fabsf(cx), 4.76837158e-7f, fabsf(ax) * 0.015625f);
        const float py = fmaf(fabsf(cy), 4.76837158e-7f, fabsf(ay) * 0.015625f);
        const float pz = fmaf(fabsf(cz), 4.76837158e-7f, fabsf(az) * 0.015625f);
        const float c0x = cx - px, c1x = cx + px, c0y = cy - py, c1y = cy + py, c0z = cz - pz, c1z = cz + pz;
        const unsigned oct4 = oct_inv * 0x01010101u;
        const unsigned m15 = sc.m15_bits;
        unsigned hitmask = 0;
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const unsigned meta4 = __float_as_uint(half ? n1.w : n1.z);
            const unsigned qlx = __float_as_uint(half ? n2.y : n2.x), qly = __float_as_uint(half ? n2.w : n2.z);
            const unsigned qlz = __float_as_uint(half ? n3.y : n3.x), qhx = __float_as_uint(half ? n3.w : n3.z);
            const unsigned qhy = __float_as_uint(half ? n4.y : n4.x), qhz = __float_as_uint(half ? n4.w : n4.z);
            // near / far plane bytes per axis depend on the direction sign
            const unsigned nx = idx < 0.f ? qhx : qlx, fx = idx < 0.f ? qlx : qhx;
            const unsigned ny = idy < 0.f ? qhy : qly, fy = idy < 0.f ? qly : qhy;
            const unsigned nz = idz < 0.f ? qhz : qlz, fz = idz < 0.f ? qlz : qhz;
            // four children at once: which are inner (meta bits 3 and 4 both set), the bit each child
            // sets in the hit mask (inner: slot -> traversal priority by octant), and the bits it sets
            const unsigned inner4 = ((meta4 & (meta4 << 1)) & 0x10101010u) >> 4;       // 0x01 per inner child
            const unsigned index4 = (meta4 ^ (oct4 & (inner4 * 7u))) & 0x1f1f1f1fu;
            const unsigned bits4 = (meta4 >> 5) & 0x07070707u;                         // 0 for an empty child
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float t0x = fmaf(byte_to_m15(nx, j, m15), ax, c0x), t1x = fmaf(byte_to_m15(fx, j, m15), ax, c1x);
                const float t0y = fmaf(byte_to_m15(ny, j, m15), ay, c0y), t1y = fmaf(byte_to_m15(fy, j, m15), ay, c1y);
                const float t0z = fmaf(byte_to_m15(nz, j, m15), az, c0z), t1z = fmaf(byte_to_m15(fz, j, m15), az, c1z);
                const float t0 = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, r.tnear));
                const float t1 = fminf(fminf(t1x, t1y), fminf(t1z, best_t));
                if (t0 <= t1)      // one PRMT extracts the byte; the shift uses the low 5 bits of the index byte
                    hitmask |= __byte_perm(bits4, 0u, 0x4440u + j) << byte_of(index4, j);
            }
        }
        ng = make_uint2(__float_as_uint(n1.x), (hitmask & 0xff000000u) | (ew >> 24));
        const uint2 ntg = make_uint2(__float_as_uint(n1.y), hitmask & 0x00ffffffu);
        if (ntg.y) {
            if (tg.y == 0u) tg = ntg;
            else if (sp < LJ_BVH8_STACK) push(ntg);
            else {                                  // stack full (never seen in practice): test the new group right away
                unsigned m = ntg.y;
                while (m) {
                    const int b = __ffs(m) - 1; m &= m - 1;
                    if (COUNT) n_prims++;
                    if (prim<ANY>(sc, (int)ntg.x + b)) return true;
                }
            }
        }
        if (REFILL) refill_from_stack();
        return false;
    }

    // Primitive phase: test ONE pending primitive of `tg`.  Requires wants_prim().  Returns true
    // when an occluder was found (ANY only).
    template <bool ANY, bool COUNT, bool REFILL = true>
    LJ_D bool prim_step(const DScene &sc, bool &found) {
        const int b = __ffs(tg.y) - 1;
        tg.y &= tg.y - 1;
        if (COUNT) n_prims++;
        if (prim<ANY>(sc, (int)tg.x + b)) { found = true; return true; }
        if (REFILL && tg.y == 0u) refill_from_stack();
        return false;
    }

    // Scalar driver (whole-path engines): run to completion, primitives right after their node.
    template <bool ANY, bool COUNT>
    LJ_D bool run(const DScene &sc) {
        bool found = false;
        while (true) {
            while (wants_prim()) if (prim_step<ANY, COUNT>(sc, found)) return true;
            if (!wants_node()) break;
            if (node_step<ANY, COUNT>(sc)) return true;
        }
        return false;
    }

    // full hit record of the winning primitive (same arithmetic as during traversal)
    LJ_D bool finish(const DScene &sc, Hit *h) const {
        if (best_slot < 0) return false;
        const float4 a = __ldg(&sc.tri_verts[3 * best_slot + 0]);
        const float4 b = __ldg(&sc.tri_verts[3 * best_slot + 1]);
        const float4 c = __ldg(&sc.tri_verts[3 * best_slot + 2]);
        h->geom = __float_as_int(a.w); h->prim = __float_as_int(b.w);
        if (__float_as_int(c.w) == 0) {
            TriHit th;
            const f3 O{r.ox, r.oy, r.oz}, D{r.dx, r.dy, r.dz};
            tri_test(O, D, r.tnear, r.tfar, f3{a.x, a.y, a.z}, f3{b.x, b.y, b.z}, f3{c.x, c.y, c.z}, th);
            h->t = th.t; h->u = th.u; h->v = th.v; h->ngx = th.ng.x; h->ngy = th.ng.y; h->ngz = th.ng.z;
        } else {
            const SpherePrim s = sc.spheres[__float_as_int(a.x)];
            double ts = sphere_test(r, s, r.tfar);
            h->t = (float)ts;
            sphere_hit_details(r, s, ts, *h);
        }
        return true;
    }
};

LJ_D bool trace_closest(const DScene &sc, const RayF &r, Hit *h) {
    if (!sc.bvh8) return traverse_bvh2<false>(sc, r, h);
    Trav8 tv; tv.init(r);
    tv.run<false, false>(sc);
    return tv.finish(sc, h);
}
LJ_D bool trace_any(const DScene &sc, const RayF &r) {
    if (!sc.bvh8) return traverse_bvh2<true>(sc, r, nullptr);
    Trav8 tv; tv.init(r);
    return tv.run<true, false>(sc);
}
