// Collapse of the binary radix tree into an 8-wide BVH with quantised child boxes
// (80-byte nodes, layout of Ylitie, Karras, Laine 2017).  Runs on the device, one
// tree level per launch: each thread owns one wide node, pulls up to 8 children out
// of the binary tree by repeatedly opening the child with the largest surface area,
// assigns them to octant-ordered slots, allocates its inner children / primitive
// ranges with atomics and quantises the child boxes conservatively.
#include "lj_bvh_build.h"
#include "lj_alloc.h"

#include <cooperative_groups.h>
#include <cfloat>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { snprintf(err, 256, "%s: %s", #x, cudaGetErrorString(e_)); return -1; } } while (0)

namespace {

constexpr int MAX_LEAF_CAP = 3;   // primitives per leaf slot (unary count in 3 bits)

struct Work { int node2; int node8; };

// number of primitives / first sorted slot under binary-tree child id (inner >= 0, leaf = ~slot)
__device__ __forceinline__ int child_count(int c, const int *count) { return c < 0 ? 1 : count[c]; }

__device__ __forceinline__ void child_box(int c, const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                                          const float4 *prim_lo, const float4 *prim_hi, float *lo, float *hi) {
    float4 l, h;
    if (c < 0) { int p = sorted_prim[~c]; l = prim_lo[p]; h = prim_hi[p]; } else { l = node_lo[c]; h = node_hi[c]; }
    lo[0] = l.x; lo[1] = l.y; lo[2] = l.z; hi[0] = h.x; hi[1] = h.y; hi[2] = h.z;
}

// One work item = one 8-wide node, built by a group of 8 lanes (four items per warp): gathers up to 8 children of binary
// node w.node2, writes node w.node8, queues the inner children.  During the gather lane i owns child i (id, primitive
// count, box); afterwards lane s owns octant slot s.  The choices are the sequential algorithm's: "largest box first,
// first one on ties" and the greedy slot assignment in (child, slot) order — taken as minima / maxima over the group
// with the lane index as the tie-break.  (One lane per item took ~80 us per tree level: a few thousand dependent
// instructions and loads; sponza's nine levels were 0.7 ms of a 1.9 ms build.)
__device__ __forceinline__ void load_child8(int c, const int *count, const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                                            const float4 *prim_lo, const float4 *prim_hi, int &cnt, float *lo, float *hi) {
    cnt = child_count(c, count);
    child_box(c, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, lo, hi);
}
// lane of the group holding the largest key (first lane on ties); keys < 0 never win: returns -1 when there is none
__device__ __forceinline__ int group_argmax8(unsigned gmask, int lane, float key) {
    int who = key >= 0.f ? lane : -1;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const float ok = __shfl_xor_sync(gmask, key, o);
        const int ow = __shfl_xor_sync(gmask, who, o);
        if (ow >= 0 && (who < 0 || ok > key || (ok == key && ow < who))) { key = ok; who = ow; }
    }
    return who;
}
__device__ void collapse_item8(const Work w, const unsigned gmask, const int lane /* 0..7 in the group */, const int glane0 /* warp lane of the group's lane 0 */,
                               Work *items_out, int *n_out,
                               const int *left, const int *right, const float4 *node_lo, const float4 *node_hi,
                               const int *count, const int *sorted_prim,
                               const float4 *prim_lo, const float4 *prim_hi, const float4 *tri_verts,
                               Bvh8Node *nodes8, int *node_counter, float4 *tri_verts8, int *prim_counter, int MAX_LEAF, int open_leaves) {
    int c = 0, cnt = 0, nchild; bool valid = false;
    float lo[3] = {0.f, 0.f, 0.f}, hi[3] = {0.f, 0.f, 0.f};
    float plo[3], phi[3];
    if (w.node2 < 0) {            // degenerate tree: a single primitive
        nchild = 1;
        child_box(w.node2, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, plo, phi);
        if (lane == 0) { c = w.node2; valid = true; load_child8(c, count, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, cnt, lo, hi); }
    } else {
        nchild = 2;
        const float4 l = node_lo[w.node2], h = node_hi[w.node2];
        plo[0] = l.x; plo[1] = l.y; plo[2] = l.z; phi[0] = h.x; phi[1] = h.y; phi[2] = h.z;
        if (lane < 2) { c = lane == 0 ? left[w.node2] : right[w.node2]; valid = true; load_child8(c, count, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, cnt, lo, hi); }
    }
    // open the largest expandable child until 8 children; then, with slots left over (the bottom of the tree: every child
    // already holds <= MAX_LEAF primitives), split multi-primitive leaf children as well, largest box first, so that each
    // primitive (or pair) gets its own, tighter child box: same number of nodes, fewer primitive tests
    for (int pass = 0; pass < 2; pass++) {
        if (pass == 1 && !open_leaves) break;
        while (nchild < 8) {
            const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
            const float area = dx * dy + dy * dz + dz * dx;
            const bool expandable = valid && c >= 0 && (pass == 1 || cnt > MAX_LEAF);
            const int best = group_argmax8(gmask, lane, expandable && area > -1.f ? area : -1.f);
            if (best < 0) break;
            const int cb = __shfl_sync(gmask, c, glane0 + best);
            const int l = left[cb], r = right[cb];
            if (lane == best) { c = l; load_child8(c, count, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, cnt, lo, hi); }
            if (lane == nchild) { c = r; valid = true; load_child8(c, count, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, cnt, lo, hi); }
            nchild++;
        }
    }
    // greedy assignment of children to octant slots: slot bit a set <=> child lies on the + side of axis a
    const float cx = 0.5f * (plo[0] + phi[0]), cy = 0.5f * (plo[1] + phi[1]), cz = 0.5f * (plo[2] + phi[2]);
    const float off0 = 0.5f * (lo[0] + hi[0]) - cx, off1 = 0.5f * (lo[1] + hi[1]) - cy, off2 = 0.5f * (lo[2] + hi[2]) - cz;
    unsigned slot_free = 0xffu; int my_slot = -1;
    for (int round = 0; round < nchild; round++) {
        float best = -FLT_MAX; int bs = -1;
        if (valid && my_slot < 0) {
#pragma unroll
            for (int s = 0; s < 8; s++) {
                if (!(slot_free & (1u << s))) continue;
                const float cost = ((s & 1) ? off0 : -off0) + ((s & 2) ? off1 : -off1) + ((s & 4) ? off2 : -off2);
                if (cost > best) { best = cost; bs = s; }
            }
        }
        // the pair the sequential scan (children outer, slots inner, strict >) would pick: largest cost, first child on ties
        int who = bs >= 0 ? lane : -1;
        float key = best;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const float ok = __shfl_xor_sync(gmask, key, o);
            const int ow = __shfl_xor_sync(gmask, who, o);
            if (ow >= 0 && (who < 0 || ok > key || (ok == key && ow < who))) { key = ok; who = ow; }
        }
        if (who < 0) break;         // (cannot happen with finite boxes)
        const int ws = __shfl_sync(gmask, bs, glane0 + who);
        slot_free &= ~(1u << ws);
        if (lane == who) my_slot = ws;
    }
    // from here on lane s owns slot s: find the child assigned to it and fetch its data
    int ci = -1;
#pragma unroll
    for (int i = 0; i < 8; i++) { const int sl = __shfl_sync(gmask, my_slot, glane0 + i); if (sl == lane) ci = i; }
    const int src = glane0 + (ci < 0 ? 0 : ci);
    const int sc_ = __shfl_sync(gmask, c, src), scnt = __shfl_sync(gmask, cnt, src);
    float slo[3], shi[3];
#pragma unroll
    for (int a = 0; a < 3; a++) { slo[a] = __shfl_sync(gmask, lo[a], src); shi[a] = __shfl_sync(gmask, hi[a], src); }
    // classify, allocate
    const bool inner = ci >= 0 && sc_ >= 0 && scnt > MAX_LEAF;
    const int leaf_cnt = (ci >= 0 && !inner) ? scnt : 0;
    const unsigned imask = (__ballot_sync(gmask, inner) >> glane0) & 0xffu;
    const int n_inner = __popc(imask), inner_rank = __popc(imask & ((1u << lane) - 1u));
    int incl = leaf_cnt;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) { const int y = __shfl_up_sync(gmask, incl, o, 8); if (lane >= o) incl += y; }
    const int prim_off = incl - leaf_cnt, n_leaf_prims = __shfl_sync(gmask, incl, glane0 + 7);
    int child_base = 0, prim_base = 0, out_base = 0;
    if (lane == 0) {
        child_base = n_inner ? atomicAdd(node_counter, n_inner) : 0;
        prim_base = n_leaf_prims ? atomicAdd(prim_counter, n_leaf_prims) : 0;
        out_base = n_inner ? atomicAdd(n_out, n_inner) : 0;
    }
    child_base = __shfl_sync(gmask, child_base, glane0); prim_base = __shfl_sync(gmask, prim_base, glane0); out_base = __shfl_sync(gmask, out_base, glane0);

    // quantisation frame: origin = node lo, per-axis power-of-two scale so that 255 steps cover the extent
    int e[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        float ext = phi[a] - plo[a];
        int ea;
        if (!(ext > 0.f)) ea = -100;
        else {
            ea = (int)ceilf(log2f(ext / 255.f));
            // make sure 255 * 2^e really reaches the far side in float arithmetic
            while (plo[a] + 255.f * exp2f((float)ea) < phi[a]) ea++;
        }
        if (ea < -126) ea = -126;
        e[a] = ea;
    }
    Bvh8Node *nd = &nodes8[w.node8];
    if (lane == 0) {
        nd->px = plo[0]; nd->py = plo[1]; nd->pz = plo[2];
        nd->ex = (uint8_t)(e[0] + 127); nd->ey = (uint8_t)(e[1] + 127); nd->ez = (uint8_t)(e[2] + 127);
        nd->imask = (uint8_t)imask; nd->child_base = child_base; nd->prim_base = prim_base;
    }
    const int s = lane;
    uint8_t q[6] = {0, 0, 0, 0, 0, 0};
    uint8_t meta = 0;
    if (ci >= 0) {
        if (inner) {
            meta = (uint8_t)((1u << 5) | (24 + s));
            Work o; o.node2 = sc_; o.node8 = child_base + inner_rank;
            items_out[out_base + inner_rank] = o;
        } else {
            const unsigned unary = (1u << scnt) - 1u;
            meta = (uint8_t)((unary << 5) | prim_off);
            // enumerate the (at most MAX_LEAF) leaves of this small subtree
            int stk[MAX_LEAF_CAP + 1]; int sp = 0, k = 0;
            stk[sp++] = sc_;
            while (sp) {
                int x = stk[--sp];
                if (x >= 0) { stk[sp++] = right[x]; stk[sp++] = left[x]; continue; }
                int srcp = ~x, dst = prim_base + prim_off + k++;
                tri_verts8[3 * dst + 0] = tri_verts[3 * srcp + 0];
                tri_verts8[3 * dst + 1] = tri_verts[3 * srcp + 1];
                tri_verts8[3 * dst + 2] = tri_verts[3 * srcp + 2];
            }
        }
#pragma unroll
        for (int a = 0; a < 3; a++) {
            float scale = exp2f((float)e[a]);
            float p = plo[a];
            int ql = (int)floorf((slo[a] - p) / scale);
            int qh = (int)ceilf((shi[a] - p) / scale);
            ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
            while (ql > 0 && p + (float)ql * scale > slo[a]) ql--;
            while (qh < 255 && p + (float)qh * scale < shi[a]) qh++;
            q[a] = (uint8_t)ql; q[3 + a] = (uint8_t)qh;
        }
    }
    nd->meta[s] = meta;
    nd->qlox[s] = q[0]; nd->qloy[s] = q[1]; nd->qloz[s] = q[2];
    nd->qhix[s] = q[3]; nd->qhiy[s] = q[4]; nd->qhiz[s] = q[5];
}

struct CollapseArgs {
    const int *left, *right; const float4 *node_lo, *node_hi; const int *count, *sorted_prim;
    const float4 *prim_lo, *prim_hi, *tri_verts;
    Bvh8Node *nodes8; float4 *tri_verts8; int n_prims_total, max_leaf, open_leaves;
};
// counters: [0] 8-wide nodes allocated, [1] primitive slots, [2..4] queue sizes of three consecutive levels (rotating), [5] levels

// one level per launch (LJ_BVH8_COOP=0): the host reads the size of the next level back after each
__global__ void __launch_bounds__(64) k_collapse_level(const Work *items, int n_items, Work *items_out, int *n_out, int *counters, CollapseArgs a) {
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const int wi = gtid >> 3, wl = threadIdx.x & 31, glane0 = wl & 24;
    if (wi >= n_items) return;
    collapse_item8(items[wi], 0xffu << glane0, wl & 7, glane0, items_out, n_out, a.left, a.right, a.node_lo, a.node_hi, a.count, a.sorted_prim,
                   a.prim_lo, a.prim_hi, a.tri_verts, a.nodes8, counters, a.tri_verts8, counters + 1, a.max_leaf, a.open_leaves);
}
__global__ void k_collapse_root(Work *q0, const int *d_root2, int n) {
    Work root; root.node2 = (n >= 2) ? *d_root2 : ~0; root.node8 = 0;
    q0[0] = root;
}

// every level in ONE cooperative launch (the default): a grid barrier between levels instead of a launch, a memset and a
// blocking read-back.  The queue size of level L is counters[2 + L % 3]; level L fills the counter of L + 1 and zeroes
// the one of L + 2 (last read before the previous barrier).
__global__ void __launch_bounds__(64) k_collapse_all(Work *q0, Work *q1, int *counters, const int *d_root2, CollapseArgs a) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, n_groups = (gridDim.x * blockDim.x) >> 3;
    const int wl = threadIdx.x & 31, glane0 = wl & 24;
    int level = 0, n_items = 1;
    while (n_items > 0 && level <= 128) {
        const Work *items = (level & 1) ? q1 : q0;
        Work *items_out = (level & 1) ? q0 : q1;
        int *n_out = counters + 2 + (level + 1) % 3;
        if (gtid == 0) counters[2 + (level + 2) % 3] = 0;
        for (int wi = gtid >> 3; wi < n_items; wi += n_groups) {
            Work w;
            if (level == 0) { w.node2 = (a.n_prims_total >= 2) ? *d_root2 : ~0; w.node8 = 0; } else w = items[wi];
            collapse_item8(w, 0xffu << glane0, wl & 7, glane0, items_out, n_out, a.left, a.right, a.node_lo, a.node_hi, a.count, a.sorted_prim,
                           a.prim_lo, a.prim_hi, a.tri_verts, a.nodes8, counters, a.tri_verts8, counters + 1, a.max_leaf, a.open_leaves);
        }
        grid.sync();
        n_items = *(volatile int *)n_out;
        level++;
    }
    if (gtid == 0) counters[5] = level;
}

} // namespace

// d_root2 (device): root id of the binary tree (ignored when n < 2).  d_counters_out (device, 6 ints): the counters described
// above, left for the caller to read back together with its own results: [0] nodes, [1] slots, [5] levels (> 128 = did not terminate).
int lj_bvh8_collapse(int n, const int *d_root2, const float4 *tri_verts, const int *left, const int *right, const int *count,
                     const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                     const float4 *prim_lo, const float4 *prim_hi, BvhBuildOutput &out, int *counters, char *err) {
    if (n <= 0) return 0;
    Work *q0 = nullptr, *q1 = nullptr;
    CK(lj_malloc(&q0, sizeof(Work) * (size_t)n)); CK(lj_malloc(&q1, sizeof(Work) * (size_t)n));
    CK(lj_malloc(&out.bvh8, sizeof(Bvh8Node) * (size_t)n));
    CK(lj_malloc(&out.tri_verts8, sizeof(float4) * 3 * (size_t)n));
    int h_counters[6] = {1 /* nodes allocated: root */, 0 /* prims */, 1, 0, 0, 0};
    CK(cudaMemcpyAsync(counters, h_counters, sizeof h_counters, cudaMemcpyHostToDevice, 0));
    CollapseArgs a;
    a.left = left; a.right = right; a.node_lo = node_lo; a.node_hi = node_hi; a.count = count; a.sorted_prim = sorted_prim;
    a.prim_lo = prim_lo; a.prim_hi = prim_hi; a.tri_verts = tri_verts; a.nodes8 = out.bvh8; a.tri_verts8 = out.tri_verts8;
    a.n_prims_total = n;
    a.max_leaf = MAX_LEAF_CAP;        // LJ_MAX_LEAF: experiment knob (1..3)
    if (const char *e = getenv("LJ_MAX_LEAF")) { int v = atoi(e); if (v >= 1 && v <= MAX_LEAF_CAP) a.max_leaf = v; }
    a.open_leaves = 1;                // LJ_OPEN_LEAVES=0: experiment knob
    if (const char *e = getenv("LJ_OPEN_LEAVES")) a.open_leaves = atoi(e) != 0;
    bool coop = true;
    if (const char *e = getenv("LJ_BVH8_COOP")) coop = atoi(e) != 0;
    if (coop) {                       // a device without cooperative launches gets one launch per level
        int dev = 0, ok = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&ok, cudaDevAttrCooperativeLaunch, dev));
        coop = ok != 0;
    }
    if (coop) {
        int per_sm = 0, dev = 0, sm_count = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse_all, 64, 0));
        int grid = per_sm * sm_count;
        if (grid < 1) { snprintf(err, 256, "bvh8 collapse: no resident block for the cooperative launch"); return -1; }
        const int want = (n / 4 * 8 + 63) / 64 + 1;     // 8 lanes per node; the widest level holds fewer than n / 4 nodes
        if (grid > want) grid = want;
        void *args[] = {&q0, &q1, &counters, &d_root2, &a};
        CK(cudaLaunchCooperativeKernel((const void *)k_collapse_all, dim3(grid), dim3(64), args, 0, 0));
    } else {
        k_collapse_root<<<1, 1>>>(q0, d_root2, n);
        int n_items = 1, levels = 0;
        while (n_items > 0) {
            int *n_out = counters + 2 + (levels + 1) % 3;
            CK(cudaMemsetAsync(n_out, 0, sizeof(int), 0));
            k_collapse_level<<<(n_items * 8 + 63) / 64, 64>>>(q0, n_items, q1, n_out, counters, a);
            CK(cudaGetLastError());
            CK(cudaMemcpy(&n_items, n_out, sizeof(int), cudaMemcpyDeviceToHost));
            Work *t = q0; q0 = q1; q1 = t;
            if (++levels > 128) break;
        }
        CK(cudaMemcpyAsync(counters + 5, &levels, sizeof(int), cudaMemcpyHostToDevice, 0));
        CK(cudaStreamSynchronize(0));       // `levels` is a stack variable
    }
    lj_free(q0); lj_free(q1);
    return 0;
}
